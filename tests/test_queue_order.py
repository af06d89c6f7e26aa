"""CPU: the index arithmetic of the tiled path's queue order (k_shade_a, rt_trace.cu; DESIGN.md 5.2 "Queue order"),
restated in numpy -- the device code is a pure permutation of the queues, these are the two facts that make it one:
  * the 16x16 patches of the level-0 CTAs cover every sample of a (ragged) chunk exactly once;
  * a CTA's shadow queries, queued light by light through per-(light, warp) ballot counts and one 64-entry scan,
    fill its reserved range exactly once, and sq_dst maps them one-to-one onto the canonical filter slots
    (q_base[ray] + rank of the light among the facing ones) that shade-B reads.
The GPU side is tests/test_gpu_stream.py::test_query_grouping_changes_no_bit (frames equal bit for bit)."""
import numpy as np


def patch_rays(LW, rows):
    tiles_x = (LW + 15) >> 4
    blocks = tiles_x * ((rows + 15) // 16)
    b, t = np.meshgrid(np.arange(blocks), np.arange(256), indexing="ij")
    x = (b % tiles_x) * 16 + (t & 15)
    y = (b // tiles_x) * 16 + (t >> 4)
    valid = (x < LW) & (y < rows)
    return (y * LW + x)[valid]


def test_patches_cover_the_chunk_once():
    for LW, rows in ((3840, 2160), (502, 262), (16, 16), (17, 1), (1, 33), (250, 150)):
        i = patch_rays(LW, rows)
        assert len(i) == LW * rows and np.array_equal(np.sort(i), np.arange(LW * rows))


def cta_queue(face, bq):
    """face: [256, L] bool.  Returns (slots, lights, rays, dst) of every query, as k_shade_a computes them."""
    n_thr, L = face.shape
    warp, lane = np.arange(n_thr) >> 5, np.arange(n_thr) & 31
    n_face = face.sum(1)
    q_base = bq + np.cumsum(n_face) - n_face                       # canonical: thread order, lights in order
    lcnt = np.zeros(64, np.int64)                                  # [light * 8 + warp]
    for l in range(L):
        for w in range(8):
            lcnt[l * 8 + w] = face[warp == w, l].sum()
    loff = np.cumsum(lcnt) - lcnt                                  # exclusive scan in (light, warp) order
    slots, lights, rays, dst = [], [], [], []
    for t in range(n_thr):
        k = 0
        for l in range(L):
            if face[t, l]:
                below = face[(warp == warp[t]) & (lane < lane[t]), l].sum()   # popc(ballot & lanes below)
                slots.append(bq + loff[l * 8 + warp[t]] + below); lights.append(l); rays.append(t)
                dst.append(q_base[t] + k); k += 1
    return np.array(slots), np.array(lights), np.array(rays), np.array(dst)


def test_light_by_light_queue_is_a_permutation():
    rng = np.random.default_rng(3)
    for L, p in ((1, 0.5), (3, 0.7), (4, 1.0), (8, 0.3), (8, 0.0), (5, 0.95)):
        face = rng.uniform(0, 1, (256, L)) < p
        face[rng.integers(0, 256, 40)] = False                     # misses and rays outside the chunk: no query
        bq = int(rng.integers(0, 10 ** 6))
        slots, lights, rays, dst = cta_queue(face, bq)
        total = int(face.sum())
        assert np.array_equal(np.sort(slots), bq + np.arange(total))
        assert np.array_equal(np.sort(dst), bq + np.arange(total))
        order = np.argsort(slots)
        key = lights[order] * 256 + rays[order]
        assert (np.diff(key) > 0).all()                            # light by light, rays in thread order within a light
        # shade-B reads filter[q_base[ray] + k] for the k-th facing light of the ray
        for t in np.unique(rays):
            mine = rays == t
            assert np.array_equal(dst[mine], dst[mine][0] + np.arange(mine.sum()))
            assert (np.diff(lights[mine]) > 0).all()
