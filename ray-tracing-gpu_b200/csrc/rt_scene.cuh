// rt_scene.cuh -- device-side scene / queue layouts shared by all kernels.
#pragma once
#include "rt_math.cuh"

namespace rtb {

// ---- unified, insertion-ordered surface list (64 B per surface, 4 x float4) -------------
// slot k of the list IS insertion index k of CScene::m_Surfaces (Scene.cpp:107), so a hit id
// is an order and ties resolve like the reference's linear scan (Scene.cpp:1713).
//   triangle: r0 = p0.xyz | type   r1 = e1.xyz | material   r2 = e2.xyz | 0   r3 = normal.xyz | 0
//   quadric : r0 = A B C  | type   r1 = yz xz xy | material r2 = lx ly lz | K r3 = 0
//   plane   : r0 = n.xyz  | type   r1 = c 0 0 | material
enum { SURF_TRIANGLE = 0, SURF_QUADRIC = 1, SURF_PLANE = 2 };
struct SurfRec { float4 r0, r1, r2, r3; };

// material: m0 = r g b ka | m1 = kd ks shininess kr | m2 = kt ior 0 0     (48 B)
struct MatRec { float4 m0, m1, m2; };
// light: l0 = position | intensity, l1 = colour | 0                         (32 B)
struct LightRec { float4 l0, l1; };

// "resident blob": [SurfRec x N][MatRec x M][LightRec x L], 16-byte aligned, one bulk copy.
struct ResidentLayout {
    int n_surf, n_mat, n_light;
    int off_mat, off_light, bytes;      // byte offsets inside the blob
};

// ---- streamed (tiled brute-force) layout ------------------------------------------------
// triangles: 3 x float4 per triangle, tile-contiguous:  p0.xyz | order,  e1.xyz | material, e2.xyz | 0
struct TriRec { float4 a, b, c; };
// quadrics: q0 = A B C | order, q1 = yz xz xy | material, q2 = lx ly lz | K
struct QuadRec { float4 q0, q1, q2; };
// planes:  n.xyz c | then order/material in a parallel int2
struct PlaneRec { float4 nc; int order, material, pad0, pad1; };

// LBVH node (rt_bvh.cu): the boxes of BOTH children + their references, 64 bytes, one fetch
// decides two subtrees.  b0 = c0.min.xyz | c0.max.x, b1 = c0.max.yz | c1.min.xy,
// b2 = c1.min.z | c1.max.xyz, c = (ref0, ref1, -, -).  ref >= 0: internal node index;
// ref < 0: leaf, ~ref = first | (count - 1) << 28 over the Morton-sorted triangle array.
struct BvhNode { float4 b0, b1, b2; int4 c; };

struct DeviceScene {
    ResidentLayout rl;
    const unsigned char* blob;      // resident blob in global memory
    const SurfRec*  surf;           // = blob
    const MatRec*   mat;
    const LightRec* light;
    // streamed layout
    int n_tri, n_quad, n_plane;
    const TriRec*   tri;            // padded to a multiple of the tile size with never-hit triangles
    const float4*   tri_normal;     // normal.xyz | unused
    const QuadRec*  quad;
    const PlaneRec* plane;
    int n_tri_padded;
    // LBVH over the triangles (null until rtb200_build_bvh)
    const BvhNode*  bvh_nodes;
    const TriRec*   bvh_tri;        // primitives in Morton order: triangles (order / material in .a.w / .b.w, .c.w = 0)
                                    // and bounded quadrics (A B C | order, lx ly lz | material, K rhs 0 | 1.0f)
    int             bvh_root;       // reference of the root (a leaf reference when there are <= BVH_LEAF_MAX primitives)
    int             bvh_n_prims;    // triangles + bounded quadrics in the tree
    const QuadRec*  bvh_quad;       // quadrics the LBVH kernels still scan linearly (unbounded / rotated ones)
    int             bvh_n_quad;
    // exactness beyond the padded radius (rt_bvh.cu, "far rays"): the boxes are padded for ray origins whose farthest
    // corner of [bvh_lo, bvh_hi] is within bvh_radius; a ray that starts farther out (a reflection off the ground plane
    // near the horizon) walks the tree with every box grown by the pad ITS origin needs
    float bvh_lo[3], bvh_hi[3], bvh_radius;
    float bvh_grow_tri;             // triangles: grow = bvh_grow_tri * reach (the pad of the largest |e1||e2|, per unit of reach)
    float bvh_kq, bvh_cmax;         // bounded quadrics in the tree: grow = sqrt(2.1e-5 kq (reach - R)(2 cmax + reach + R))
};

// ---- per-frame constants ------------------------------------------------------------------
struct FrameParams {
    int   W, H;                 // pixels
    int   k;                    // spp grid
    int   LW, LH;               // sample lattice = k*W, k*H
    float inv_w, inv_h;         // 1/(k*W), 1/(k*H)  (Scene.cpp:678-679 at lattice resolution)
    float half_w, half_h;
    float cam_pos[3], cam_u[3], cam_v[3], cam_n[3];
    float background[3];
    int   max_bounces;
    float min_energy, scene_ior;
    int   secondary;            // 0: never spawn
    // banding (multi-GPU): local lattice row -> global lattice row
    int   band_lrows;           // band height in lattice rows (band_rows * k)
    int   band_lshift, band_pshift;   // log2 of the band height in lattice / pixel rows, -1 if not a power of two
    int   n_ranks, rank;
    int   fb_global;            // 1: the framebuffers are FULL frames (rank 0's, possibly peer-mapped over
                                //    NVLink) and pixel rows are addressed globally; 0: compact local rows
};

// ---- wavefront queues -----------------------------------------------------------------------
// level k >= 1 ray: o.xyz | energy,  d.xyz | ior.  (bounce count == level)
struct RayQueue {
    float4* od0;
    float4* od1;
    float4* colour;             // result of each ray (level >= 1)
    int4*   links;              // parents at THIS level that spawned: parent, refl child, refr child, material
    unsigned int* counters;     // [0] rays in the NEXT level, [1] links at this level (device)
};

RT_DEV V3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }
RT_DEV V3 f4_xyz(float4 v) { return mk3(v.x, v.y, v.z); }

// Interleaved banding (multi-GPU): rank r owns bands r, r + n, r + 2n, ... of `band` rows each and
// stores them compactly.  Global row of local row `local` -- the ONE formula used by the kernels,
// the host-side gather and rtb200_band_map.
__host__ __device__ inline int band_global_row(int band, int n_ranks, int rank, int local)
{
    if (n_ranks <= 1) return local;
    const int b = local / band, r = local - b * band;
    return (b * n_ranks + rank) * band + r;
}

// Global lattice row of local lattice row `ly` under interleaved banding.
RT_DEV int global_lattice_row(const FrameParams& f, int ly)
{
    if (f.n_ranks <= 1) return ly;
    if (f.band_lshift >= 0) {                              // same formula, shifts instead of a division per sample
        const int b = ly >> f.band_lshift;
        return (((b * f.n_ranks + f.rank) << f.band_lshift)) + (ly - (b << f.band_lshift));
    }
    return band_global_row(f.band_lrows, f.n_ranks, f.rank, ly);
}

// Row of the framebuffer a LOCAL pixel row is stored at (see FrameParams::fb_global).
RT_DEV int framebuffer_row(const FrameParams& f, int local_row)
{
    if (!f.fb_global || f.n_ranks <= 1) return local_row;
    if (f.band_pshift >= 0) {
        const int b = local_row >> f.band_pshift;
        return (((b * f.n_ranks + f.rank) << f.band_pshift)) + (local_row - (b << f.band_pshift));
    }
    return band_global_row(f.band_lrows / f.k, f.n_ranks, f.rank, local_row);
}

// Peer frame (rt_comm.cu): called by EVERY CTA of the grid after its last framebuffer store.  The
// stores went into rank 0's memory over NVLink; each CTA makes them visible system-wide and counts
// itself done, and the last one to do so writes the frame's sequence number into this rank's word
// of rank 0's flag array (release at system scope): rank 0's stream waits on that word.
__device__ __forceinline__ void peer_frame_signal(unsigned int* done, unsigned int* flag, unsigned int seq)
{
    __syncthreads();                                        // every thread of the CTA has issued its stores
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        __threadfence_system();
        const unsigned int n = gridDim.x * gridDim.y * gridDim.z;
        if (atomicAdd(done, 1u) == n - 1u) {
            *done = 0u;                                     // ready for the next launch (stream order)
            __threadfence_system();
            asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(flag), "r"(seq) : "memory");
        }
    }
}

// Peer frame: thread 0 of a CTA waits until rank 0's acknowledgement word (LOCAL memory, written by
// rank 0 over NVLink) has reached `want` (cyclic compare).  Normally true on the first load.  Gives
// up after 2 s of GPU time (a hung GPU is worse than a failed call) and says so in *timeouts: the host turns a
// non-zero count into RTB200_E_NCCL -- the frame may be torn and must not be used.  Followed by a barrier.
__device__ __forceinline__ void peer_frame_wait_ack(const unsigned int* ack, unsigned int want, unsigned long long* timeouts)
{
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        unsigned long long t0 = 0;
        for (;;) {
            unsigned int v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(ack) : "memory");
            if ((int)(v - want) >= 0) break;
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > 2000000000ull) { if (timeouts) atomicAdd(timeouts, 1ull); break; }
            __nanosleep(200);
        }
    }
    __syncthreads();
}

// Primary ray through lattice sample (sx, sy): Scene.cpp:1543-1552.
template <bool E> RT_DEV void primary_ray(const FrameParams& f, int sx, int sy, V3& o, V3& d)
{
    typedef FP<E> F;
    // (2*PixX*InvW - 1) * HalfW : int multiply, then float
    float vx = F::mul(F::sub(F::mul((float)(2 * sx), f.inv_w), 1.0f), f.half_w);
    float vy = F::mul(F::sub(F::mul((float)(2 * sy), f.inv_h), 1.0f), f.half_h);
    float vz = -1.0f;
    // v * Orientation, rows U,V,N, zero translation row (Matrice4.h:297-306)
    V3 r;
    r.x = F::add(F::madd(f.cam_n[0], vz, F::madd(f.cam_v[0], vy, F::mul(f.cam_u[0], vx))), 0.0f);
    r.y = F::add(F::madd(f.cam_n[1], vz, F::madd(f.cam_v[1], vy, F::mul(f.cam_u[1], vx))), 0.0f);
    r.z = F::add(F::madd(f.cam_n[2], vz, F::madd(f.cam_v[2], vy, F::mul(f.cam_u[2], vx))), 0.0f);
    o = mk3(f.cam_pos[0], f.cam_pos[1], f.cam_pos[2]);
    d = normalise3<E>(r);
}

} // namespace rtb
