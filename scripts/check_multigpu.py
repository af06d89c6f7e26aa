"""Multi-GPU parity check (run under torchrun, one rank per GPU): the frame assembled on rank 0 --
through the peer-frame path (kernels store into rank 0's memory over NVLink) and through the NCCL
gather -- must be bit-identical to the frame one GPU renders alone.  Resident, streamed and LBVH
paths, with and without supersampling, odd heights included."""
import json, os, sys
from importlib import import_module
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from __graft_entry__ import load_package
import torch, torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group(backend="cpu:gloo,cuda:nccl")
rt = load_package(); synth = import_module("rtb200.synth")
r = rt.Renderer(local)
uid = [r.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(uid, src=0)
r.comm_init(rank, world, uid[0])
solo = rt.Renderer(local) if rank == 0 else None
cases = [
    ("scene7 resident d4", rt.PackedScene.load(os.path.join(ROOT, "tests/golden/packed/scene7.npz")), 333, 217, 4, 1, 0, False),
    ("scene1 resident direct", rt.PackedScene.load(os.path.join(ROOT, "tests/golden/packed/scene1.npz")), 640, 480, 2, 1, 0, False),
    ("scene9 resident spp4", rt.PackedScene.load(os.path.join(ROOT, "tests/golden/packed/scene9.npz")), 200, 100, 3, 2, 0, False),
    ("soup streamed d3", synth.soup(2000, 16, 3, seed=7), 256, 150, 3, 1, rt.F_FORCE_STREAM, False),
    ("heightfield lbvh d3 spp4", synth.heightfield(20000, 300.0, 3, seed=3), 160, 90, 3, 2, rt.F_USE_BVH, True),
]
bad = 0
for name, ps, W, H, depth, k, flags, bvh in cases:
    for gather_f32 in (False, True):
        r.set_scene(ps)
        if bvh: r.build_bvh()
        fl = flags | (rt.F_GATHER_F32 if gather_f32 else 0)
        fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=k, flags=fl)
        f32, u8, st = r.render(fr, want_f32=gather_f32 and rank == 0, want_rgba8=rank == 0)
        dist.barrier()
        if rank == 0:
            solo.set_scene(ps)
            if bvh: solo.build_bvh()
            wf, w8, _ = solo.render(rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=k, flags=flags))
            ok8 = bool(np.array_equal(u8, w8))
            okf = True if not gather_f32 else bool(np.array_equal(f32.view(np.uint32), wf.view(np.uint32)))
            bad += (not ok8) + (not okf)
            print(json.dumps({"case": name, "ranks": world, "gather_f32": gather_f32, "p2p": not bool(os.environ.get("RTB200_NO_P2P")),
                              "rgba8_equal": ok8, "f32_equal": okf, "ms_gather": st.ms_gather, "ms_render": st.ms_render}), flush=True)
# a pipelined camera path (rtb200_render_sequence): the LAST frame left in rank 0's memory must equal
# the frame one GPU renders alone for the last camera
ps = rt.PackedScene.load(os.path.join(ROOT, "tests/golden/packed/scene1.npz"))
r.set_scene(ps)
W, H, n = 640, 363, 24
frames = []
for i in range(n):
    fr = rt.make_frame(ps.frame, W, H, max_bounces=2)
    fr.cam_pos[0] += 3.0 * i
    frames.append(fr)
_, sts = r.render_sequence(frames)
dist.barrier()
if rank == 0:
    solo.set_scene(ps)
    want_all = [solo.render(fr, want_f32=False)[1].copy() for fr in frames]
    _, got = r.read_frame(H, W, want_f32=False)
    ok = bool(np.array_equal(got, want_all[-1]))
    bad += not ok
    print(json.dumps({"case": "scene1 sequence of 24 cameras, frame left on rank 0", "ranks": world, "rgba8_equal": ok,
                      "ms_per_frame": sts[0].ms_render}), flush=True)
# the same path with a host destination on rank 0 (every frame must arrive, untorn), twice in a row
for rep in range(2):
    out = rt.pinned_empty((n, H, W, 4), np.uint8) if rank == 0 else None
    _, sts = r.render_sequence(frames, out_rgba8=out)
    if rank == 0:
        eq = [bool(np.array_equal(out[i], want_all[i])) for i in range(n)]
        bad += not all(eq)
        print(json.dumps({"case": "scene1 sequence of 24 cameras -> host", "ranks": world, "frames_equal": sum(eq), "of": n,
                          "ms_per_frame": sts[0].ms_render}), flush=True)
# synchronous calls with a moving camera and NO barrier between frames: a peer that runs ahead must
# not overwrite the frame rank 0 is still reading (double-buffered halves + ack)
got_all = []
for fr in frames:
    _, u8, _ = r.render(fr, want_f32=False, want_rgba8=rank == 0)
    if rank == 0:
        got_all.append(u8.copy())
if rank == 0:
    eq = [bool(np.array_equal(got_all[i], want_all[i])) for i in range(n)]
    bad += not all(eq)
    print(json.dumps({"case": "scene1 24 synchronous frames, moving camera, no barrier", "ranks": world, "frames_equal": sum(eq), "of": n}), flush=True)
flag = torch.tensor([bad], device="cuda"); dist.broadcast(flag, src=0)
dist.destroy_process_group()
sys.exit(1 if int(flag.item()) else 0)
