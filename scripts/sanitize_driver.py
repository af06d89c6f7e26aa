#!/usr/bin/env python
"""Small frames through every path of the library, meant to run under compute-sanitizer (memcheck / racecheck /
synccheck): resident kernels with the dependent-launch chain, a pipelined sequence, the plain tiled kernels (TMA
double buffering), the culled scan (per-warp queues and staging), the LBVH (build + walk + far rays), scene packing on
the device, the fix-up pass.  With RANK/WORLD_SIZE set (torchrun --no-python compute-sanitizer ... python this): the
peer-frame and host-band paths on several GPUs."""
import os
import sys
from importlib import import_module

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

rt = load_package()
synth = import_module("rtb200.synth")
world = int(os.environ.get("WORLD_SIZE", 1))
rank = int(os.environ.get("RANK", 0))
local = int(os.environ.get("LOCAL_RANK", 0))
dist = None
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group(backend="gloo")
r = rt.Renderer(local)
if world > 1:
    uid = [r.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    r.comm_init(rank, world, uid[0])


def scene(n):
    return rt.PackedScene.load(os.path.join(ROOT, "tests", "golden", "packed", f"scene{n}.npz"))


def say(what, st):
    if rank == 0:
        print(f"[sanitize] {what}: path {st.path}, {st.rays} rays, {st.kernel_launches} launches", flush=True)


want = rank == 0
# resident path: direct frame, level loop with dependent launches (twice: the second frame takes the guessed capacities + tail levels)
ps = scene(1); r.set_scene(ps)
_, _, st = r.render(rt.make_frame(ps.frame, 96, 64, max_bounces=2), want_f32=want, want_rgba8=want); say("scene1 resident", st)
ps = scene(7); r.set_scene(ps)
for _ in range(2):
    _, _, st = r.render(rt.make_frame(ps.frame, 96, 64, max_bounces=20), want_f32=want, want_rgba8=want)
say("scene7 resident depth 20 (async level loop)", st)
ps = scene(1); r.set_scene(ps)
frames = []
for i in range(4):
    fr = rt.make_frame(ps.frame, 96, 64, max_bounces=2); fr.cam_pos[0] += 2.0 * i; frames.append(fr)
r.render_sequence(frames, out_rgba8=(rt.pinned_empty((4, 64, 96, 4), np.uint8) if (want and world == 1) else None))
if rank == 0:
    print("[sanitize] scene1 pipelined sequence", flush=True)
# streamed paths on a soup: culled scan (default), plain tiled scan, exact mode; supersampling; device-side packing (>= 2048 triangles)
ps = synth.soup(2200, 12, 3, seed=5); r.set_scene(ps)
for name, flags, env in (("culled scan", rt.F_FORCE_STREAM, None), ("plain tiled scan", rt.F_FORCE_STREAM, "RTB200_NO_BUNDLE_CULL"),
                         ("exact mode", rt.F_FORCE_STREAM | rt.F_EXACT_FP, None)):
    if env:
        os.environ[env] = "1"
    _, _, st = r.render(rt.make_frame(ps.frame, 80, 48, max_bounces=2, spp_grid=2 if name == "culled scan" else 1, flags=flags), want_f32=want, want_rgba8=want)
    if env:
        os.environ.pop(env)
    say(f"soup streamed, {name}", st)
if world == 1:
    rng = np.random.default_rng(1)
    o = np.tile(np.asarray(ps.frame["cam_pos"], np.float32), (500, 1)); d = rng.normal(size=(500, 3)).astype(np.float32)
    r.trace_nearest(o, d, flags=rt.F_FORCE_STREAM); r.trace_shadow(o + d * 50, o, flags=rt.F_FORCE_STREAM)
    print("[sanitize] ray queries through the streamed kernels", flush=True)
# LBVH: build, walk, bounded quadrics in the tree
r.build_bvh()
_, _, st = r.render(rt.make_frame(ps.frame, 80, 48, max_bounces=2, flags=rt.F_USE_BVH), want_f32=want, want_rgba8=want); say("soup LBVH", st)
hf = synth.heightfield(3000, 200.0, 2, seed=2); r.set_scene(hf); r.build_bvh()
_, _, st = r.render(rt.make_frame(hf.frame, 80, 48, max_bounces=2, flags=rt.F_USE_BVH), want_f32=want, want_rgba8=want); say("height field LBVH", st)
if world > 1:
    # every rank writes its bands into one shared host frame
    from multiprocessing import shared_memory
    name = [None]
    shm = shared_memory.SharedMemory(create=True, size=64 * 96 * 4) if rank == 0 else None
    name[0] = shm.name if rank == 0 else None
    dist.broadcast_object_list(name, src=0)
    if rank != 0:
        shm = shared_memory.SharedMemory(name=name[0])
    arr = np.ndarray((64, 96, 4), np.uint8, buffer=shm.buf)
    r.host_register(arr)
    ps = scene(1); r.set_scene(ps)
    r.render(rt.make_frame(ps.frame, 96, 64, max_bounces=2, flags=rt.F_HOST_BANDS), want_f32=False, out_rgba8=arr)
    dist.barrier()
    r.host_unregister(arr)
    if rank == 0:
        print("[sanitize] host bands on", world, "GPUs", flush=True)
    del arr
    shm.close()
    dist.barrier()
    if rank == 0:
        shm.unlink()
r.close()
if dist is not None:
    dist.destroy_process_group()
if rank == 0:
    print("[sanitize] done", flush=True)
