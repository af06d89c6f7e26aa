#!/usr/bin/env python
"""bench.py -- headline benchmark of the per-pixel render path (BASELINE.json metric:
Mrays/s and ms/frame beside the CPU reference renderer).

Workloads (`--workload`, default c2 = the configuration the contract names for N = 1):
  c2  BASELINE config 2: bundled scene1 (1 plane, 1 triangle, 3 quadrics, 1 light) at 1920x1080,
      brute force, shadow + reflection/refraction code enabled (depth 2).   [scene-resident kernels]
  c3  config 3: seeded soup of 10 000 triangles + 64 quadrics + ground plane, 4 lights, 3840x2160,
      depth 3, brute-force shared-memory tiled (TMA).                        [streamed kernels]
  c4  config 4: seeded 1 M-triangle height field + ground plane, 8 lights, 3840x2160, depth 4,
      GPU-built LBVH.
  c5  config 5: 100 k-triangle height field, 4 lights, depth 2, 8192x8192 at 16 spp (4x4 grid),
      LBVH, image bands sharded over the GPUs, NCCL framebuffer gather.
N > 1: one process per GPU (torchrun); the image is sharded in interleaved 16-row bands, the scene
is replicated, the framebuffer is gathered to rank 0 with NCCL inside rtb200_render.  c2-c4 scale
WEAKLY (image height x N), c5 STRONGLY (the config fixes the image).

One "step" = one frame.  The timed region is K frames bracketed by a barrier and a device
synchronisation on both sides, timed with CUDA events on the library's own stream
(rtb200_timer_mark), max over ranks; the K frames are submitted as one camera path through
rtb200_render_sequence (pipelined where the path allows it -- config 2 --, frame by frame
otherwise; `ms_per_frame_synchronous_calls` is the same loop through K synchronous rtb200_render
calls).  `value` = rays of the whole job / that time, scene resident in HBM.  `e2e` = the same through the C ABI with HOST buffers: every step uploads the packed scene
from host memory (rtb200_set_scene, plus rtb200_build_bvh for the LBVH workloads) and reads the
RGBA8 frame back into pinned host memory.  A ray = one nearest-hit query or one shadow-filter
query.  `--impl reference` times the reference's own CPU renderer instead.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from importlib import import_module

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

FLOP_TRI, FLOP_QUAD, FLOP_PLANE = 46.0, 65.0, 12.0          # SURVEY.md 8(d) per-test figures
NODE_BYTES, TRI_BYTES = 64.0, 48.0                          # LBVH traversal byte model
SCENE_DAT = os.path.join(ROOT, "tests", "golden", "scenes", "scene1.dat")
SCENE_NPZ = os.path.join(ROOT, "tests", "golden", "packed", "scene1.npz")


def _workloads(rt):
    synth = import_module("rtb200.synth")
    return {
        "c2": dict(name="C2: scene1.dat {W}x{H} depth 2, brute force, shadow + reflection/refraction enabled",
                   scene=lambda: rt.PackedScene.load(SCENE_NPZ), W=1920, H=1080, depth=2, spp=1, flags=0, bvh=False,
                   scaling="weak", bound="fp32", steps=50, crop=None, dat=SCENE_DAT,
                   data="bundled scene1 (packed fixture generated from the reference's own parser + Pretraitement)"),
        "c3": dict(name="C3: synthetic soup 10k tris + 64 quadrics + plane, 4 lights, {W}x{H} depth 3, brute force tiled",
                   scene=lambda: synth.soup(10000, 64, 4, seed=1234), W=3840, H=2160, depth=3, spp=1, flags=rt.F_FORCE_STREAM,
                   bvh=False, scaling="weak", bound="fp32", steps=5, crop=(256, 144), dat=None,
                   data="synthetic, seed 1234 (rtb200.synth.soup)"),
        "c4": dict(name="C4: synthetic 1M-triangle height field + plane, 8 lights, {W}x{H} depth 4, GPU-built LBVH",
                   scene=lambda: synth.heightfield(1000000, 1000.0, 8, seed=4321), W=3840, H=2160, depth=4, spp=1,
                   flags=rt.F_USE_BVH, bvh=True, scaling="weak", bound="hbm", steps=20, crop=(48, 27), dat=None,
                   data="synthetic, seed 4321 (rtb200.synth.heightfield)"),
        "c5": dict(name="C5: synthetic 100k-triangle height field + plane, 4 lights, {W}x{H} at 16 spp, depth 2, LBVH, bands over the GPUs",
                   scene=lambda: synth.heightfield(100000, 350.0, 4, seed=5), W=8192, H=8192, depth=2, spp=4,
                   flags=rt.F_USE_BVH, bvh=True, scaling="strong", bound="hbm", steps=3, crop=(32, 32), dat=None,
                   data="synthetic, seed 5 (rtb200.synth.heightfield)"),
    }


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))


def ncu_traffic(workload):
    """DRAM bytes of ONE launch of the workload's dominant kernel, from the committed ncu capture."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")))[workload]
        return float(t["dram_read_bytes"] + t["dram_write_bytes"]), t["source"]
    except Exception:
        return None, "no ncu --set full capture of this workload's kernel committed"


def hbm_peak():
    """Measured HBM copy bandwidth of this pool's B200 (driver-written), else the profiling guide's fallback."""
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6500.0, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's own renderer (oracle/_ref) on the host cores of this box
# ---------------------------------------------------------------------------------------------
def cpu_reference_arm(rt, wl, ps, W, H, steps, warmup, all_cores):
    """Times the reference's CPU renderer (compiled from the unmodified sources; the C port only
    if _ref is absent).  c2: full frames.  c3-c5: a centred crop of the frame (bounded sample),
    rays counted by the oracle on exactly that crop, ms/frame extrapolated by pixel count."""
    from oracle import refbind
    from oracle.portbind import PortOracle
    port = PortOracle(rt.abi)
    cores = os.cpu_count() or 1
    threads = cores if all_cores else 1
    depth, k = wl["depth"], wl["spp"]
    fr = rt.make_frame(ps.frame, W, H, max_bounces=depth, spp_grid=k)
    crop = wl["crop"]
    if crop is None:
        cw, ch, x0, y0 = W, H, 0, 0
    else:
        cw, ch = crop
        x0, y0 = (W - cw) // 2, (H - ch) // 2
    # the oracle's region entry points work on the k x k sample lattice (k*W x k*H primary rays)
    lx0, ly0, lw, lh = x0 * k, y0 * k, cw * k, ch * k
    use_ref = refbind.available("rec_ndebug")
    times = []
    if use_ref:
        ref = refbind.RefRenderer("rec_ndebug")
        kind = "reference"
        for i in range(warmup + steps):
            if wl["dat"]:
                ref.load(wl["dat"], max_bounces=depth)
            elif i == 0:
                ref.new_scene(camera=(ps.frame["camera_pos"], ps.frame["camera_look"], ps.frame["camera_up"]),
                              background=ps.frame["background"], max_bounces=depth)
                ref.add_packed(ps)
            if crop is None and not all_cores:
                _, secs = ref.render_full(W, H)            # canonical: CScene::LancerRayons(), 1 thread
            else:
                if wl["dat"] or i == 0:
                    ref.prepare(W, H)
                _, secs = ref.render_region(W, H, k, lx0, ly0, lw, lh, threads=threads)
            if i >= warmup:
                times.append(secs)
        ref.free()
    else:
        kind = "port"
        for i in range(warmup + steps):
            _, secs = port.render_region(ps, fr, lx0, ly0, lw, lh, threads=threads)
            if i >= warmup:
                times.append(secs)
    # rays of exactly the timed sample, by the oracle's own recursion (same definition as the GPU's)
    port.render_region(ps, fr, lx0, ly0, lw, lh, threads=cores)
    c = port.ray_counts()
    n_rays = c["primary"] + c["shadow"] + c["reflect"] + c["refract"]
    secs = float(np.mean(times))
    scale = (W * H) / float(cw * ch)
    how = ("OpenMP rows over the reference's ObtenirCouleur" if (all_cores or crop is not None)
           else "CScene::LancerRayons() as shipped, single thread")
    sample = (f"full {W}x{H} frame" if crop is None else f"centred {cw}x{ch}-pixel crop of the {W}x{H} frame ({k * k} spp)")
    return {"value": n_rays / secs / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
            "ms_per_frame": secs * 1e3 * scale, "extrapolated": crop is not None, "sample_seconds": secs, "sample_rays": n_rays,
            "sample": f"{sample}, depth {depth}, {len(times)} run(s), {how}"}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    rt = load_package()
    wl = _workloads(rt)[args.workload]
    ps = wl["scene"]()
    n = max(1, args.gpus)
    W, H = wl["W"], wl["H"] * (n if wl["scaling"] == "weak" else 1)
    steps = max(1, args.steps if wl["crop"] is None else min(args.steps, 3))
    base = cpu_reference_arm(rt, wl, ps, W, H, steps, max(0, min(args.warmup, 1)), all_cores=True)
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": base["value"], "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": base["sample_seconds"] * 1e3,
        "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f32", "data": wl["data"],
        "config": {"workload": wl["name"].format(W=W, H=H) + ", reference CPU renderer"},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def run_gpu(args):
    rank, world, local_rank = dist_env()
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group(backend="cpu:gloo,cuda:nccl")
    rt = load_package()
    wl = _workloads(rt)[args.workload]
    steps = args.steps if args.steps > 0 else wl["steps"]
    warmup = max(3, args.warmup)
    ps = wl["scene"]()
    W, H = wl["W"], wl["H"] * (world if wl["scaling"] == "weak" else 1)
    r = rt.Renderer(local_rank)
    if world > 1:
        uid = [r.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        r.comm_init(rank, world, uid[0])
    r.set_scene(ps)
    bvh_info = r.build_bvh() if wl["bvh"] else None
    fr = rt.make_frame(ps.frame, W, H, max_bounces=wl["depth"], spp_grid=wl["spp"], flags=wl["flags"])
    peak_tf, peak_mhz = r.fp32_peak()

    def barrier():
        if dist is not None:
            dist.barrier()

    out8 = rt.pinned_empty((H, W, 4), np.uint8)      # pinned host frame for the end-to-end leg

    # ---- device-resident leg: scene already in HBM, frame stays on the device ----------------
    for _ in range(warmup):
        r.render(fr, want_f32=False, want_rgba8=False)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.25)
    barrier()
    dev_ms, prim_ms, launches = [], [], 0
    stage = np.zeros(3)
    st = None
    # frame-by-frame through the synchronous call (per-kernel device times for the roofline) ...
    r.timer_mark(2)
    for _ in range(steps):
        _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
        dev_ms.append(st.ms_render + st.ms_gather)
        prim_ms.append(st.ms_primary_kernel)
        stage += (st.ms_trace_nearest, st.ms_trace_shadow, st.ms_shade)
        launches += st.kernel_launches
    r.timer_mark(3)
    sync_loop_ms = r.timer_elapsed(2, 3) / steps
    barrier()
    # ... and the K timed steps as ONE submission (rtb200_render_sequence: the frames of a camera
    # path enqueued back to back where the path allows it, else the same frame-by-frame loop)
    r.render_sequence([fr] * min(steps, 3), want_stats=False)
    barrier()
    r.timer_mark(0)
    _, seq_stats = r.render_sequence([fr] * steps)
    r.timer_mark(1)
    loop_ms = r.timer_elapsed(0, 1) / steps                 # K frames as ONE device-timed region
    assert all(s.rays_primary == st.rays_primary and s.rays_shadow == st.rays_shadow for s in seq_stats)
    barrier()
    # keep the GPU under load long enough for the clock sampler to see it.  Every frame contains a
    # collective at N > 1, so all ranks must run the SAME number of extra frames: it is derived from
    # the max-over-ranks frame time, not from each rank's own clock.
    frame_ms = loop_ms
    if dist is not None:
        import torch
        tmax = torch.tensor([loop_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        frame_ms = float(tmax[0])
    n_load = int(min(max(1000.0 / max(frame_ms, 1e-3), 1), 20000))
    for i in range(0, n_load, 256):
        r.render_sequence([fr] * min(256, n_load - i), want_stats=False)
    barrier()
    clocks = sampler.stop()
    rays_local = float(st.rays)
    # streamed path: the tiled kernels cull whole triangles per warp (rt_trace.cu, "warp bundles"); the FP32 roofline
    # of the north-star is about the brute-force scan itself, so that kernel is timed too, with the culling off
    # (two extra frames on every rank, outside the timed region)
    brute = None
    if st.path == 1:
        prev = os.environ.get("RTB200_NO_BUNDLE_CULL")
        os.environ["RTB200_NO_BUNDLE_CULL"] = "1"
        r.render(fr, want_f32=False, want_rgba8=False)
        _, _, sb = r.render(fr, want_f32=False, want_rgba8=False)
        if prev is None:
            os.environ.pop("RTB200_NO_BUNDLE_CULL")
        else:
            os.environ["RTB200_NO_BUNDLE_CULL"] = prev
        brute = {"kernel_ms": sb.ms_primary_kernel, "frame_ms": sb.ms_render + sb.ms_gather,
                 "stage_ms": {"trace_nearest": sb.ms_trace_nearest, "trace_shadow": sb.ms_trace_shadow, "shade": sb.ms_shade}}
        barrier()
    # ---- end-to-end leg: H2D scene upload (+ LBVH build) + render + D2H of the RGBA8 frame, every step
    e2e_steps = max(1, min(steps, 20))
    for _ in range(2):
        r.set_scene(ps)
        if wl["bvh"]:
            r.build_bvh()
        r.render(fr, want_f32=False, want_rgba8=True, out_rgba8=out8)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        r.set_scene(ps)
        if wl["bvh"]:
            r.build_bvh()
        r.render(fr, want_f32=False, want_rgba8=(rank == 0), out_rgba8=out8)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    # informational: the same frames as ONE camera-path submission with a host destination (scene
    # uploaded once per submission, every RGBA8 frame copied to pinned host memory while the next renders)
    e2e_seq_ms, seq_n = None, 0
    if args.workload == "c2":
        seq_n = 8
        outN = rt.pinned_empty((seq_n, H, W, 4), np.uint8) if rank == 0 else None
        for rep in range(3):
            barrier()
            t0 = time.perf_counter()
            r.set_scene(ps)
            r.render_sequence([fr] * seq_n, out_rgba8=outN, want_stats=False)
            barrier()
            e2e_seq_ms = (time.perf_counter() - t0) * 1e3 / seq_n
        if rank == 0:
            assert np.array_equal(outN[0], out8) and np.array_equal(outN[seq_n - 1], out8)

    # ---- max over ranks / sum of rays --------------------------------------------------------
    rays_total, ms_max, e2e_max, dev_max = rays_local, loop_ms, e2e_ms, float(np.mean(dev_ms))
    if dist is not None:
        import torch
        t = torch.tensor([loop_ms, e2e_ms, dev_max, e2e_seq_ms or 0.0], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_max, e2e_max, dev_max = float(t[0]), float(t[1]), float(t[2])
        if e2e_seq_ms is not None:
            e2e_seq_ms = float(t[3])
        s = torch.tensor([rays_local], dtype=torch.float64, device="cuda")
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        rays_total = float(s[0])
    if rank != 0:
        r.close()
        if dist is not None:
            dist.destroy_process_group()
        return 0

    info = r.device_info()
    T, Q, P, L = ps.counts
    traffic, traffic_src = ncu_traffic(args.workload)
    if wl["bound"] == "fp32":
        # dominant kernel: the level-0 nearest-hit kernel (c3) / the fused level-0 kernel (c2)
        k_ms = float(np.mean(prim_ms))
        frame_ms_roof = dev_max
        if brute is not None:
            k_ms, frame_ms_roof = brute["kernel_ms"], brute["frame_ms"]
        if st.path == 0:
            flops = st.tests_triangle * FLOP_TRI + st.tests_quadric * FLOP_QUAD + st.tests_plane * FLOP_PLANE
            kernel = "k_resident_level<reference-rounding, primary> (trace + shade + shadow fused)"
        else:
            flops = st.rays_primary * (T * FLOP_TRI + Q * FLOP_QUAD + P * FLOP_PLANE)
            kernel = "k_trace_nearest<filter + reference-rounded survivors, primary>, plain scan (RTB200_NO_BUNDLE_CULL=1)"
        achieved = flops / (k_ms * 1e-3) / 1e12 if k_ms > 0 else 0.0
        frame_flops = st.tests_triangle * FLOP_TRI + st.tests_quadric * FLOP_QUAD + st.tests_plane * FLOP_PLANE
        roof = {"bound": "fp32_fma", "kernel": kernel, "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": achieved / peak_tf if peak_tf else None, "traffic": traffic, "traffic_source": traffic_src, "kernel_ms": k_ms,
                "algorithmic_flop_per_launch": flops,
                "whole_frame": {"flop": frame_flops, "TFLOP/s": frame_flops / (frame_ms_roof * 1e-3) / 1e12,
                                "frac": frame_flops / (frame_ms_roof * 1e-3) / 1e12 / peak_tf if peak_tf else None},
                "peak_source": f"measured: FFMA chain on all {info['sm_count']} SMs (rtb200_measure_fp32_peak), "
                               f"ran at {peak_mhz:.0f} MHz; MEASURED_PEAKS.json carries no FP32 figure"}
        if brute is not None:
            executed = st.tests_triangle_executed / max(st.tests_triangle, 1)
            roof["note"] = ("achieved / frac / whole_frame are the PLAIN brute-force scan (every ray against every surface, "
                            "warp-bundle culling off: two extra frames outside the timed region); `value` and ms_per_step are "
                            "the default path, whose warps skip the triangles outside the cone of their rays")
            roof["plain_scan"] = brute
            roof["with_bundle_culling"] = {"kernel_ms": float(np.mean(prim_ms)), "frame_ms": dev_max,
                                           "triangle_tests_executed_frac": executed,
                                           "algorithmic_TFLOP/s": frame_flops / (dev_max * 1e-3) / 1e12}
    else:
        # dominant kernel: the LBVH shadow-filter traversal (most rays are shadow queries)
        peak, src = hbm_peak()
        k_ms = stage[1] / steps
        byts = st.shadow_bvh_nodes * NODE_BYTES + st.shadow_tests_triangle * TRI_BYTES
        achieved = byts / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
        all_b = st.bvh_nodes_visited * NODE_BYTES + st.tests_triangle * TRI_BYTES
        roof = {"bound": "hbm", "kernel": "k_bvh_shadow (all launches of a frame)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "kernel_ms": k_ms, "algorithmic_bytes_per_frame": byts,
                "note": "bytes REQUESTED by the traversal (nodes x 64 B + leaf triangles x 48 B); the working set "
                        "(nodes + triangles) largely lives in the 126 MB L2, so a fraction above 1 means L2 hits, not HBM traffic",
                "whole_frame": {"bytes": all_b, "GB/s": all_b / (dev_max * 1e-3) / 1e9,
                                "nodes_per_ray": st.bvh_nodes_visited / max(st.rays, 1), "tri_tests_per_ray": st.tests_triangle / max(st.rays, 1)},
                "peak_source": src}
    line = {
        "metric": "Mrays/s", "value": rays_total / (ms_max * 1e-3) / 1e6, "unit": "Mrays/s",
        "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_max,
        "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f32", "data": wl["data"],
        "config": {"workload": wl["name"].format(W=W, H=H), "triangles": T, "quadrics": Q, "planes": P, "lights": L,
                   "spp": wl["spp"] ** 2, "depth": wl["depth"],
                   "l2_policy": ("frame buffers (33 MB f32+rgba8 per 1080p frame) rewritten every step; the scene is 0.6 KB and "
                                 "lives in shared memory, so there is no L2-resident input to flush") if args.workload == "c2" else
                                ("per-frame ray / hit / shadow-query buffers (hundreds of MB, larger than the 126 MB L2) are rewritten "
                                 "every step; the scene itself is meant to stay cache resident, as in production"),
                   "parallelism": (f"interleaved 16-row image bands x{world}, scene replicated; the producing kernels store into rank 0's frame "
                                   "over NVLink and signal completion themselves (peer frame), NCCL gather as fallback") if world > 1 else
                                  "one GPU (image bands x1)",
                   "arithmetic": "reference rounding sequence (bit-exact mode)",
                   "switches": {k: v for k, v in os.environ.items() if k.startswith("RTB200_")}},
        "rays_per_frame": rays_total, "ms_per_frame": ms_max, "device_ms_per_frame": dev_max,
        "ms_per_frame_synchronous_calls": sync_loop_ms,
        "stage_ms_per_frame": {"trace_nearest": stage[0] / steps, "trace_shadow": stage[1] / steps, "shade": stage[2] / steps},
        "rays": {"primary": st.rays_primary, "shadow": st.rays_shadow, "reflect": st.rays_reflect, "refract": st.rays_refract, "rank": 0},
        "e2e": {"value": rays_total / (e2e_max * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_max,
                "h2d_bytes_per_step": int(ps.nbytes()), "d2h_bytes_per_step": int(out8.nbytes), "steps": e2e_steps,
                "includes": "rtb200_set_scene" + (" + rtb200_build_bvh" if wl["bvh"] else "") + " + rtb200_render + RGBA8 read-back"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof,
        "device": info,
    }
    if e2e_seq_ms is not None:
        line["e2e_sequence"] = {"value": rays_total / (e2e_seq_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_seq_ms, "frames": seq_n,
                                "includes": "rtb200_set_scene once + rtb200_render_sequence with a pinned host destination (every frame copied out, "
                                            "overlapped with the next frame's kernel); not the contract's e2e, which uploads the scene every step"}
    if bvh_info:
        line["bvh"] = bvh_info
    if world == 1 and not args.no_cpu:
        if args.workload == "c2":
            line["cpu_baseline"] = cpu_reference_arm(rt, wl, ps, W, H, steps=6, warmup=1, all_cores=False)
            line["cpu_baseline_all_cores"] = cpu_reference_arm(rt, wl, ps, W, H, steps=6, warmup=1, all_cores=True)
        else:
            line["cpu_baseline"] = cpu_reference_arm(rt, wl, ps, W, H, steps=1, warmup=0, all_cores=True)
    print(json.dumps(line))
    r.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed frames (0 = the workload's default)")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="rtb200", choices=["rtb200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps <= 0:
            args.steps = 3
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
