#!/usr/bin/env python
"""bench.py -- headline benchmark of the per-pixel render path (BASELINE.json metric:
Mrays/s and ms/frame beside the CPU reference renderer).

Workload (N = 1): BASELINE config 2 = bundled scene1 (1 plane, 1 triangle, 3 quadrics, 1 light)
at 1920x1080, brute-force intersection, shadow + reflection/refraction code enabled (depth 2).
N > 1 ("weak" scaling): the same scene with the image height multiplied by N, sharded in
interleaved 16-row bands over N B200 (one process per GPU), framebuffer gathered to rank 0
with NCCL inside rtb200_render.

One "step" = one frame.  `value` = Mrays/s with the scene resident in HBM (device-timed, max
over ranks); `e2e` = the same metric through the C ABI with host buffers: every step uploads
the packed scene from pinned host memory (rtb200_set_scene) and reads the RGBA8 frame back
into pinned host memory.  A ray = one nearest-hit query or one shadow-filter query
(SURVEY.md 8d).  `--impl reference` times the reference's own CPU renderer instead.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

FLOP_TRI, FLOP_QUAD, FLOP_PLANE = 46.0, 65.0, 12.0          # SURVEY.md 8(d) per-test figures
SCENE_DAT = os.path.join(ROOT, "tests", "golden", "scenes", "scene1.dat")
SCENE_NPZ = os.path.join(ROOT, "tests", "golden", "packed", "scene1.npz")
BASE_W, BASE_H, DEPTH = 1920, 1080, 2


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


def count_rays_cpu(rt, ps, W, H):
    """Ray count of the workload by the oracle's own recursion (same definition as the GPU's)."""
    from oracle.portbind import PortOracle
    port = PortOracle(rt.abi)
    fr = rt.make_frame(ps.frame, W, H, max_bounces=DEPTH)
    port.render(ps, fr, threads=port.max_threads)
    c = port.ray_counts()
    return c["primary"] + c["shadow"] + c["reflect"] + c["refract"], c


def cpu_reference_arm(rt, ps, W, H, steps, warmup, all_cores):
    """Times the reference's own CPU renderer (oracle/_ref, compiled from the unmodified
    sources) on this box's host cores; falls back to the C port only if _ref is absent."""
    from oracle import refbind
    n_rays, _ = count_rays_cpu(rt, ps, W, H)
    cores = os.cpu_count() or 1
    if refbind.available("rec_ndebug") and os.path.exists(SCENE_DAT):
        ref = refbind.RefRenderer("rec_ndebug")
        kind = "reference"
        times = []
        for i in range(warmup + steps):
            ref.load(SCENE_DAT, max_bounces=DEPTH)
            if all_cores:
                ref.prepare(W, H)
                _, secs = ref.render_region(W, H, 1, 0, 0, W, H, threads=cores)
            else:
                _, secs = ref.render_full(W, H)        # canonical: CScene::LancerRayons(), 1 thread
            if i >= warmup:
                times.append(secs)
        ref.free()
    else:
        from oracle.portbind import PortOracle
        port = PortOracle(rt.abi)
        kind = "port"
        fr = rt.make_frame(ps.frame, W, H, max_bounces=DEPTH)
        times = []
        for i in range(warmup + steps):
            _, secs = port.render(ps, fr, threads=cores if all_cores else 1)
            if i >= warmup:
                times.append(secs)
    secs = float(np.mean(times))
    return {"value": n_rays / secs / 1e6, "unit": "Mrays/s", "cores": cores if all_cores else 1, "kind": kind,
            "ms_per_frame": secs * 1e3,
            "sample": f"full {W}x{H} frame of scene1.dat, depth {DEPTH}, {len(times)} frame(s), "
                      + ("OpenMP rows over the reference's ObtenirCouleur" if all_cores else "CScene::LancerRayons() as shipped, single thread")}


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    rt = load_package()
    ps = rt.PackedScene.load(SCENE_NPZ)
    W, H = BASE_W, BASE_H * max(1, args.gpus)
    base = cpu_reference_arm(rt, ps, W, H, max(1, args.steps), max(0, args.warmup), all_cores=True)
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": base["value"], "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": base["ms_per_frame"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "bundled scene file",
        "config": {"workload": f"scene1.dat {W}x{H} depth {DEPTH} brute force, reference CPU renderer"},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def run_gpu(args):
    rank, world, local_rank = dist_env()
    n = max(1, args.gpus)
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group(backend="cpu:gloo,cuda:nccl")
    rt = load_package()
    ps = rt.PackedScene.load(SCENE_NPZ)
    W, H = BASE_W, BASE_H * world
    r = rt.Renderer(local_rank)
    if world > 1:
        uid = [r.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        r.comm_init(rank, world, uid[0])
    r.set_scene(ps)
    fr = rt.make_frame(ps.frame, W, H, max_bounces=DEPTH)
    peak_tf, peak_mhz = r.fp32_peak()

    def barrier():
        if dist is not None:
            dist.barrier()

    # pinned host buffers for the end-to-end leg
    out8 = rt.pinned_empty((H, W, 4), np.uint8)

    # ---- device-resident leg: scene already in HBM, frame stays on the device ----------------
    for _ in range(max(3, args.warmup)):
        r.render(fr, want_f32=False, want_rgba8=False)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.25)
    ms, kms, launches = [], [], 0
    st = None
    for _ in range(args.steps):
        _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
        ms.append(st.ms_render + st.ms_gather)
        kms.append(st.ms_trace_nearest)
        launches += st.kernel_launches
    # keep the GPU under load long enough for the clock sampler to see it
    t_end = time.time() + 1.0
    while time.time() < t_end:
        r.render(fr, want_f32=False, want_rgba8=False)
    barrier()
    clocks = sampler.stop()
    ms_step = float(np.mean(ms))
    rays_local = float(st.rays)
    # ---- end-to-end leg: H2D scene upload + render + D2H of the RGBA8 frame, every step --------
    for _ in range(2):
        r.set_scene(ps)
        r.render(fr, want_f32=False, want_rgba8=True, out_rgba8=out8)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r.set_scene(ps)
        r.render(fr, want_f32=False, want_rgba8=(rank == 0), out_rgba8=out8)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps

    # ---- max over ranks / sum of rays --------------------------------------------------------
    rays_total, ms_max, e2e_max = rays_local, ms_step, e2e_ms
    if dist is not None:
        import torch
        t = torch.tensor([ms_step, e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_max, e2e_max = float(t[0]), float(t[1])
        s = torch.tensor([rays_local], dtype=torch.float64, device="cuda")
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        rays_total = float(s[0])
    if rank != 0:
        r.close()
        if dist is not None:
            dist.destroy_process_group()
        return 0

    flops = st.tests_triangle * FLOP_TRI + st.tests_quadric * FLOP_QUAD + st.tests_plane * FLOP_PLANE
    k_ms = float(np.mean(kms))
    achieved = flops / (k_ms * 1e-3) / 1e12 if k_ms > 0 else 0.0
    info = r.device_info()
    line = {
        "metric": "Mrays/s", "value": rays_total / (ms_max * 1e-3) / 1e6, "unit": "Mrays/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_max,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "bundled scene1 (packed fixture generated from the reference's own parser + Pretraitement)",
        "config": {"workload": f"C2: scene1.dat {W}x{H} depth {DEPTH}, brute force, shadow + reflection/refraction enabled",
                   "surfaces": int(sum(ps.counts[:3])), "lights": int(ps.counts[3]), "spp": 1,
                   "l2_policy": "frame buffers (33 MB f32+rgba8 per 1080p frame) rewritten every step; scene is 0.6 KB and "
                                "lives in shared memory, so there is no L2-resident input to flush",
                   "parallelism": f"image bands x{world}", "arithmetic": "reference rounding sequence (bit-exact mode)"},
        "rays_per_frame": rays_total, "ms_per_frame": ms_max,
        "e2e": {"value": rays_total / (e2e_max * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_max,
                "h2d_bytes_per_step": int(ps.nbytes()), "d2h_bytes_per_step": int(out8.nbytes)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "fp32_fma", "kernel": "k_resident_level<reference-rounding, primary>",
                     "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": achieved / peak_tf if peak_tf else None, "traffic": None,
                     "kernel_ms": k_ms, "algorithmic_flop_per_launch": flops,
                     "peak_source": f"measured: FFMA chain on all {info['sm_count']} SMs (rtb200_measure_fp32_peak), "
                                    f"ran at {peak_mhz:.0f} MHz; MEASURED_PEAKS.json carries no FP32 figure"},
        "device": info,
    }
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_reference_arm(rt, ps, W, H, steps=6, warmup=1, all_cores=False)
        line["cpu_baseline_all_cores"] = cpu_reference_arm(rt, ps, W, H, steps=6, warmup=1, all_cores=True)
    print(json.dumps(line))
    r.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="rtb200", choices=["rtb200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
