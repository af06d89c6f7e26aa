#!/usr/bin/env python
"""bench.py -- headline benchmark of the per-pixel render path (BASELINE.json metric:
Mrays/s and ms/frame, 1080p-8K, at 1/2/4/8 B200 beside the CPU reference renderer).

Workloads (BASELINE.json `configs`, recipes in SURVEY.md 8d):
  c1  config 1: bundled scene1 at 640x480 on the reference's own C++ CPU renderer, single thread,
      as shipped (CScene::LancerRayons, Scene.cpp:1535-1563).  No GPU arm; a block of the default line.
  c2  config 2: the same scene at 1920x1080, brute force, shadow + reflection/refraction code
      enabled (depth 2).                                                    [scene-resident kernels]
  c3  config 3: seeded soup of 10 000 triangles + 64 quadrics + ground plane, 4 lights, 3840x2160,
      depth 3, brute-force shared-memory tiled (TMA).                        [streamed kernels]
  c4  config 4: seeded 1 M-triangle height field + ground plane, 8 lights, 3840x2160, depth 4,
      GPU-built LBVH.
  c5  config 5: 100 k-triangle height field, 4 lights, depth 2, 8192x8192 at 16 spp (4x4 grid),
      LBVH, image bands sharded over the GPUs.
Which one is the headline of the JSON line (`--workload auto`, the default; the rule is printed in
`config.workload_rule`):
  * one process on a one-GPU box: C3 -- the config the north-star's ">= 50 % of the FP32 FMA peak on
    the brute-force intersection kernel" is written on -- with C1, C2, C4 and C5 as blocks of the same
    line (`configs`), each with ms_per_frame / value / e2e / roofline / cpu_baseline / parity;
  * several ranks, or one process on a box with several GPUs (= the N = 1 point of a scaling curve):
    C5 STRONG scaling, the config the north-star's "near-linear 8-GPU scaling at 8K" is written on.
`--workload cK` forces one workload.  c2-c4 scale WEAKLY when forced at N > 1 (image height x N).

One "step" = one frame (config 2, whose frame is 85 us: `frames_per_step` frames, see the line).  The
timed region is K steps bracketed by a barrier and a device synchronisation on both sides, timed with
CUDA events on the library's own stream (rtb200_timer_mark), max over ranks, submitted as one camera
path through rtb200_render_sequence.  `value` = rays of the whole job / that time, scene resident in
HBM.  `e2e` = the same through the C ABI with HOST buffers: every step uploads the packed scene from
host memory (rtb200_set_scene, plus rtb200_build_bvh for the LBVH workloads) and reads the RGBA8 frame
back into pinned host memory; on several GPUs every rank writes its own bands into the caller's frame
(a shared-memory segment) over its own PCIe link.  A ray = one nearest-hit query or one shadow-filter
query.  `parity` = a crop of the frame rendered live by the reference's own CPU renderer
(oracle/_ref) on this box, compared with the same crop of the GPU frame.  `--impl reference` times the
reference's CPU renderer alone on the same config.
"""
import argparse
import json
import math
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
MIN_REGION_MS = 300.0                                        # a timed region shorter than this is repeated frames


def _workloads(rt):
    synth = import_module("rtb200.synth")
    return {
        "c2": dict(name="C2: scene1.dat {W}x{H} depth 2, brute force, shadow + reflection/refraction enabled",
                   scene=lambda: rt.PackedScene.load(SCENE_NPZ), W=1920, H=1080, depth=2, spp=1, flags=0, bvh=False,
                   scaling="weak", bound="fp32", crop=None, ref_crop=None, dat=SCENE_DAT,
                   data="bundled scene1 (packed fixture generated from the reference's own parser + Pretraitement)"),
        "c3": dict(name="C3: synthetic soup 10k tris + 64 quadrics + plane, 4 lights, {W}x{H} depth 3, brute force tiled",
                   scene=lambda: synth.soup(10000, 64, 4, seed=1234), W=3840, H=2160, depth=3, spp=1, flags=rt.F_FORCE_STREAM,
                   bvh=False, scaling="weak", bound="fp32", crop=(256, 144), ref_crop=(160, 90), dat=None,
                   data="synthetic, seed 1234 (rtb200.synth.soup)"),
        "c4": dict(name="C4: synthetic 1M-triangle height field + plane, 8 lights, {W}x{H} depth 4, GPU-built LBVH",
                   scene=lambda: synth.heightfield(1000000, 1000.0, 8, seed=4321), W=3840, H=2160, depth=4, spp=1,
                   flags=rt.F_USE_BVH, bvh=True, scaling="weak", bound="issue", crop=(64, 64), ref_crop=(16, 16), dat=None,
                   data="synthetic, seed 4321 (rtb200.synth.heightfield)"),
        "c5": dict(name="C5: synthetic 100k-triangle height field + plane, 4 lights, {W}x{H} at 16 spp, depth 2, LBVH, bands over the GPUs",
                   scene=lambda: synth.heightfield(100000, 350.0, 4, seed=5), W=8192, H=8192, depth=2, spp=4,
                   flags=rt.F_USE_BVH, bvh=True, scaling="strong", bound="issue", crop=(32, 128), ref_crop=(16, 16), dat=None,
                   data="synthetic, seed 5 (rtb200.synth.heightfield)"),
    }


def physical_gpus():
    """GPUs of the box (not of CUDA_VISIBLE_DEVICES): decides whether an N = 1 run is a point of a scaling curve."""
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=20).stdout
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


def pick_workload(name, world):
    if name != "auto":
        return name, [], f"--workload {name} given"
    if world > 1:
        return "c5", [], "N > 1: config 5 (8192x8192 x 16 spp), STRONG scaling -- the config the north-star scales"
    n = physical_gpus()
    if n > 1:
        return "c5", ["c3"], (f"N = 1 on a box with {n} GPUs is read as the first point of the scaling curve: config 5, as at N > 1, with "
                              "config 3 (the FP32-roofline config) as a block under `configs` "
                              "(on a one-GPU box the headline is config 3 with configs 1, 2, 4, 5 as blocks)")
    return "c3", ["c1", "c2", "c4", "c5"], ("N = 1 on a one-GPU box: config 3 (the brute-force kernel the FP32 target is written on); "
                                            "configs 1, 2, 4 and 5 are blocks of this line under `configs`")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons, sampled every 50 ms; stop(t0, t1) keeps the samples whose
    wall-clock stamp lies inside the timed region (B200_PROFILING.md)."""
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
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append((time.time(), [c.strip() for c in line.split(",")]))
        except Exception:
            pass

    def stop(self, windows):
        if self.proc:
            self.proc.terminate()
        if self.is_alive():
            self.join(timeout=2)

        def inside(t):
            return any(a - 0.03 <= t <= b + 0.03 for a, b in windows)
        rows_in = [r for t, r in self.rows if inside(t)]
        where = "inside the timed region(s)"
        if len(rows_in) < 3:                       # region shorter than the sampling period: all samples of the run
            rows_in = [r for _, r in self.rows]
            where = "whole run (timed region shorter than 3 samples)"
        sm = [float(r[1]) for r in rows_in if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows_in if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in rows_in:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "sampled": where}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))


def ncu_traffic(workload):
    """DRAM bytes of ONE launch of the workload's dominant kernel, from the committed ncu capture."""
    for name in ("r2_ncu_traffic.json", "r1_ncu_traffic.json"):
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", name)))[workload]
            return float(t["dram_read_bytes"] + t["dram_write_bytes"]), t["source"]
        except Exception:
            continue
    return None, "no ncu --set full capture of this workload's kernel committed"


def ncu_counters(workload):
    """Issue-side counters of the LBVH traversal kernels from the committed ncu captures (what binds that kernel)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r2_ncu_lbvh_counters.json")))[workload]
    except Exception:
        return None


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6500.0, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"


# ---------------------------------------------------------------------------------------------
# CPU side: the reference's own renderer (oracle/_ref) on the host cores of this box
# ---------------------------------------------------------------------------------------------
def crop_rect(wl, W, H, crop):
    """(x0, y0, w, h) of the centred crop in PIXELS; None = the full frame."""
    if crop is None:
        return 0, 0, W, H
    cw, ch = min(crop[0], W), min(crop[1], H)
    return (W - cw) // 2, (H - ch) // 2, cw, ch


class CpuRef:
    """The reference CPU renderer holding one workload's scene (built once, rendered many times)."""

    def __init__(self, rt, wl, ps, W, H):
        from oracle import refbind
        from oracle.portbind import PortOracle
        self.rt, self.wl, self.ps, self.W, self.H = rt, wl, ps, W, H
        self.port = PortOracle(rt.abi)
        self.cores = os.cpu_count() or 1
        self.k = wl["spp"]
        self.frame = rt.make_frame(ps.frame, W, H, max_bounces=wl["depth"], spp_grid=self.k)
        self.use_ref = refbind.available("rec_ndebug")
        self.kind = "reference" if self.use_ref else "port"
        self.ref = None
        if self.use_ref:
            self.ref = refbind.RefRenderer("rec_ndebug")
            if wl["dat"]:
                self.ref.load(wl["dat"], max_bounces=wl["depth"])
            else:
                self.ref.new_scene(camera=(ps.frame["camera_pos"], ps.frame["camera_look"], ps.frame["camera_up"]),
                                   background=ps.frame["background"], max_bounces=wl["depth"])
                self.ref.add_packed(ps)
            self.ref.prepare(W, H)

    def region(self, rect, threads):
        """float RGB of the k*k lattice samples of the pixel rectangle and the seconds it took."""
        x0, y0, w, h = rect
        k = self.k
        if self.use_ref:
            return self.ref.render_region(self.W, self.H, k, x0 * k, y0 * k, w * k, h * k, threads=threads)
        return self.port.render_region(self.ps, self.frame, x0 * k, y0 * k, w * k, h * k, threads=threads)

    def full_as_shipped(self):
        """CScene::LancerRayons() itself (single thread, whole frame) -- reloads the scene: Pretraitement is not idempotent."""
        self.ref.load(self.wl["dat"], max_bounces=self.wl["depth"])
        out, secs = self.ref.render_full(self.W, self.H)
        self.ref.load(self.wl["dat"], max_bounces=self.wl["depth"])
        self.ref.prepare(self.W, self.H)
        return out, secs

    def rays(self, rect):
        """rays of exactly that rectangle, by the oracle's own recursion (same definition as the GPU's)"""
        x0, y0, w, h = rect
        k = self.k
        self.port.render_region(self.ps, self.frame, x0 * k, y0 * k, w * k, h * k, threads=self.cores)
        c = self.port.ray_counts()
        return c["primary"] + c["shadow"] + c["reflect"] + c["refract"]

    def pixels(self, lattice, rect):
        """pixel colours of the rectangle from its lattice samples: mean of k x k"""
        _, _, w, h = rect
        k = self.k
        if k == 1:
            return lattice
        return lattice.reshape(h, k, w, k, 3).astype(np.float64).mean(axis=(1, 3)).astype(np.float32)

    def close(self):
        if self.ref is not None:
            self.ref.free()


def cpu_baseline_block(cpu, rect, all_cores, runs, warmup):
    threads = cpu.cores if all_cores else 1
    full = rect[2] == cpu.W and rect[3] == cpu.H
    times, img = [], None
    for i in range(warmup + runs):
        if full and not all_cores and cpu.use_ref and cpu.wl["dat"]:
            img, secs = cpu.full_as_shipped()
        else:
            img, secs = cpu.region(rect, threads)
        if i >= warmup:
            times.append(secs)
    n_rays = cpu.rays(rect)
    secs = float(np.mean(times))
    scale = (cpu.W * cpu.H) / float(rect[2] * rect[3])
    how = ("CScene::LancerRayons() as shipped, single thread" if (full and not all_cores and cpu.use_ref and cpu.wl["dat"])
           else f"OpenMP rows over the reference's ObtenirCouleur, {threads} thread(s)")
    sample = (f"full {cpu.W}x{cpu.H} frame" if full else
              f"centred {rect[2]}x{rect[3]}-pixel crop of the {cpu.W}x{cpu.H} frame ({cpu.k * cpu.k} spp)")
    return {"value": n_rays / secs / 1e6, "unit": "Mrays/s", "cores": threads, "kind": cpu.kind,
            "ms_per_frame": secs * 1e3 * scale, "extrapolated": not full, "sample_seconds": secs, "sample_rays": n_rays,
            "sample": f"{sample}, depth {cpu.wl['depth']}, {len(times)} run(s), {how}"}, img


def parity_block(cpu, rect, want_lattice, got_f32, got8):
    """GPU frame (crop of the float frame + of the RGBA8 frame) against the reference CPU renderer's
    render of the same pixels.  North-star tolerance: RGBA8 within +-1 on >= 99.9 % of the pixels,
    the rest confined to silhouette / shadow-terminator pixels and counted."""
    x0, y0, w, h = rect
    want = cpu.pixels(want_lattice, rect)
    want8, _ = cpu.port.quantise(want)
    d8 = np.abs(got8[..., :3].astype(np.int32) - want8[..., :3].astype(np.int32)).max(axis=2)
    bad = d8 > 1
    out = {"against": f"oracle/_ref ({cpu.kind}) rendered live on this box", "crop": [int(x0), int(y0), int(w), int(h)],
           "pixels": int(w * h), "within_pm1_frac": float((d8 <= 1).mean()), "rgba8_equal_frac": float((d8 == 0).mean()),
           "mismatches": int(bad.sum()), "max_abs_diff_8bit": int(d8.max())}
    if got_f32 is not None:
        bits = (got_f32.view(np.uint32) != want.view(np.uint32)).any(axis=2)
        out["float_bit_equal_frac"] = float(1.0 - bits.mean())
        out["max_abs_diff_float"] = float(np.abs(got_f32 - want).max())
    off = 0
    if bad.any():
        # a mismatching pixel is "on a silhouette / shadow terminator" when the oracle's primary hit or any light's
        # shadow filter changes inside its 8-neighbourhood (SURVEY.md 8d "Parity gate")
        ys, xs = np.nonzero(bad)
        for py, px in zip(ys[:256], xs[:256]):
            nb = [(px + dx, py + dy) for dy in (-1, 0, 1) for dx in (-1, 0, 1)]
            o, d = primary_rays(cpu, [x0 + a for a, _ in nb], [y0 + b for _, b in nb])
            ids, t, _ = cpu.port.nearest(cpu.ps, o, d, threads=1)
            edge = len(set(ids.tolist())) > 1
            if not edge and ids[4] >= 0:
                pts = (o + d * t[:, None]).astype(np.float32)
                for l in cpu.ps["lights"]:
                    f = cpu.port.shadow(cpu.ps, pts, np.tile(l[0:3], (9, 1)).astype(np.float32), threads=1)
                    if not np.array_equal(f, np.tile(f[4], (9, 1))):
                        edge = True
                        break
            off += 0 if edge else 1
    out["mismatches_off_silhouette"] = int(off)
    out["pass"] = bool(out["within_pm1_frac"] >= 0.999 and off == 0)
    return out


def primary_rays(cpu, xs, ys):
    """Primary rays of pixel corners (xs, ys), Scene.cpp:1543-1552 in binary32 (classification of mismatches only)."""
    F = np.float32
    fr = cpu.frame
    W, H = cpu.W, cpu.H
    xs, ys = np.asarray(xs), np.asarray(ys)
    U, V, N = (np.asarray(v[:], F) for v in (fr.cam_u, fr.cam_v, fr.cam_n))
    vx = ((F(2.0) * xs.astype(F) * (F(1.0) / F(W))).astype(F) - F(1.0)).astype(F) * F(fr.half_w)
    vy = ((F(2.0) * ys.astype(F) * (F(1.0) / F(H))).astype(F) - F(1.0)).astype(F) * F(fr.half_h)
    d = vx[:, None] * U[None, :] + vy[:, None] * V[None, :] - N[None, :]
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(F)
    o = np.tile(np.asarray(fr.cam_pos[:], F), (len(xs), 1))
    return o, d


def c1_block(rt):
    """BASELINE config 1: scene1.dat at 640x480 on the reference CPU renderer, single thread, as shipped
    (Main.cpp:169-198 times exactly this call).  Also pins the oracle build on THIS host: the float checksum
    must be the one BASELINE.md section 2 records."""
    from oracle import refbind
    if not refbind.available("rec_ndebug"):
        return {"unavailable": "oracle/_ref not built on this box"}
    wl = dict(_workloads(rt)["c2"], W=640, H=480)
    ps = wl["scene"]()
    cpu = CpuRef(rt, wl, ps, 640, 480)
    times, img = [], None
    for i in range(6):
        img, secs = cpu.full_as_shipped()
        if i >= 1:
            times.append(secs)
    n_rays = cpu.rays((0, 0, 640, 480))
    cpu.close()
    s = float(img.astype(np.float64).sum())
    secs = float(np.mean(times))
    return {"workload": "C1: scene1.dat 640x480 depth 2, reference C++ CPU renderer, single thread (CScene::LancerRayons as shipped)",
            "ms_per_frame": secs * 1e3, "value": n_rays / secs / 1e6, "unit": "Mrays/s", "cores": 1, "kind": cpu.kind, "runs": len(times),
            "rays_per_frame": n_rays, "float_rgb_sum": s, "float_rgb_sum_baseline_md": 238466.994696,
            "checksum_matches": bool(abs(s - 238466.994696) < 5e-6 * 238466.994696)}


def run_reference(args):
    """The reference arm: the reference's own CPU renderer, all host cores, on the headline config; every step renders a
    bounded sample (a centred crop) of that workload."""
    rank, world, _ = dist_env()
    if rank != 0:
        return 0
    rt = load_package()
    key, _, rule = pick_workload(args.workload, max(world, args.gpus))
    if key == "c1":
        key = "c2"
    wl = _workloads(rt)[key]
    ps = wl["scene"]()
    n = max(1, args.gpus)
    W, H = wl["W"], wl["H"] * (n if wl["scaling"] == "weak" else 1)
    cpu = CpuRef(rt, wl, ps, W, H)
    rect = crop_rect(wl, W, H, wl["ref_crop"])
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    times = []
    for i in range(warmup + steps):
        _, secs = cpu.region(rect, cpu.cores)
        if i >= warmup:
            times.append(secs)
    n_rays = cpu.rays(rect)
    cpu.close()
    secs = float(np.mean(times))
    full = rect[2] == W and rect[3] == H
    base = {"value": n_rays / secs / 1e6, "unit": "Mrays/s", "cores": cpu.cores, "kind": cpu.kind,
            "ms_per_frame": secs * 1e3 * (W * H) / float(rect[2] * rect[3]), "extrapolated": not full,
            "sample_seconds": secs, "sample_rays": n_rays,
            "sample": (f"full {W}x{H} frame" if full else f"centred {rect[2]}x{rect[3]}-pixel crop of the {W}x{H} frame ({wl['spp'] ** 2} spp)")
                      + f", depth {wl['depth']}, {steps} step(s), OpenMP rows over the reference's ObtenirCouleur, {cpu.cores} thread(s)"}
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": base["value"], "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": secs * 1e3,
        "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f32", "data": wl["data"],
        "config": {"workload": wl["name"].format(W=W, H=H), "workload_rule": rule},
        "ms_per_frame": base["ms_per_frame"],
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
class Job:
    """What the workloads of one bench.py process share: the context, the ranks, the clock sampler."""

    def __init__(self, args):
        self.args = args
        self.rank, self.world, self.local_rank = dist_env()
        self.dist = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(self.local_rank)
            dist.init_process_group(backend="cpu:gloo,cuda:nccl")
            self.dist = dist
        self.rt = load_package()
        self.r = self.rt.Renderer(self.local_rank)
        if self.world > 1:
            uid = [self.r.comm_unique_id() if self.rank == 0 else None]
            self.dist.broadcast_object_list(uid, src=0)
            self.r.comm_init(self.rank, self.world, uid[0])
        self.windows = []
        self.sampler = ClockSampler(self.local_rank)
        self.sampler.start()
        self.peak = self.r.fp32_peak_detail()

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def reduce(self, values, op="max"):
        if self.dist is None:
            return [float(v) for v in values]
        import torch
        t = torch.tensor(list(values), dtype=torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return [float(v) for v in t]

    def shared_frame(self, H, W):
        """The caller's RGBA8 frame of the end-to-end leg.  One rank: pinned host memory.  Several ranks (one process per
        GPU): ONE POSIX shared-memory segment mapped by every rank, so that each GPU can write its own bands into the
        caller's frame over its own PCIe link (rtb200_render with RTB200_F_HOST_BANDS)."""
        rt = self.rt
        if self.world == 1:
            return rt.pinned_empty((H, W, 4), np.uint8), None
        from multiprocessing import shared_memory
        name = [None]
        shm = None
        if self.rank == 0:
            shm = shared_memory.SharedMemory(create=True, size=H * W * 4)
            name[0] = shm.name
        self.dist.broadcast_object_list(name, src=0)
        if self.rank != 0:
            shm = shared_memory.SharedMemory(name=name[0])
        arr = np.ndarray((H, W, 4), np.uint8, buffer=shm.buf)
        self.r.host_register(arr)
        self.barrier()
        return arr, shm

    def close(self):
        self.r.close()
        if self.dist is not None:
            self.dist.destroy_process_group()


def measure(job, key, steps, warmup, headline, want_cpu, want_parity):
    """One workload on the job's GPUs: device-resident leg, end-to-end leg, roofline, CPU baseline + parity (rank 0)."""
    rt, r, args = job.rt, job.r, job.args
    rank, world = job.rank, job.world
    wl = _workloads(rt)[key]
    ps = wl["scene"]()
    W, H = wl["W"], wl["H"] * (world if wl["scaling"] == "weak" else 1)
    r.set_scene(ps)
    bvh_info = r.build_bvh() if wl["bvh"] else None
    fr = rt.make_frame(ps.frame, W, H, max_bounces=wl["depth"], spp_grid=wl["spp"], flags=wl["flags"])
    out8, shm = job.shared_frame(H, W)

    # ---- device-resident leg: scene already in HBM, frame stays on the device ----------------
    st = None
    for _ in range(warmup):
        _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
    est_ms = max(st.ms_render + st.ms_gather, 1e-3)
    # a step shorter than MIN_REGION_MS / K is made of several frames, so that the timed region is never a few ms
    fps = max(1, int(math.ceil(MIN_REGION_MS / steps / est_ms))) if est_ms * steps < MIN_REGION_MS else 1
    if world > 1:
        fps = int(job.reduce([fps])[0])
    n_frames = steps * fps
    job.barrier()
    # per-kernel device times (roofline) from a short loop of synchronous calls ...
    prim_ms, stage, launches_sync = [], np.zeros(3), 0
    n_sync = min(n_frames, max(3, min(steps, 8)))
    r.timer_mark(2)
    for _ in range(n_sync):
        _, _, st = r.render(fr, want_f32=False, want_rgba8=False)
        prim_ms.append(st.ms_primary_kernel)
        stage += (st.ms_trace_nearest, st.ms_trace_shadow, st.ms_shade)
        launches_sync += st.kernel_launches
    r.timer_mark(3)
    sync_loop_ms = r.timer_elapsed(2, 3) / n_sync
    # ... and the K timed steps as ONE submission (rtb200_render_sequence: frames enqueued back to back
    # where the path allows it, else the same frame-by-frame loop)
    r.render_sequence([fr] * min(n_frames, 3), want_stats=False)
    # the timed region: exactly K steps.  A region shorter than a second is timed three times (host jitter moved the 0.3 s
    # region of config 2 by 25 % between two boxes) and the MEDIAN region is the one reported; all are listed.
    n_regions = 3 if est_ms * n_frames < 1000.0 else 1
    if world > 1:
        n_regions = int(job.reduce([n_regions])[0])
    regions = []
    for _ in range(n_regions):
        job.barrier()
        t_wall0 = time.time()
        r.timer_mark(0)
        done, seq_launches = 0, 0
        while done < n_frames:
            n = min(512, n_frames - done)
            _, seq_stats = r.render_sequence([fr] * n)
            assert all(s.rays_primary == st.rays_primary and s.rays_shadow == st.rays_shadow for s in seq_stats)
            seq_launches += sum(int(s.kernel_launches) for s in seq_stats)
            done += n
        r.timer_mark(1)
        regions.append(r.timer_elapsed(0, 1))
        job.barrier()
        job.windows.append((t_wall0, time.time()))
    if world > 1:                                  # the same region on every rank: the one whose max over ranks is the median
        allr = [job.reduce([x])[0] for x in regions]
        region_ms = regions[int(np.argsort(allr)[len(allr) // 2])]
        regions = allr
    else:
        region_ms = float(np.median(regions))
    step_ms = region_ms / steps
    launches = seq_launches if seq_launches > 0 else int(round(launches_sync / n_sync * n_frames))
    rays_local = float(st.rays)

    # streamed path: the tiled kernels cull whole triangles per warp; the FP32 roofline of the north-star is about the
    # brute-force scan itself, so that kernel is timed too, with the culling off (extra frames outside the timed region)
    brute = None
    if st.path == 1:
        prev = os.environ.get("RTB200_NO_BUNDLE_CULL")
        os.environ["RTB200_NO_BUNDLE_CULL"] = "1"
        r.render(fr, want_f32=False, want_rgba8=False)
        ks, fs, stg = [], [], np.zeros(3)
        for _ in range(3):
            _, _, sb = r.render(fr, want_f32=False, want_rgba8=False)
            ks.append(sb.ms_primary_kernel)
            fs.append(sb.ms_render + sb.ms_gather)
            stg += (sb.ms_trace_nearest, sb.ms_trace_shadow, sb.ms_shade)
        if prev is None:
            os.environ.pop("RTB200_NO_BUNDLE_CULL")
        else:
            os.environ["RTB200_NO_BUNDLE_CULL"] = prev
        brute = {"kernel_ms": float(np.mean(ks)), "frame_ms": float(np.mean(fs)), "frames": 3,
                 "stage_ms": {"trace_nearest": stg[0] / 3, "trace_shadow": stg[1] / 3, "shade": stg[2] / 3}}
        job.barrier()

    # ---- end-to-end leg: H2D scene upload (+ LBVH build) + render + D2H of the RGBA8 frame, every step ----
    e2e_frames = steps if est_ms * steps >= 50.0 else min(n_frames, 200)
    host_flags = rt.F_HOST_BANDS if world > 1 else 0
    fr_e2e = rt.make_frame(ps.frame, W, H, max_bounces=wl["depth"], spp_grid=wl["spp"], flags=wl["flags"] | host_flags)
    for _ in range(2):
        r.set_scene(ps)
        if wl["bvh"]:
            r.build_bvh()
        r.render(fr_e2e, want_f32=False, want_rgba8=True, out_rgba8=out8)
    job.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_frames):
        r.set_scene(ps)
        if wl["bvh"]:
            r.build_bvh()
        r.render(fr_e2e, want_f32=False, want_rgba8=(rank == 0 or world > 1), out_rgba8=out8)
    job.barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_frames
    frame8 = np.array(out8) if rank == 0 else None     # the assembled frame as the caller sees it

    rays_total = job.reduce([rays_local], "sum")[0]
    ms_max, e2e_max, dev_max = job.reduce([step_ms / fps, e2e_ms, region_ms / n_frames])

    # ---- several GPUs: the same frame on ONE GPU of this box (rank 0 alone): equal bytes, and the strong-scaling base ----
    single = None
    if world > 1 and headline:
        if rank == 0:
            r1 = rt.Renderer(job.local_rank)
            r1.set_scene(ps)
            if wl["bvh"]:
                r1.build_bvh()
            one8 = rt.pinned_empty((H, W, 4), np.uint8)
            r1.render(fr, want_f32=False, want_rgba8=True, out_rgba8=one8)
            ms1 = []
            for _ in range(3):
                _, _, s1 = r1.render(fr, want_f32=False, want_rgba8=False)
                ms1.append(s1.ms_render)
            single = {"ms_per_frame": float(np.mean(ms1)), "value": float(s1.rays) / (float(np.mean(ms1)) * 1e-3) / 1e6, "unit": "Mrays/s",
                      "frames": 3, "assembled_frame_equals_single_gpu_frame": bool(np.array_equal(one8, frame8)),
                      "differing_bytes": int((one8 != frame8).sum()),
                      "note": "rank 0 renders the whole frame alone (second context, no communicator) after the timed regions"}
            r1.close()
        job.barrier()

    block = None
    if rank == 0:
        info = r.device_info()
        T, Q, P, L = ps.counts
        traffic, traffic_src = ncu_traffic(key)
        peak_tf = job.peak["tflops"]
        if wl["bound"] == "fp32":
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
                    "frac": achieved / peak_tf if peak_tf else None, "traffic": traffic, "traffic_source": traffic_src,
                    "kernel_ms": k_ms, "algorithmic_flop_per_launch": flops,
                    "whole_frame": {"flop": frame_flops, "TFLOP/s": frame_flops / (frame_ms_roof * 1e-3) / 1e12,
                                    "frac": frame_flops / (frame_ms_roof * 1e-3) / 1e12 / peak_tf if peak_tf else None},
                    "peak_source": "measured here: register-resident FFMA chains on every SM (rtb200_fp32_peak_detail); "
                                   "TFLOP/s = 2 x ffma_per_thread x threads_per_block x blocks / elapsed_ms; MEASURED_PEAKS.json carries no FP32 figure",
                    "peak_parts": job.peak}
            if brute is not None:
                executed = st.tests_triangle_executed / max(st.tests_triangle, 1)
                roof["note"] = ("achieved / frac / whole_frame are the PLAIN brute-force scan (every ray against every surface, "
                                "culling off: extra frames outside the timed region); `value` and ms_per_step are the default "
                                "path, whose warps skip the triangles outside the cones of their rays")
                roof["plain_scan"] = brute
                roof["default_path_with_culling"] = {"level0_kernel_ms": float(np.mean(prim_ms)), "frame_ms": dev_max,
                                                     "triangle_tests_executed_frac": executed,
                                                     "algorithmic_TFLOP/s": frame_flops / (dev_max * 1e-3) / 1e12}
        else:
            # LBVH traversal: a per-thread graph walk.  What binds it is instruction issue with partly idle warps (ncu);
            # the SURVEY's byte model (nodes x 64 B + leaf triangles x 48 B against HBM) is kept as information only --
            # those bytes are served by L1 / L2, so it is NOT a roofline fraction.
            peak, src = hbm_peak()
            k_ms = stage[1] / n_sync
            byts = st.shadow_bvh_nodes * NODE_BYTES + st.shadow_tests_triangle * TRI_BYTES
            all_b = st.bvh_nodes_visited * NODE_BYTES + st.tests_triangle * TRI_BYTES
            ctr = ncu_counters(key)
            frac = None
            if ctr and ctr.get("issue_active_pct") is not None:
                frac = ctr["issue_active_pct"] / 100.0 * ctr.get("active_threads_per_inst", 32.0) / 32.0
            roof = {"bound": "issue", "kernel": "k_bvh_shadow (all launches of a frame)", "kernel_ms": k_ms,
                    "achieved": frac, "peak": 1.0, "unit": "useful issue slots (issue active x active threads / 32)", "frac": frac,
                    "ncu": ctr, "traffic": traffic, "traffic_source": traffic_src,
                    "dram_frac_of_hbm_peak": (traffic / (k_ms * 1e-3) / 1e9 / peak) if (traffic and k_ms > 0) else None,
                    "requested_bytes_model": {"bytes_per_frame_shadow": byts, "GB/s": byts / (k_ms * 1e-3) / 1e9 if k_ms > 0 else None,
                                              "hbm_peak_GB/s": peak, "peak_source": src,
                                              "note": "bytes REQUESTED by the walk (nodes x 64 B + leaf triangles x 48 B): served by L1 / L2, not a roofline"},
                    "whole_frame": {"bytes_requested": all_b, "nodes_per_ray": st.bvh_nodes_visited / max(st.rays, 1),
                                    "tri_tests_per_ray": st.tests_triangle / max(st.rays, 1)}}
        h2d = int(ps.nbytes())
        d2h = int(out8.nbytes)
        block = {
            "workload": wl["name"].format(W=W, H=H), "value": rays_total / (ms_max * 1e-3) / 1e6, "unit": "Mrays/s",
            "ms_per_frame": ms_max, "frames_per_step": fps, "timed_frames": n_frames, "timed_region_ms": region_ms,
            "timed_regions_ms": [float(x) for x in regions],
            "device_ms_per_frame": dev_max, "ms_per_frame_synchronous_calls": sync_loop_ms,
            "stage_ms_per_frame": {"trace_nearest": stage[0] / n_sync, "trace_shadow": stage[1] / n_sync, "shade": stage[2] / n_sync},
            "rays_per_frame": rays_total,
            "rays": {"primary": st.rays_primary, "shadow": st.rays_shadow, "reflect": st.rays_reflect, "refract": st.rays_refract, "rank": 0},
            "scene": {"triangles": T, "quadrics": Q, "planes": P, "lights": L, "spp": wl["spp"] ** 2, "depth": wl["depth"]},
            "e2e": {"value": rays_total / (e2e_max * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_max * fps, "ms_per_frame": e2e_max,
                    "h2d_bytes_per_step": h2d * fps, "d2h_bytes_per_step": d2h * fps, "frames": e2e_frames,
                    "includes": "rtb200_set_scene" + (" + rtb200_build_bvh" if wl["bvh"] else "") + " + rtb200_render + RGBA8 read-back, every frame"
                                + ("; every rank copies its own bands into the caller's shared host frame over its own PCIe link" if world > 1 else "")},
            "gpu_launches": int(launches),
            "roofline": roof,
            "data": wl["data"],
        }
        if bvh_info:
            block["bvh"] = bvh_info
        if key == "c3" and world == 1:
            # information: the SAME frame through the library's LBVH path (RTB200_F_USE_BVH instead of the tiled / culled scan
            # this config names) -- what a caller who only wants the frame should use; bytes compared with the headline's
            r.set_scene(ps)
            r.build_bvh()
            fb = rt.make_frame(ps.frame, W, H, max_bounces=wl["depth"], spp_grid=wl["spp"], flags=rt.F_USE_BVH)
            alt8 = rt.pinned_empty((H, W, 4), np.uint8)
            r.render(fb, want_f32=False, want_rgba8=True, out_rgba8=alt8)
            ms = []
            for _ in range(5):
                _, _, sb = r.render(fb, want_f32=False, want_rgba8=False)
                ms.append(sb.ms_render)
            block["same_frame_through_lbvh_path"] = {"ms_per_frame": float(np.mean(ms)), "value": rays_total / (float(np.mean(ms)) * 1e-3) / 1e6,
                                                     "unit": "Mrays/s", "frames": 5, "rgba8_equal_to_headline_frame": bool(np.array_equal(alt8, frame8))}
            del alt8
            r.set_scene(ps)
        if single:
            block["single_gpu_same_box"] = single
            block["speedup_vs_single_gpu_same_box"] = single["ms_per_frame"] / dev_max
        # ---- CPU baseline + parity: the reference's own renderer, live, on this box's cores ------------------
        if want_cpu or want_parity:
            cpu = CpuRef(rt, wl, ps, W, H)
            rect = crop_rect(wl, W, H, wl["crop"])
            full = wl["crop"] is None
            if full:
                block["cpu_baseline"], _ = cpu_baseline_block(cpu, rect, all_cores=False, runs=4, warmup=1)
                block["cpu_baseline_all_cores"], img = cpu_baseline_block(cpu, rect, all_cores=True, runs=4, warmup=1)
            else:
                block["cpu_baseline"], img = cpu_baseline_block(cpu, rect, all_cores=True, runs=1, warmup=0)
            if want_parity:
                x0, y0, w, h = rect
                got8 = frame8[y0:y0 + h, x0:x0 + w]
                got_f32 = None
                if world == 1:
                    if H * W <= (4096 * 2304):
                        f32, _, _ = r.render(fr, want_f32=True, want_rgba8=False)
                    else:
                        r.render(fr, want_f32=False, want_rgba8=False)
                        f32, _ = r.read_frame(H, W)
                    got_f32 = np.ascontiguousarray(f32[y0:y0 + h, x0:x0 + w])
                    del f32
                block["parity"] = parity_block(cpu, rect, img, got_f32, got8)
            cpu.close()
    if shm is not None:
        job.barrier()
        r.host_unregister(out8)
        del out8
        shm.close()
        if rank == 0:
            shm.unlink()
    return block, wl


def run_gpu(args):
    job = Job(args)
    rank, world = job.rank, job.world
    key, nested, rule = pick_workload(args.workload, world)
    if key == "c1":
        if rank == 0:
            print(json.dumps(c1_block(job.rt)))
        job.close()
        return 0
    steps = args.steps if args.steps > 0 else 20
    warmup = max(3, args.warmup)
    want_cpu = not args.no_cpu
    head, wl = measure(job, key, steps, warmup, headline=True, want_cpu=want_cpu, want_parity=want_cpu)
    configs = {}
    for k in nested:
        if k == "c1":
            if rank == 0 and want_cpu:
                configs["c1"] = c1_block(job.rt)
            continue
        if args.no_nested:
            continue
        b, _ = measure(job, k, steps if k != "c5" else min(steps, 5), 3, headline=False, want_cpu=want_cpu, want_parity=want_cpu)
        if rank == 0:
            configs[k] = b
    clocks = job.sampler.stop(job.windows)
    if rank != 0:
        job.close()
        return 0
    fps = head["frames_per_step"]
    line = {
        "metric": "Mrays/s", "value": head["value"], "unit": "Mrays/s",
        "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": head["ms_per_frame"] * fps,
        "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f32", "data": head["data"],
        "config": dict({"workload": head["workload"], "workload_rule": rule, "frames_per_step": fps}, **head["scene"],
                       l2_policy=("frame buffers rewritten every step; the scene is 0.6 KB and lives in shared memory, so there is no "
                                  "L2-resident input to flush") if key == "c2" else
                                 ("per-frame ray / hit / shadow-query buffers (hundreds of MB, larger than the 126 MB L2) are rewritten "
                                  "every step; the scene itself is meant to stay cache resident, as in production"),
                       parallelism=(f"interleaved 16-row image bands x{world}, scene replicated; device-resident frames are assembled in rank 0's "
                                    "HBM over NVLink by the producing kernels (peer frame; NCCL gather as fallback), host frames are written band "
                                    "by band by every rank over its own PCIe link") if world > 1 else "one GPU (image bands x1)",
                       arithmetic="reference rounding sequence (bit-exact mode)",
                       switches={k: v for k, v in os.environ.items() if k.startswith("RTB200_")}),
        "gpu_launches": head["gpu_launches"],
        "clocks": clocks,
        "device": job.r.device_info(),
    }
    for k, v in head.items():
        if k not in ("workload", "value", "unit", "data", "scene", "gpu_launches"):
            line[k] = v
    if configs:
        line["configs"] = configs
    print(json.dumps(line))
    job.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed steps (0 = 20)")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="rtb200", choices=["rtb200", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline + parity legs (and config 1)")
    ap.add_argument("--no-nested", action="store_true", help="headline workload only (no `configs` blocks)")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps <= 0:
            args.steps = 3
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
