#!/usr/bin/env python
"""One bench.py JSON line -> a short table (headline + every block under `configs`).  usage: python scripts/bench_summary.py line.json"""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])


def brief(b, name):
    print(f"== {name}: {b.get('workload', d['config']['workload'])[:90]}")
    e = b.get("e2e", {})
    print(f"   ms/frame {b.get('ms_per_frame'):.4f}  value {b.get('value'):.1f} Mrays/s  e2e {e.get('value', 0):.1f} Mrays/s ({e.get('ms_per_frame', 0):.4f} ms/frame)  launches {b.get('gpu_launches')}")
    if "stage_ms_per_frame" in b:
        print("   stages", {k: round(v, 3) for k, v in b["stage_ms_per_frame"].items()})
    r = b.get("roofline")
    if r:
        print(f"   roofline {r['bound']}: frac {r['frac']}  kernel_ms {r.get('kernel_ms')}  traffic {r.get('traffic')}")
        if "default_path_with_culling" in r:
            print("   default path", r["default_path_with_culling"])
    for k in ("cpu_baseline", "cpu_baseline_all_cores"):
        if k in b:
            c = b[k]
            print(f"   {k}: {c['value']:.4f} {c['unit']} on {c['cores']} core(s), {c['sample_seconds']:.2f} s, {c['ms_per_frame']:.1f} ms/frame{' (extrapolated)' if c['extrapolated'] else ''}")
    if "parity" in b:
        print("   parity", {k: v for k, v in b["parity"].items() if k not in ("against",)})
    for k in ("same_frame_through_lbvh_path", "single_gpu_same_box", "speedup_vs_single_gpu_same_box"):
        if k in b:
            print(f"   {k}:", b[k])


print({k: d[k] for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "scaling", "gpu_launches", "clocks")})
print("workload_rule:", d["config"]["workload_rule"])
brief(d, "headline")
for k, v in d.get("configs", {}).items():
    if k == "c1":
        print("== c1", v)
    else:
        brief(v, k)
