#!/usr/bin/env python
""".ncu-rep captures (ncu --set full, one launch each) -> the two small JSON files bench.py reads:
  profiles/r2_ncu_traffic.json        DRAM bytes per launch of each workload's dominant kernel (roofline.traffic)
  profiles/r2_ncu_lbvh_counters.json  what binds the LBVH walk: issue active, active threads per instruction, stalls,
                                      L1 / L2 hit rates, L2 throughput (roofline of configs 4 / 5)
and one text summary per capture (profiles/<tag>_<name>_ncu_full.txt).
usage: python scripts/ncu_extract.py gpurun_out r2p [output directory, default profiles/]
(scripts/r2_profile.sh runs it ON the GPU box with gpurun_out/extracted as output: ten .ncu-rep files exceed what a call may bring back)"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import ncu_summary  # noqa: E402

WANT = {
    "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "active_threads_per_inst": "smsp__thread_inst_executed_per_inst_executed.ratio",
    "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "l1_hit_pct": "l1tex__t_sector_hit_rate.pct",
    "l2_hit_pct": "lts__t_sector_hit_rate.pct",
    "stall_long_scoreboard": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "stall_wait": "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "stall_not_selected": "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "stall_no_instruction": "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "stall_barrier": "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "fma_pipe_pct": "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "registers": "launch__registers_per_thread",
    "duration_ms": "gpu__time_duration.sum",
    "inst_executed": "smsp__inst_executed.sum",
    "dram_read": "dram__bytes_read.sum",
    "dram_write": "dram__bytes_write.sum",
    "l2_throughput_pct": "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "bank_conflicts_shared": "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1_data_pipe_pct": "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1_wavefronts": "l1tex__data_pipe_lsu_wavefronts.sum",
    "global_load_requests": "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "global_load_sectors": "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
}
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}


def read(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, r = rows[0], rows[1], rows[2]
    out = {"kernel": r[hdr.index("Kernel Name")]}
    for k, m in WANT.items():
        if m in hdr:
            i = hdr.index(m)
            v = float(r[i].replace(",", ""))
            if k in ("dram_read", "dram_write", "duration_ms"):
                v *= UNIT.get(units[i], 1.0)
            out[k] = v
    return out


def main():
    d, tag = sys.argv[1], sys.argv[2]
    out_dir = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "profiles")
    os.makedirs(out_dir, exist_ok=True)
    caps = {n[len(tag) + 1:-8]: os.path.join(d, n) for n in sorted(os.listdir(d)) if n.startswith(tag + "_") and n.endswith(".ncu-rep")}
    got = {}
    for name, rep in caps.items():
        got[name] = read(rep)
        sys.argv = ["ncu_summary.py", rep, os.path.join(out_dir, f"r2_{name}_ncu_full.txt")]
        try:
            import io, contextlib
            with contextlib.redirect_stdout(io.StringIO()):
                ncu_summary.main()
        except Exception as e:
            print("summary failed for", name, e)
    traffic = {}
    for wl, name in (("c2", "c2_resident"), ("c3", "c3_plain_nearest"), ("c4", "c4_bvh_shadow"), ("c5", "c5_bvh_shadow")):
        if name in got:
            g = got[name]
            traffic[wl] = {"dram_read_bytes": g["dram_read"], "dram_write_bytes": g["dram_write"], "kernel": g["kernel"],
                           "source": f"profiles/r2_{name}_ncu_full.txt (ncu --set full --clock-control none, one launch)"}
    json.dump(traffic, open(os.path.join(out_dir, "r2_ncu_traffic.json"), "w"), indent=1)
    lb = {}
    for wl, name in (("c4", "c4_bvh_shadow"), ("c5", "c5_bvh_shadow")):
        if name in got:
            lb[wl] = dict(got[name], source=f"profiles/r2_{name}_ncu_full.txt")
    if "c4_bvh_nearest" in got:
        lb["c4_nearest"] = dict(got["c4_bvh_nearest"], source="profiles/r2_c4_bvh_nearest_ncu_full.txt")
    json.dump(lb, open(os.path.join(out_dir, "r2_ncu_lbvh_counters.json"), "w"), indent=1)
    for k, v in got.items():
        print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items() if a not in ("kernel",)})


if __name__ == "__main__":
    main()
