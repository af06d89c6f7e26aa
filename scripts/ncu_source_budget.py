#!/usr/bin/env python
"""Executed warp instructions of one profiled kernel by SOURCE line of the kernel body (the outermost frame of the
inlining chain), from an `ncu --set full --import-source on` report and the cubin of the same build.
usage: python scripts/ncu_source_budget.py report.ncu-rep object.o <mangled kernel name> <units> > profiles/x.txt
(units = what to divide by, e.g. warps of work = pixels / 32)"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile


def main():
    rep, obj, fun, units = sys.argv[1], os.path.abspath(sys.argv[2]), sys.argv[3], float(sys.argv[4])
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, data = rows[1], rows[2:]
    ie = hdr.index("Instructions Executed")
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=tmp, capture_output=True)
        cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
        dis = subprocess.run(["nvdisasm", "-gi", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    seq, chain, pending, inside = [], [], [], False
    for line in dis.splitlines():
        if line.startswith("//---") and ".text." in line:
            inside = (".text." + fun + " ") in line + " " or line.rstrip("- ").endswith(".text." + fun)
            continue
        if not inside:
            continue
        m = re.search(r'//## File "(.*?)", line (\d+)', line)
        if m:
            pending.append((os.path.basename(m.group(1)), int(m.group(2))))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", line)
        if m:
            if pending:
                chain, pending = pending, []
            seq.append((m.group(2).strip(), list(chain)))
    assert len(seq) >= len(data), (len(seq), len(data))
    outer, ops, total = collections.Counter(), collections.defaultdict(collections.Counter), 0
    for (op, ch), d in zip(seq, data):
        c = int(d[ie])
        total += c
        outer[ch[-1]] += c
        ops[ch[-1]][re.sub(r"^(@!?U?P\d\s+)", "", op).split()[0].split(".")[0]] += c
    print(f"# {fun}: {total} warp instructions executed = {total / units:.1f} per unit ({units:.0f} units)")
    print("# per unit   share   kernel-body line            top opcodes (per unit)")
    for k, v in outer.most_common(28):
        top = ", ".join(f"{a} {b / units:.0f}" for a, b in ops[k].most_common(7))
        print(f"{v / units:9.1f}  {100 * v / total:5.1f} %  {k[0]}:{k[1]:<6d}  {top}")


if __name__ == "__main__":
    main()
