#!/bin/bash
# per-launch durations of ONE config-3 frame (ncu launch list; cold-cache, serialised times: shares, not absolutes)
OUT=gpurun_out; TAG=${1:-r2l}; mkdir -p $OUT
cat > /tmp/one_frame.py <<'PY'
import os, sys
from importlib import import_module
sys.path.insert(0, os.getcwd())
from __graft_entry__ import load_package
rt = load_package(); synth = import_module("rtb200.synth")
ps = synth.soup(10000, 64, 4, seed=1234)
r = rt.Renderer(0); r.set_scene(ps)
fr = rt.make_frame(ps.frame, 3840, 2160, max_bounces=3, flags=rt.F_FORCE_STREAM)
for _ in range(2): r.render(fr, want_f32=False, want_rgba8=False)
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_c3_frame_launches.csv python /tmp/one_frame.py > $OUT/${TAG}_c3_frame_launches.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("$OUT/${TAG}_c3_frame_launches.csv")) if len(r)>10 and r[0].isdigit()]
half=len(rows)//2
# second frame only: find the index of the second k_cull_nearest primary... simply print the last frame's kernels
names=[(r[4].split('(')[0][:60], float(r[-1].replace(',',''))) for r in rows]
# locate start of last frame: last occurrence of a memset-less marker: first k_cull_nearest with Lb1 (primary)
idx=[i for i,(n,_) in enumerate(names) if 'k_cull_nearest' in n and 'true, false' in rows[i][4]]
start=idx[-1] if idx else 0
tot=0
for n,t in names[start:]:
    print(f"{t/1e6 if t>1e4 else t/1e6:10.4f} ms  {n}")
    tot+=t
print('total', tot/1e6)
PY
