#!/bin/bash
# usage: tools/ncu_c2.sh <tag> [lib]  -- ncu duration + instruction count of the c2 kernels for one library build
tag=$1; lib=${2:-}
[ -n "$lib" ] && export KP_LIB_PATH=$lib
python tools/snap_one.py c2 > gpurun_out/plain_$tag.log 2>&1 && ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__cycles_active.avg,sm__cycles_elapsed.max --clock-control none -k regex:"place_|plan_count" -s 2 -c 4 --csv --log-file gpurun_out/ncu_$tag.csv python tools/snap_one.py c2 > gpurun_out/ncu_$tag.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(l for l in open("gpurun_out/ncu_$tag.csv") if not l.startswith("=="))]
h=rows[0]
for r in rows[1:]:
    d=dict(zip(h,r)); print("$tag", d["Kernel Name"][:28], d["Metric Name"], d["Metric Value"])
PY
