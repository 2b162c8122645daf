#!/bin/bash
# usage: tools/ncu_stalls.sh <frames> <tag> : stall-reason + i-cache metrics of the growth kernels
F=$1; tag=$2
M=sm__icc_request_hit_rate.pct,sm__icc_requests.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active
for r in no_instruction long_scoreboard short_scoreboard barrier wait branch_resolving sleeping membar lg_throttle mio_throttle not_selected; do
  M=$M,smsp__average_warps_issue_stalled_${r}_per_issue_active.ratio
done
ncu --metrics $M --clock-control none -k regex:grow_kernel -c 2 --csv --log-file gpurun_out/stalls_$tag.csv python bench.py --frames $F --steps 1 --warmup 3 --cpu-sample 1 > gpurun_out/stalls_$tag.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/stalls_$tag.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); mi=hdr.index("Metric Name"); vi=hdr.index("Metric Value"); ii=hdr.index("ID")
out={}
for r in rows[1:]:
    out.setdefault((r[ii],r[ki][:40]),{})[r[mi].replace("smsp__average_warps_issue_stalled_","").replace("_per_issue_active.ratio","")]=r[vi]
for k,v in out.items():
    print("$tag",k); print("    ",v)
PY
