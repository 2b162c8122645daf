#!/bin/bash
# usage: tools/bench_short.sh <tag> [env assignments...]; prints the key fields of a 5-step bench
tag=$1; shift
env "$@" python bench.py --steps 6 $BENCH_ARGS > gpurun_out/b_$tag.log 2>&1
python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/b_$tag.log") if x.startswith("{")][-1]
    d=json.loads(l)
    print("$tag", "value %.0f ms %.2f e2e %.0f e2e+n %.0f lat %.2f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("e2e_with_normals", {}).get("value", 0), d["latency_ms_single_frame"]), {k: round(v, 2) for k, v in d["stage_ms"].items()}, d["parity"])
except Exception as e:
    print("$tag failed", e); print(open("gpurun_out/b_$tag.log").read()[-1500:])
PY
