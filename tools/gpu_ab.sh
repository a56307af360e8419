#!/bin/bash
# A/B on ONE box: bash tools/gpu_ab.sh <tag> <debug-flag word of the B arm> <workloads...>; the A arm runs with no debug flags.
tag=$1; fb=$2; shift 2
out=gpurun_out/$tag
mkdir -p $out
for rep in 1 2; do
for w in "$@"; do
  for f in 0 $fb; do
    timeout 600 python bench.py --workload $w --no-cpu-baseline --no-subrecords --debug-flags $f > $out/${w}_${f}_$rep.json 2> $out/${w}_${f}_$rep.err
    python - "$out/${w}_${f}_$rep.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[1], "us/step", round(d["ms_per_step"] * 1e3, 2), "min", round(d["ms_per_step_min"] * 1e3, 2), "kernels", d["kernels_us"], "clocks", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
  done
done
done
