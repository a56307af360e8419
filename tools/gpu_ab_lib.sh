#!/bin/bash
# A/B of two BUILDS on one box: bash tools/gpu_ab_lib.sh <tag> <path of the B library> <workloads...>; A = the in-tree library.
tag=$1; libb=$2; shift 2
out=gpurun_out/$tag
mkdir -p $out
for rep in $(seq 1 ${REPS:-3}); do
for w in "$@"; do
  for arm in a b; do
    if [ $arm = b ]; then export MLAPO_LIB=$libb; else unset MLAPO_LIB; fi
    timeout 600 python bench.py --workload $w --no-cpu-baseline --no-subrecords > $out/${w}_${arm}_$rep.json 2> $out/${w}_${arm}_$rep.err
    python - "$out/${w}_${arm}_$rep.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[1], "us/step", round(d["ms_per_step"] * 1e3, 2), "min", round(d["ms_per_step_min"] * 1e3, 2), "kernels", d["kernels_us"], "clocks", d["clocks"]["sm_mhz"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
  done
done
done
