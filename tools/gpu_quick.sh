#!/bin/bash
# Quick GPU-box check during development: selected tests + selected bench lines.  Usage: bash tools/gpu_quick.sh <tag> "<pytest -k expr>" <workloads...>
tag=$1; kexpr=$2; shift 2
out=gpurun_out/$tag
mkdir -p $out
if [ -n "$kexpr" ]; then
  timeout 1500 python -m pytest tests -m gpu -q --maxfail=30 -p no:cacheprovider -k "$kexpr" > $out/pytest.log 2>&1; echo "pytest rc=$?" >> $out/pytest.log
  tail -15 $out/pytest.log
fi
for w in "$@"; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline --no-subrecords > $out/bench_$w.json 2> $out/bench_$w.err
  python - "$out/bench_$w.json" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[1], "us/step", round(d["ms_per_step"] * 1e3, 2), "min", round(d["ms_per_step_min"] * 1e3, 2), "kernels", d["kernels_us"], "clocks", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
