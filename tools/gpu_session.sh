#!/bin/bash
# One GPU-box session of the development loop: parity tests, the bench line, A/B lines, micro-benchmarks, timelines.
# Usage (from the repo root, under gpurun): bash tools/gpu_session.sh [tag]
tag=${1:-s}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > $out/smi.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -q --maxfail=30 -p no:cacheprovider > $out/pytest.log 2>&1; echo "pytest rc=$?" >> $out/pytest.log
timeout 900 python bench.py > $out/bench_c2.json 2> $out/bench_c2.err; echo "rc=$?" >> $out/bench_c2.err
timeout 300 python bench.py --safe-start --no-subrecords --no-cpu-baseline > $out/bench_c2_safe.json 2>&1
for w in c1 c5; do
  timeout 300 python bench.py --workload $w --no-cpu-baseline > $out/bench_${w}.json 2>&1
done
timeout 300 python bench.py --debug-flags 4096 --no-subrecords --no-cpu-baseline > $out/bench_c2_l2prefetch.json 2>&1
timeout 120 python tools/microbench_run.py > $out/microbench.json 2>&1
timeout 120 python tools/timeline.py c2 pdl > $out/tl_c2_pdl.txt 2>&1
timeout 120 python tools/timeline.py c1 pdl > $out/tl_c1_pdl.txt 2>&1
tail -3 $out/pytest.log; head -c 600 $out/bench_c2.json; echo; cat $out/microbench.json
