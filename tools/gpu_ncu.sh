#!/bin/bash
# ncu evidence for the default workload (run under gpurun): launch list with per-launch device time and DRAM bytes, then one
# full capture of the two largest kernels.  The same command runs once without ncu first and must exit 0.
out=gpurun_out/ncu
mkdir -p $out
CMD="python bench.py --steps 20 --warmup 5 --no-graph --no-cpu-baseline --no-subrecords"
$CMD > $out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:"quant_hidden|mm1_|norm_rope|q_fused" -s 40 -c 16 --csv --log-file $out/launches_c2.csv $CMD > $out/ncu_launches.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"q_fused|mm1_pair" -s 16 -c 2 -o $out/full_c2 -f $CMD > $out/ncu_full.log 2>&1
echo "rc=$?" > $out/rc.txt
tail -2 $out/plain.log | cut -c1-300; tail -20 $out/launches_c2.csv | cut -c1-200; cat $out/rc.txt; ls -la $out
