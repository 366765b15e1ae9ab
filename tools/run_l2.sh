#!/bin/bash
# ncu launch list (time + DRAM bytes per launch) of the final tree, for profiles/traffic.json
mkdir -p gpurun_out/r2
CMD="python bench.py --workload kron26 --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/r2/l2_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 900 --csv \
    --log-file gpurun_out/r2/l2_launches_kron26_first900.csv $CMD > gpurun_out/r2/l2_ncu1.log 2>&1
echo "launch list rc=$?"
