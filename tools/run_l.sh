#!/bin/bash
# 1-GPU evidence for round 2: the default bench line, the ncu launch list (time + DRAM bytes per launch) of the same command,
# and one --set full capture of the dominant kernel.
mkdir -p gpurun_out/r2
python bench.py --steps 5 --warmup 3 > gpurun_out/r2/l_bench_kron26.json 2> gpurun_out/r2/l_bench_kron26.err; echo "bench rc=$?"
CMD="python bench.py --workload kron26 --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/r2/l_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 900 --csv \
    --log-file gpurun_out/r2/l_launches_kron26_first900.csv $CMD > gpurun_out/r2/l_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:bu_thread_kernel -c 3 -o gpurun_out/r2/l_prof_bu_thread $CMD > gpurun_out/r2/l_ncu2.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out/r2/ | grep " l_"
