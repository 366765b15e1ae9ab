#!/bin/bash
# one --set full capture of the lane-major BFS kernel (lattice 1024^2: the whole batch is one launch)
mkdir -p gpurun_out/r2
CMD="python bench.py --workload lattice1024 --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/r2/lm_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:lm_bfs_kernel -s 1 -c 1 -o gpurun_out/r2/prof_lm_bfs_lattice1024 $CMD > gpurun_out/r2/lm_ncu.log 2>&1
echo "rc=$?"; ls -la gpurun_out/r2/prof_lm_bfs_lattice1024.ncu-rep
