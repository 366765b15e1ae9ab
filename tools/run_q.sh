#!/bin/bash
mkdir -p gpurun_out/r2
python tools/opt_sweep.py kron26 --levels "base" "alpha=24" "alpha=16" "alpha=12" "alpha=9" "alpha=6" > gpurun_out/r2/q_alpha_kron26.txt 2>&1
for w in rmat22 er24; do python tools/opt_sweep.py $w "base" "alpha=24" "alpha=16" "alpha=12" "alpha=9" "alpha=6"; done > gpurun_out/r2/q_alpha_others.txt 2>&1
python tools/opt_sweep.py lattice4096 --steps 1 "base" "alpha=16" "alpha=9" > gpurun_out/r2/q_alpha_lattice.txt 2>&1
grep -h "ms/step" gpurun_out/r2/q_alpha_kron26.txt gpurun_out/r2/q_alpha_others.txt gpurun_out/r2/q_alpha_lattice.txt | cut -c1-200
grep -h "batch 7" gpurun_out/r2/q_alpha_kron26.txt | cut -c1-420
