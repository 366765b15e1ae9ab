#!/bin/bash
# 1-GPU validation of the count / finalize changes: full GPU test suite, then option sweeps.
mkdir -p gpurun_out/r2
python -m pytest tests -m gpu -x -q > gpurun_out/r2/i_pytest.log 2>&1; tail -3 gpurun_out/r2/i_pytest.log
python tools/opt_sweep.py kron26 --levels "base" "count_missing=0" > gpurun_out/r2/i_sweep_kron26.txt 2>&1
for w in rmat22 er24 rmat16; do python tools/opt_sweep.py $w "base" "count_missing=0"; done > gpurun_out/r2/i_sweep_others.txt 2>&1
python tools/opt_sweep.py lattice4096 --steps 1 "base" "order_key=1" "reorder=0" > gpurun_out/r2/i_sweep_lattice.txt 2>&1
cut -c1-330 gpurun_out/r2/i_sweep_kron26.txt gpurun_out/r2/i_sweep_others.txt gpurun_out/r2/i_sweep_lattice.txt
