#!/bin/bash
mkdir -p gpurun_out/r2
python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/r2/o_pytest.log 2>&1; tail -2 gpurun_out/r2/o_pytest.log
python tools/opt_sweep.py kron26 --levels "base" "dense_prefix=0" > gpurun_out/r2/o_sweep_kron26.txt 2>&1
for w in rmat22 er24; do python tools/opt_sweep.py $w "base" "dense_prefix=0"; done > gpurun_out/r2/o_sweep_others.txt 2>&1
cut -c1-420 gpurun_out/r2/o_sweep_kron26.txt gpurun_out/r2/o_sweep_others.txt
