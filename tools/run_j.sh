#!/bin/bash
# 2-GPU check of the drop-in's NCCL path: tests, then the executable on RMAT-22 with and without the sharded read.
mkdir -p gpurun_out/r2
python -m pytest tests/test_cli.py tests/test_gpu_parity.py -q -m gpu -k "nccl" > gpurun_out/r2/j_pytest_nccl.log 2>&1; tail -3 gpurun_out/r2/j_pytest_nccl.log
python tools/drop_in_at_scale.py --workload rmat22 --ranks 1,2 --repeat 2 --out gpurun_out/r2/j_dropin_rmat22.json > gpurun_out/r2/j_dropin_rmat22.txt 2>&1
MSBFS_NO_SHARDED_LOAD=1 python tools/drop_in_at_scale.py --workload rmat22 --ranks 2 --repeat 2 --out gpurun_out/r2/j_dropin_rmat22_whole.json > gpurun_out/r2/j_dropin_rmat22_whole.txt 2>&1
cut -c1-330 gpurun_out/r2/j_dropin_rmat22.txt gpurun_out/r2/j_dropin_rmat22_whole.txt
