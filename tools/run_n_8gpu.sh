#!/bin/bash
# 8-GPU, final round-2 code: NCCL tests of the drop-in, then the executable at the headline scale
# (sharded read; NCCL start-up before the clock), np = 8 also with whole-file reads and with NCCL start-up inside the window.
mkdir -p gpurun_out/r2
python -m pytest tests/test_cli.py tests/test_gpu_parity.py -q -m gpu -k "nccl" > gpurun_out/r2/n_pytest_nccl.log 2>&1; tail -2 gpurun_out/r2/n_pytest_nccl.log
python tools/drop_in_at_scale.py --workload kron26 --ranks 1,2,4,8 --repeat 1 --out gpurun_out/r2/n_dropin_kron26.json > gpurun_out/r2/n_dropin_kron26.txt 2>&1
MSBFS_NO_SHARDED_LOAD=1 python tools/drop_in_at_scale.py --workload kron26 --ranks 8 --repeat 1 --out gpurun_out/r2/n_dropin_kron26_whole.json > gpurun_out/r2/n_dropin_kron26_whole.txt 2>&1
MSBFS_NCCL_IN_WINDOW=1 python tools/drop_in_at_scale.py --workload kron26 --ranks 8 --repeat 1 --out gpurun_out/r2/n_dropin_kron26_ncclin.json > gpurun_out/r2/n_dropin_kron26_ncclin.txt 2>&1
grep -h '"np"' gpurun_out/r2/n_dropin_kron26.txt gpurun_out/r2/n_dropin_kron26_whole.txt gpurun_out/r2/n_dropin_kron26_ncclin.txt | cut -c1-330
