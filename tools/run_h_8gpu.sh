#!/bin/bash
# 8-GPU evidence run (round 2): NCCL tests, the drop-in executable at the headline scale, scaling bench.
mkdir -p gpurun_out/r2
nvidia-smi -L > gpurun_out/r2/h_gpus.txt; nproc >> gpurun_out/r2/h_gpus.txt; free -g >> gpurun_out/r2/h_gpus.txt
python -m pytest tests/test_cli.py tests/test_gpu_parity.py -q -m gpu -k "nccl" > gpurun_out/r2/h_pytest_nccl.log 2>&1
tail -3 gpurun_out/r2/h_pytest_nccl.log
NCCL_DEBUG=INFO NCCL_DEBUG_FILE=/tmp/nccl.%p.log python tools/drop_in_at_scale.py --workload kron26 --ranks 1,2,4,8 --repeat 1 \
  --out gpurun_out/r2/h_dropin_kron26.json > gpurun_out/r2/h_dropin_kron26.txt 2>&1
cat /tmp/nccl.*.log 2>/dev/null | grep -E "Init COMPLETE|NVLS|via P2P|Connected all" | head -80 > gpurun_out/r2/h_nccl_info.txt
tail -30 gpurun_out/r2/h_dropin_kron26.txt | cut -c1-300
for cfg in "8 x" "8 --no-replan" "4 x"; do
  set -- $cfg; extra=""; [ "$2" != "x" ] && extra="$2"
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 2951$1 bench.py --gpus $1 --steps 5 --warmup 3 $extra \
    > gpurun_out/r2/h_bench_n$1$extra.json 2> gpurun_out/r2/h_bench_n$1$extra.err
  cut -c1-200 gpurun_out/r2/h_bench_n$1$extra.json
done
