#!/bin/bash
# 8-GPU: the drop-in executable at the headline scale with the sharded read (and np = 8 with whole-file reads), final scaling bench.
mkdir -p gpurun_out/r2
python tools/drop_in_at_scale.py --workload kron26 --ranks 1,2,4,8 --repeat 1 --out gpurun_out/r2/k_dropin_kron26.json > gpurun_out/r2/k_dropin_kron26.txt 2>&1
MSBFS_NO_SHARDED_LOAD=1 python tools/drop_in_at_scale.py --workload kron26 --ranks 8 --repeat 1 --out gpurun_out/r2/k_dropin_kron26_whole.json > gpurun_out/r2/k_dropin_kron26_whole.txt 2>&1
grep -h '"np"' gpurun_out/r2/k_dropin_kron26.txt gpurun_out/r2/k_dropin_kron26_whole.txt | cut -c1-330
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 5 --warmup 3 \
    > gpurun_out/r2/k_bench_n$n.json 2> gpurun_out/r2/k_bench_n$n.err
  cut -c1-160 gpurun_out/r2/k_bench_n$n.json
done
