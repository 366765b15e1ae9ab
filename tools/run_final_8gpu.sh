#!/bin/bash
# 8-GPU box, final round-2 tree: strong scaling of the headline workload (other workloads: pass them as arguments "N workload")
mkdir -p gpurun_out/r2
run() { n=$1; w=$2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 5 --warmup 3 --no-cpu --workload $w \
  > gpurun_out/r2/final_bench_${w}_n$n.json 2> gpurun_out/r2/final_bench_${w}_n$n.err
  python -c "
import json
d=json.loads(open('gpurun_out/r2/final_bench_${w}_n$n.json').read().strip().splitlines()[-1]); print('$w', $n, round(d['ms_per_step'],2), round(d['queries_per_sec']), d['config']['lanes'], d['per_rank_device_ms'], d['plan'][:200])"; }
if [ $# -ge 2 ]; then while [ $# -ge 2 ]; do run $1 $2; shift 2; done; else run 8 kron26; run 4 kron26; run 2 kron26; fi
