#!/bin/bash
mkdir -p gpurun_out/r2
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2/u_pytest.log 2>&1; tail -4 gpurun_out/r2/u_pytest.log
timeout 300 python bench.py --workload lattice4096 --steps 3 --warmup 3 > gpurun_out/r2/u_bench_lattice.json 2> gpurun_out/r2/u_bench_lattice.err; tail -2 gpurun_out/r2/u_bench_lattice.err; cut -c1-300 gpurun_out/r2/u_bench_lattice.json
