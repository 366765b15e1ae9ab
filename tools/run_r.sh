#!/bin/bash
mkdir -p gpurun_out/r2
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_cli.py -x -q -m gpu > gpurun_out/r2/r_pytest.log 2>&1; tail -3 gpurun_out/r2/r_pytest.log
timeout 300 python -m pytest tests/test_zy_fullsize_parity.py -x -q -m gpu -k lattice > gpurun_out/r2/r_pytest_lattice.log 2>&1; tail -3 gpurun_out/r2/r_pytest_lattice.log
timeout 300 python tools/opt_sweep.py lattice4096 --steps 1 "base" "lane_major=0" > gpurun_out/r2/r_sweep_lattice.txt 2>&1; cut -c1-400 gpurun_out/r2/r_sweep_lattice.txt
timeout 300 python tools/opt_sweep.py lattice1024 --steps 2 "base" "lane_major=0" > gpurun_out/r2/r_sweep_lattice1024.txt 2>&1; cut -c1-300 gpurun_out/r2/r_sweep_lattice1024.txt
