#!/bin/bash
# final-tree check on one GPU: whole GPU suite, smoke(), the default bench line and the reference arm as the driver runs them
mkdir -p gpurun_out/r2
python -m pytest tests -m gpu -x -q > gpurun_out/r2/p_pytest.log 2>&1; tail -2 gpurun_out/r2/p_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2/p_smoke.log 2>&1; tail -1 gpurun_out/r2/p_smoke.log
( time python bench.py --impl reference --steps 20 --warmup 5 ) > gpurun_out/r2/p_ref.json 2> gpurun_out/r2/p_ref.err; tail -4 gpurun_out/r2/p_ref.err; cut -c1-300 gpurun_out/r2/p_ref.json
( time python bench.py --steps 20 --warmup 5 ) > gpurun_out/r2/p_bench.json 2> gpurun_out/r2/p_bench.err; tail -4 gpurun_out/r2/p_bench.err; cut -c1-300 gpurun_out/r2/p_bench.json
