#!/bin/bash
# r1t: final state check -- whole GPU suite, smoke(), default bench
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_r1t.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r1t.log; tail -5 gpurun_out/pytest_gpu_r1t.log)
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r1t.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke_r1t.log; tail -3 gpurun_out/smoke_r1t.log)
(timeout 600 python bench.py > gpurun_out/bench_r1t.json 2> gpurun_out/bench_r1t.err; echo "bench exit $?"; cut -c1-400 gpurun_out/bench_r1t.json; tail -2 gpurun_out/bench_r1t.err)
