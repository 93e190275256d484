#!/bin/bash
# Final 1-GPU pass of the round: full GPU suite, smoke(), default bench line.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_final.log
tail -8 gpurun_out/pytest_gpu_final.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke_final.log
tail -3 gpurun_out/smoke_final.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench exit $?"
cut -c1-400 gpurun_out/bench_final.json
tail -3 gpurun_out/bench_final.err
