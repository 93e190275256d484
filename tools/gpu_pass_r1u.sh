#!/bin/bash
# r1u: multigrid with z-marching residual / prolongation on 3-D levels -- parity (all paths), time-to-solution
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_multigrid.py -q > gpurun_out/pytest_mg_r1u.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_mg_r1u.log; tail -12 gpurun_out/pytest_mg_r1u.log)
(timeout 600 python tools/mg_compare.py --shapes 128,128,128 256,256,256 512,512,512 --methods multigrid > gpurun_out/mg_compare_r1u.jsonl 2> gpurun_out/mg_compare_r1u.err; cut -c1-560 gpurun_out/mg_compare_r1u.jsonl; tail -5 gpurun_out/mg_compare_r1u.err)
