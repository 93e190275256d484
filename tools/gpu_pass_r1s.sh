#!/bin/bash
# r1s: multigrid with the one-pass two-colour sweep -- parity (3 paths), time-to-solution
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_multigrid.py -q -x > gpurun_out/pytest_mg_r1s.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_mg_r1s.log; tail -12 gpurun_out/pytest_mg_r1s.log)
(timeout 600 python tools/mg_compare.py --shapes 129,129,1 1025,1025,1 128,128,128 256,256,256 512,512,512 --methods multigrid > gpurun_out/mg_compare_r1s.jsonl 2> gpurun_out/mg_compare_r1s.err; cut -c1-560 gpurun_out/mg_compare_r1s.jsonl; tail -5 gpurun_out/mg_compare_r1s.err)
