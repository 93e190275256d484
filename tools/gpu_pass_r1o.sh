#!/bin/bash
# r1o: geometric multigrid -- parity tests, then time-to-solution against Red-Black SOR / Jacobi
mkdir -p gpurun_out
nvidia-smi -L
(timeout 600 python -m pytest tests/test_gpu_multigrid.py -q > gpurun_out/pytest_mg.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_mg.log; tail -25 gpurun_out/pytest_mg.log)
(timeout 600 python tools/mg_compare.py --shapes 129,129,1 1025,1025,1 128,128,128 256,256,256 --methods multigrid,rbsor > gpurun_out/mg_compare.jsonl 2> gpurun_out/mg_compare.err; cut -c1-420 gpurun_out/mg_compare.jsonl; tail -5 gpurun_out/mg_compare.err)
(timeout 300 python tools/mg_compare.py --shapes 512,512,512 --methods multigrid > gpurun_out/mg_compare_512.jsonl 2>> gpurun_out/mg_compare.err; cut -c1-700 gpurun_out/mg_compare_512.jsonl; tail -3 gpurun_out/mg_compare.err)
