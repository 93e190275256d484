#!/bin/bash
# r1p: multigrid projection-step parity, ncu capture of the multigrid kernels, then the whole GPU suite
mkdir -p gpurun_out
(timeout 300 python -m pytest tests/test_gpu_multigrid.py -q -k projection > gpurun_out/pytest_mg2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_mg2.log; tail -6 gpurun_out/pytest_mg2.log)
(timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_mg --launch-skip 700 --launch-count 14 -o gpurun_out/prof_r1p python tools/mg_compare.py --shapes 256,256,256 --methods multigrid > gpurun_out/ncu_r1p.log 2>&1; tail -2 gpurun_out/ncu_r1p.log)
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_r1p.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r1p.log; tail -6 gpurun_out/pytest_gpu_r1p.log)
