#!/bin/bash
# r1q: multigrid with the V-cycle as a CUDA graph -- parity both ways, time-to-solution, ncu of the fine level
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_multigrid.py -q > gpurun_out/pytest_mg_r1q.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_mg_r1q.log; tail -6 gpurun_out/pytest_mg_r1q.log)
(timeout 600 python tools/mg_compare.py --shapes 129,129,1 1025,1025,1 128,128,128 256,256,256 --methods multigrid,rbsor,jacobi > gpurun_out/mg_compare_r1q.jsonl 2> gpurun_out/mg_compare_r1q.err; cut -c1-330 gpurun_out/mg_compare_r1q.jsonl; tail -5 gpurun_out/mg_compare_r1q.err)
(timeout 300 python tools/mg_compare.py --shapes 512,512,512 --methods multigrid >> gpurun_out/mg_compare_r1q.jsonl 2>> gpurun_out/mg_compare_r1q.err; tail -1 gpurun_out/mg_compare_r1q.jsonl | cut -c1-700)
(CFD_B200_NO_MG_GRAPH=1 timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_mg --launch-skip 1 --launch-count 12 -o gpurun_out/prof_r1q python tools/mg_compare.py --shapes 512,512,512 --methods multigrid > gpurun_out/ncu_r1q.log 2>&1; tail -2 gpurun_out/ncu_r1q.log)
