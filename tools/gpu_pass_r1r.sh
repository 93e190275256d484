#!/bin/bash
# r1r: multigrid with the interior fast path -- parity, time-to-solution incl. the Krylov methods and 1024^3
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_multigrid.py -q > gpurun_out/pytest_mg_r1r.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_mg_r1r.log; tail -4 gpurun_out/pytest_mg_r1r.log)
(timeout 600 python tools/mg_compare.py --shapes 129,129,1 256,256,256 512,512,512 --methods multigrid,cg,bicgstab --cap 5000 > gpurun_out/mg_compare_r1r.jsonl 2> gpurun_out/mg_compare_r1r.err; cut -c1-300 gpurun_out/mg_compare_r1r.jsonl; tail -5 gpurun_out/mg_compare_r1r.err)
(timeout 400 python tools/mg_compare.py --shapes 1024,1024,1024 --methods multigrid >> gpurun_out/mg_compare_r1r.jsonl 2>> gpurun_out/mg_compare_r1r.err; tail -1 gpurun_out/mg_compare_r1r.jsonl | cut -c1-700; tail -3 gpurun_out/mg_compare_r1r.err)
