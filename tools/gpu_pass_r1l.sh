#!/bin/bash
# Round-1l GPU pass: 16-CTA (non-portable) cluster for the small-grid solves -- parity, then timing.
mkdir -p gpurun_out
CFD_B200_CLUSTER_CTAS=16 timeout 200 python -m pytest tests/test_gpu_poisson.py -q -k "cluster" > gpurun_out/pytest_c16.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_c16.log
tail -6 gpurun_out/pytest_c16.log | cut -c1-300
{
  for sh in 129,129,1 200,200,1; do
    CFD_B200_CLUSTER_CTAS=16 timeout 60 python tools/kernel_table.py --shape $sh --method jacobi --iters 2000 --steps 5
    CFD_B200_CLUSTER_CTAS=16 CFD_B200_CLUSTER_ASYNC=1 timeout 60 python tools/kernel_table.py --shape $sh --method jacobi --iters 2000 --steps 5
    CFD_B200_CLUSTER_CTAS=16 timeout 60 python tools/kernel_table.py --shape $sh --method cg --iters 200 --steps 5
  done
} > gpurun_out/kernel_tables_r1l.jsonl 2> gpurun_out/kernel_tables_r1l.err
cut -c1-300 gpurun_out/kernel_tables_r1l.jsonl
tail -3 gpurun_out/kernel_tables_r1l.err
