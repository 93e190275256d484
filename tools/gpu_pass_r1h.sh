#!/bin/bash
# Round-1h GPU pass: full GPU suite with the cluster (P)CG path, small-grid solve timings with / without
# the cluster kernels.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r1h.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r1h.log
tail -15 gpurun_out/pytest_gpu_r1h.log
{
  for sh in 33,33,1 129,129,1; do
    timeout 120 python tools/kernel_table.py --shape $sh --method jacobi --iters 2000 --steps 5
    timeout 120 python tools/kernel_table.py --shape $sh --method cg --iters 200 --steps 5
    CFD_B200_NO_CLUSTER_CG=1 timeout 120 python tools/kernel_table.py --shape $sh --method cg --iters 200 --steps 5
    timeout 120 python tools/kernel_table.py --shape $sh --method pcg --iters 200 --steps 5
  done
  timeout 120 python tools/kernel_table.py --shape 257,257,1 --method cg --iters 200 --steps 5
} > gpurun_out/kernel_tables_r1h.jsonl 2> gpurun_out/kernel_tables_r1h.err
cut -c1-330 gpurun_out/kernel_tables_r1h.jsonl
tail -3 gpurun_out/kernel_tables_r1h.err
