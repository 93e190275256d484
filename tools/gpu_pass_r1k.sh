#!/bin/bash
# Round-1k GPU pass: st.async neighbour-signalling variant of the cluster Jacobi (opt-in) -- parity, then timing.
mkdir -p gpurun_out
CFD_B200_CLUSTER_ASYNC=1 timeout 150 python -m pytest tests/test_gpu_poisson.py -q -k "jacobi" > gpurun_out/pytest_async.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_async.log
tail -6 gpurun_out/pytest_async.log | cut -c1-300
{
  for sh in 65,65,1 129,129,1 200,200,1; do
    CFD_B200_CLUSTER_ASYNC=1 timeout 60 python tools/kernel_table.py --shape $sh --method jacobi --iters 2000 --steps 5
  done
} > gpurun_out/kernel_tables_r1k.jsonl 2> gpurun_out/kernel_tables_r1k.err
cut -c1-300 gpurun_out/kernel_tables_r1k.jsonl
tail -3 gpurun_out/kernel_tables_r1k.err
nvidia-smi --query-gpu=name,memory.used --format=csv,noheader
