#!/bin/bash
# Round-1j GPU pass: full GPU suite (not stopping at the first failure), register-resident cluster Jacobi timing.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_r1j.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r1j.log
tail -12 gpurun_out/pytest_gpu_r1j.log | cut -c1-300
{
  for sh in 33,33,1 65,65,1 129,129,1 200,200,1; do
    timeout 120 python tools/kernel_table.py --shape $sh --method jacobi --iters 2000 --steps 5
  done
} > gpurun_out/kernel_tables_r1j.jsonl 2> gpurun_out/kernel_tables_r1j.err
cut -c1-300 gpurun_out/kernel_tables_r1j.jsonl
tail -3 gpurun_out/kernel_tables_r1j.err
