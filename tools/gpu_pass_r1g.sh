#!/bin/bash
# Round-1g GPU pass: full GPU suite after the group-clamp / density-reciprocal change, RK / Euler kernel
# tables (before: profiles/r1_kernel_tables.jsonl, r1f_kernel_tables.jsonl), tiny-grid cluster Jacobi,
# stand-alone Poisson objects against the reference CUDA backend with a linear fit over four iteration counts.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r1g.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r1g.log
tail -4 gpurun_out/pytest_gpu_r1g.log
{
  timeout 120 python tools/kernel_table.py --shape 512,512,512 --stepper rk4 --steps 10
  timeout 120 python tools/kernel_table.py --shape 512,512,512 --stepper rk2 --steps 10
  timeout 120 python tools/kernel_table.py --shape 512,512,512 --stepper euler --steps 10
  timeout 120 python tools/kernel_table.py --shape 128,128,128 --stepper rk4 --steps 50
  timeout 120 python tools/kernel_table.py --shape 4096,4096,1 --stepper rk4 --steps 10
  timeout 120 python tools/kernel_table.py --shape 4096,4096,1 --stepper euler --steps 10
  timeout 120 python tools/kernel_table.py --shape 33,33,1 --method jacobi --iters 2000 --steps 5
  timeout 120 python tools/kernel_table.py --shape 129,129,1 --method jacobi --iters 2000 --steps 5
} > gpurun_out/kernel_tables_r1g.jsonl 2> gpurun_out/kernel_tables_r1g.err
cut -c1-330 gpurun_out/kernel_tables_r1g.jsonl
tail -3 gpurun_out/kernel_tables_r1g.err
timeout 600 python tools/ref_gpu_compare.py --n 512 --iters 40 --workloads poisson_bicgstab,poisson_cg,poisson_jacobi,poisson_rbsor --out gpurun_out/ref_gpu_compare_poisson.jsonl > gpurun_out/ref_gpu_compare_poisson.log 2>&1
cut -c1-500 gpurun_out/ref_gpu_compare_poisson.log
