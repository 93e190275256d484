#!/bin/bash
# Round-1f GPU pass (run under gpurun): cluster-Jacobi parity + timing, explicit-Euler kernel table and
# ncu capture, longer RK comparison against the reference CUDA backend, stand-alone BiCGSTAB timing probe.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_poisson.py -x -q > gpurun_out/pytest_poisson.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_poisson.log
tail -4 gpurun_out/pytest_poisson.log
{
  for sh in 129,129,1 33,33,1 65,65,1; do
    timeout 120 python tools/kernel_table.py --shape $sh --method jacobi --iters 2000 --steps 5
    CFD_B200_NO_CLUSTER_JACOBI=1 timeout 120 python tools/kernel_table.py --shape $sh --method jacobi --iters 2000 --steps 5
  done
  timeout 120 python tools/kernel_table.py --shape 512,512,512 --stepper euler --steps 10
  timeout 120 python tools/kernel_table.py --shape 128,128,128 --stepper euler --steps 50
  timeout 120 python tools/kernel_table.py --shape 4096,4096,1 --stepper euler --steps 10
  timeout 200 python tools/kernel_table.py --shape 512,512,512 --stepper poisson --method bicgstab --iters 50 --steps 1 --warmup 1
  timeout 200 python tools/kernel_table.py --shape 512,512,512 --stepper poisson --method bicgstab --iters 150 --steps 1 --warmup 1
  timeout 200 python tools/kernel_table.py --shape 512,512,512 --stepper poisson --method bicgstab --iters 150 --steps 1 --warmup 1 --rhs smooth
} > gpurun_out/kernel_tables_r1f.jsonl 2> gpurun_out/kernel_tables_r1f.err
cut -c1-420 gpurun_out/kernel_tables_r1f.jsonl
tail -3 gpurun_out/kernel_tables_r1f.err
timeout 400 python tools/ref_gpu_compare.py --n 512 --steps 24 --workloads rk4,explicit_euler --out gpurun_out/ref_gpu_compare_rk.jsonl > gpurun_out/ref_gpu_compare_rk.log 2>&1
cat gpurun_out/ref_gpu_compare_rk.log | cut -c1-400
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'k_euler_tma|k_jacobi_cluster' -c 3 -o gpurun_out/prof_r1f -f \
    bash -c 'python tools/kernel_table.py --shape 512,512,512 --stepper euler --steps 1 --warmup 0; python tools/kernel_table.py --shape 129,129,1 --method jacobi --iters 2000 --steps 1 --warmup 0' > gpurun_out/ncu_r1f.log 2>&1
tail -2 gpurun_out/ncu_r1f.log
ncu -i gpurun_out/prof_r1f.ncu-rep --page raw --csv > gpurun_out/prof_r1f_raw.csv 2>/dev/null
ls -la gpurun_out | head -30
