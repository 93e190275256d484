#!/bin/bash
# Round-1i GPU pass: full GPU suite (cluster CG, staged pageable transfers), end-to-end comparison against
# the reference CUDA backend with pageable host arrays, transfer-thread sweep.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r1i.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_r1i.log
tail -12 gpurun_out/pytest_gpu_r1i.log
nproc
timeout 500 python tools/ref_gpu_compare.py --n 512 --steps 4 --workloads projection,rk4 --out gpurun_out/ref_gpu_compare_e2e_staged.jsonl > gpurun_out/ref_gpu_compare_e2e_staged.log 2>&1
cut -c1-420 gpurun_out/ref_gpu_compare_e2e_staged.log
for t in 0 2 4 8; do
  echo "CFD_B200_XFER_THREADS=$t"
  CFD_B200_XFER_THREADS=$t timeout 200 python tools/ref_gpu_compare.py --n 512 --steps 1 --worker cfd_b200 rk2 2>&1 | tail -1 | cut -c1-300
done > gpurun_out/xfer_thread_sweep.log 2>&1
cat gpurun_out/xfer_thread_sweep.log
