#!/bin/bash
# 2-GPU pass: decomposed == undecomposed oracle (incl. explicit Euler), 2-GPU bench line at 512^3 per GPU.
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 tests/multigpu_parity.py > gpurun_out/multigpu_parity_2gpu.log 2>&1; echo "parity exit $?" >> gpurun_out/multigpu_parity_2gpu.log
grep -v "^W\|^\[W\|Warning\|warn" gpurun_out/multigpu_parity_2gpu.log | tail -24 | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --steps 3 --warmup 3 --box 512 --no-cpu > gpurun_out/bench_2gpu_512.json 2> gpurun_out/bench_2gpu_512.err; echo "bench exit $?"
grep "^{" gpurun_out/bench_2gpu_512.json | cut -c1-300
