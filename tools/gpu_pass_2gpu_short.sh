#!/bin/bash
# short 2-GPU sanity pass: bench line at 256^3 per GPU, then the decomposed-vs-oracle parity worker
mkdir -p gpurun_out
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --steps 2 --warmup 3 --box 256 --no-cpu > gpurun_out/bench_2gpu_256.json 2> gpurun_out/bench_2gpu_256.err; echo "bench exit $?"
grep "^{" gpurun_out/bench_2gpu_256.json | cut -c1-260
timeout 90 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 tests/multigpu_parity.py > gpurun_out/multigpu_parity_2gpu.log 2>&1; echo "parity exit $?" >> gpurun_out/multigpu_parity_2gpu.log
grep -v "^W\|^\[W\|Warning\|warn" gpurun_out/multigpu_parity_2gpu.log | tail -8 | cut -c1-160
