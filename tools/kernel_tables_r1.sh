#!/bin/bash
# Per-kernel tables (tools/kernel_table.py) of every solver / stepper on the BASELINE.json grid shapes
# that fit the box; one JSON line per case -> gpurun_out/kernel_tables.jsonl
out=gpurun_out/kernel_tables.jsonl; : > $out
run() { timeout 120 python tools/kernel_table.py "$@" 2>/dev/null | tail -1 >> $out; }
for m in jacobi rbsor cg pcg bicgstab; do run --shape 512,512,512 --stepper projection --method $m --iters 50; done
run --shape 512,512,512 --stepper rk2
run --shape 512,512,512 --stepper rk4
run --shape 512,512,512 --stepper projection --method cg --iters 50 --energy
run --shape 128,128,128 --stepper rk4
run --shape 128,128,128 --stepper projection --method cg --iters 50
run --shape 4096,4096,1 --stepper projection --method pcg --iters 50
run --shape 129,129,1 --stepper projection --method jacobi --iters 50
run --shape 1024,1024,1024 --stepper projection --method rbsor --iters 20 --steps 1
wc -l $out
