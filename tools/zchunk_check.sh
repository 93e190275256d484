#!/bin/bash
# default chunk choice after the small-grid rule: the C2 / small-grid lines of tools/zchunk_sweep.sh without the knob
out=gpurun_out/zchunk_default.jsonl; mkdir -p gpurun_out; : > $out
for cfg in "128,128,128 --stepper rk4 --steps 200" "128,128,128 --stepper rk4 --steps 200 --fast" "256,256,256 --stepper rk4 --steps 50" "128,128,128 --stepper projection --method jacobi --iters 50 --steps 20" "128,128,128 --stepper projection --method cg --iters 50 --steps 20" "128,128,128 --stepper euler --steps 200" "64,64,64 --stepper rk4 --steps 200" "192,192,192 --stepper rk4 --steps 50"; do
  timeout 120 python tools/kernel_table.py --shape $cfg --no-profile 2>/dev/null | tail -1 |
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print(json.dumps({'shape': d['shape'], 'arith': d['arithmetic'], 'stepper': d['stepper'], 'method': d['method'], 'ms_per_step': round(d['ms_per_step'],4), 'MLUPS': round(d['MLUPS'])}))" >> $out
done
cat $out
