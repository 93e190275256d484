#!/bin/bash
# z-chunk length of the TMA plane pipelines on cache-resident grids (CFD_B200_ZCHUNK_TMA): whole-step time without
# per-launch events, one JSON line per (shape, stepper, chunk).  Run on the GPU box: bash tools/zchunk_sweep.sh
out=gpurun_out/zchunk_sweep.jsonl; mkdir -p gpurun_out; : > $out
for cfg in "128,128,128 --stepper rk4 --steps 200" "256,256,256 --stepper rk4 --steps 50" "128,128,128 --stepper projection --method jacobi --iters 50 --steps 20" "128,128,128 --stepper euler --steps 200"; do
  for zc in 0 4 7 9 14 18 21 32 42 63 126; do
    if [ $zc = 0 ]; then unset CFD_B200_ZCHUNK_TMA; else export CFD_B200_ZCHUNK_TMA=$zc; fi
    timeout 120 python tools/kernel_table.py --shape $cfg --no-profile 2>/dev/null | tail -1 |
      python -c "import json,sys; d=json.loads(sys.stdin.read()); print(json.dumps({'zchunk': $zc, 'shape': d['shape'], 'stepper': d['stepper'], 'ms_per_step': round(d['ms_per_step'],4), 'MLUPS': round(d['MLUPS'])}))" >> $out
  done
done
cat $out
