#!/bin/bash
# One `ncu --set full` pass over the 512^3 kernels that are not part of the headline bench
# (Jacobi, fused Red-Black SOR, CG, RK stage, predictor/corrector).  Run on the GPU box via gpurun,
# only after the plain command has exited 0.  Reports land in gpurun_out/.
set -e
cat > /tmp/prof512.py <<'PY'
import subprocess, sys
for args in (["--stepper", "projection", "--method", "jacobi", "--iters", "3"],
             ["--stepper", "projection", "--method", "rbsor", "--iters", "3"],
             ["--stepper", "projection", "--method", "cg", "--iters", "3"],
             ["--stepper", "rk4"]):
    r = subprocess.run([sys.executable, "tools/kernel_table.py", "--shape", "512,512,512", "--steps", "1"] + args,
                       capture_output=True, text=True)
    if r.returncode != 0:
        print(r.stdout[-2000:], r.stderr[-2000:]); sys.exit(1)
PY
timeout 200 python /tmp/prof512.py
timeout 500 ncu --set full --clock-control none --target-processes all \
    -k regex:'jacobi_tma|rbsor_fused|cg_dir_tma|cg_update|rk_stage_tma|predictor_tma|corrector_tma|divergence_tma|linf_residual_tma' \
    --launch-skip 0 -c 40 -o gpurun_out/prof_r1d_512 -f python /tmp/prof512.py > gpurun_out/ncu_r1d.log 2>&1
tail -2 gpurun_out/ncu_r1d.log
ncu -i gpurun_out/prof_r1d_512.ncu-rep --page raw --csv > gpurun_out/prof_r1d_512_raw.csv 2>/dev/null
ls -la gpurun_out/
# the raw CSV is what profiles/ keeps; drop the report if it would not fit the 64 MiB return limit
if [ $(stat -c %s gpurun_out/prof_r1d_512.ncu-rep) -gt 50000000 ]; then rm gpurun_out/prof_r1d_512.ncu-rep; fi
