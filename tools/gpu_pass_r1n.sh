#!/bin/bash
# Round-1n: ncu --set full of the kernels changed late in the round (RK stage with group clamps, explicit
# Euler, cluster Jacobi with st.async signalling, cluster CG), each after its plain command exited 0.
mkdir -p gpurun_out
cat > /tmp/prof_r1n.py <<'PY'
import subprocess, sys
for args in (["--shape", "512,512,512", "--stepper", "rk4", "--steps", "1", "--warmup", "0"],
             ["--shape", "512,512,512", "--stepper", "euler", "--steps", "1", "--warmup", "0"],
             ["--shape", "129,129,1", "--method", "jacobi", "--iters", "2000", "--steps", "1", "--warmup", "0"],
             ["--shape", "129,129,1", "--method", "cg", "--iters", "200", "--steps", "1", "--warmup", "0"]):
    r = subprocess.run([sys.executable, "tools/kernel_table.py"] + args, capture_output=True, text=True)
    if r.returncode != 0:
        print(r.stdout[-2000:], r.stderr[-2000:]); sys.exit(1)
PY
timeout 200 python /tmp/prof_r1n.py || exit 1
timeout 500 ncu --set full --clock-control none --import-source on --target-processes all \
    -k regex:'k_rk_stage_tma|k_euler_tma|k_jacobi_cluster|k_cg_cluster' -c 7 -o gpurun_out/prof_r1n -f python /tmp/prof_r1n.py > gpurun_out/ncu_r1n.log 2>&1
tail -2 gpurun_out/ncu_r1n.log
ls -la gpurun_out/prof_r1n.ncu-rep
