#!/bin/bash
# One parametrised GPU job script (run on the box through gpurun):
#
#   gpurun --timeout 1500 -- 'bash tools/gpu_pass.sh TAG stage [stage ...]'
#
# Every stage writes into gpurun_out/ with the TAG prefix; what is worth keeping is copied to profiles/ by hand.
#   tests        pytest -m gpu (whole suite)                      tests:<expr>  pytest -m gpu -k <expr>
#   smoke        __graft_entry__.smoke()
#   bench        python bench.py (default run)                    bench:<args>  extra bench.py arguments
#   parity:N     tests/multigpu_parity.py under torchrun, N ranks (needs gpurun --gpus N)
#   scale:N      torchrun bench.py --gpus N
#   tables       tools/kernel_table.py over the BASELINE.json shapes
#   (box time on one B200: tests 4 min; smoke + bench 1.5 min; launches + ncu 20 min -- the two ncu stages replay
#    9-ms kernels ~40 times each; every table stage 1-2 min)
#   launches     ncu launch list (gpu__time_duration.sum) of the default bench command + share summary
#   ncu          ncu --set full of the headline kernels -> profiles-ready CSV (tools/ncu_summary.py)
#   ncu512       ncu --set full of the 512^3 kernels that are not part of the headline
#   abtile       kernel tables with the shipped library and with the legacy tile geometry (libcfd_b200_legacy.so)
#   mgpcg        time to solution of mgpcg / pcg / multigrid at 4096^2, 1024^2, 256^3, 512^3 (not run in round 2: no budget left)
#   c4tables     kernel tables of the CG / PCG paths at 4096^2 and 512^3
#   ncuc4        ncu --set full of the C4 kernels (4096^2, PCG + energy)
#   ncurb        ncu --set full of the one-pass Red-Black SOR sweeps (4096^2, 256^3)
#   c4waves      2-D pointwise kernels: blocks per launch (CFD_B200_PW_WAVES) at the C4 slab sizes
#   c3slabs:N    tools/c3_slabs.py on 1 and N GPUs: one-pass Red-Black SOR on slabs against the half-sweeps
#   rbtables     Red-Black SOR: one-pass sweeps (3-D, 2-D) against the half-sweeps, 2-D tiles per block
#   sass         SASS opcode histogram of the shipped cubins (runs anywhere)
TAG=$1; shift
mkdir -p gpurun_out
O=gpurun_out/$TAG
PORT_BASE=29500
torchrun_() { n=$1; shift; PORT_BASE=$((PORT_BASE + 17)); python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((PORT_BASE + n)) "$@"; }
for stage in "$@"; do
  arg=${stage#*:}; [ "$arg" = "$stage" ] && arg=""
  case ${stage%%:*} in
    tests)
      if [ -n "$arg" ]; then timeout 1500 python -m pytest tests -m gpu -q -k "$arg" > ${O}_pytest.log 2>&1
      else timeout 1500 python -m pytest tests -m gpu -q > ${O}_pytest.log 2>&1; fi
      echo "pytest exit $?" >> ${O}_pytest.log; tail -15 ${O}_pytest.log | cut -c1-400 ;;
    smoke)
      timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > ${O}_smoke.log 2>&1; echo "smoke exit $?" >> ${O}_smoke.log
      tail -3 ${O}_smoke.log | cut -c1-300 ;;
    bench)
      timeout 900 python bench.py $arg > ${O}_bench.json 2> ${O}_bench.err; echo "bench exit $?"
      cut -c1-600 ${O}_bench.json; tail -3 ${O}_bench.err ;;
    parity)
      timeout 900 bash -c "$(declare -f torchrun_); torchrun_ $arg tests/multigpu_parity.py" > ${O}_parity_${arg}gpu.log 2>&1
      echo "parity exit $?" >> ${O}_parity_${arg}gpu.log; grep -E "ok|FAIL|PARITY|exit|Error" ${O}_parity_${arg}gpu.log | tail -50 | cut -c1-200 ;;
    scale)
      n=${arg%% *}; extra=${arg#$n}; tagx=$(echo "$extra" | grep -o "config [a-z0-9]*" | tr -d " ")
      timeout 1200 bash -c "$(declare -f torchrun_); torchrun_ $n bench.py --gpus $n $extra" > ${O}_bench_${n}gpu${tagx}.json 2> ${O}_bench_${n}gpu${tagx}.err
      echo "scale exit $?"; cut -c1-500 ${O}_bench_${n}gpu${tagx}.json; grep -v "^\[rank\|^ *File\|^W1\|^E1\|^\*\*\*\|Setting OMP" ${O}_bench_${n}gpu${tagx}.err | tail -4 | cut -c1-300 ;;
    c3slabs)
      out=${O}_c3_slabs.jsonl; : > $out
      timeout 300 python tools/c3_slabs.py 2>/dev/null | grep '^{' >> $out
      if [ -n "$arg" ] && [ "$arg" != "1" ]; then
        timeout 600 bash -c "$(declare -f torchrun_); PORT_BASE=29600; torchrun_ $arg tools/c3_slabs.py" 2>/dev/null | grep '^{' >> $out
        timeout 600 bash -c "$(declare -f torchrun_); PORT_BASE=29700; torchrun_ $arg tools/c3_slabs.py --size 512 --steps 5" 2> ${O}_c3_512.err | grep '^{' >> $out
      fi
      python -c "
import json,sys
for l in open('$out'):
    d=json.loads(l); print(d['workload'][:22], d['n_gpus'], d['path'], d['transport'], d['ms_per_step'], [(k['kernel'],k['launches'],k['avg_us']) for k in d['kernels'][:4]])" ;;
    abtile)
      # A/B of the tile geometry: the shipped library (own cells of a tile start on a 512-byte boundary) against
      # cfd_b200/libcfd_b200_legacy.so (built with -DCFDB_TILE_I0=1: the round-1 geometry, one element off it)
      out=${O}_ab_tile.jsonl; : > $out
      run() { timeout 200 python tools/kernel_table.py "$@" 2>/dev/null | tail -1 >> $out; }
      cp cfd_b200/libcfd_b200.so gpurun_out/_new.so
      for which in aligned legacy; do
        if [ $which = legacy ]; then cp cfd_b200/libcfd_b200_legacy.so cfd_b200/libcfd_b200.so; fi
        echo "{\"library\": \"$which\"}" >> $out
        run --shape 1024,1024,1024 --stepper projection --method bicgstab --iters 10 --steps 1 --warmup 1
        run --shape 512,512,512 --stepper rk4
        run --shape 512,512,512 --stepper projection --method jacobi --iters 50
        run --shape 512,512,512 --stepper projection --method rbsor --iters 50
        run --shape 512,512,512 --stepper projection --method cg --iters 50 --energy
        run --shape 4096,4096,1 --stepper projection --method pcg --iters 50 --energy
        run --shape 256,256,256 --stepper projection --method rbsor --iters 50
        run --shape 128,128,128 --stepper rk4 --steps 20
      done
      cp gpurun_out/_new.so cfd_b200/libcfd_b200.so; rm gpurun_out/_new.so
      python - $out <<'PY'
import json, sys
lib = None; rows = {}
for l in open(sys.argv[1]):
    d = json.loads(l)
    if "library" in d: lib = d["library"]; continue
    key = (tuple(d["shape"]), d["stepper"], d["method"] if d["stepper"] == "projection" else "")
    rows.setdefault(key, {})[lib] = d
for key, v in rows.items():
    if "aligned" not in v or "legacy" not in v: print(key, "incomplete"); continue
    a, b = v["aligned"], v["legacy"]
    print(key, "ms/step %.3f vs %.3f (legacy)" % (a["ms_per_step"], b["ms_per_step"]))
    kb = {k["kernel"]: k for k in b["kernels"]}
    for k in a["kernels"][:6]:
        if k["kernel"] in kb: print("    %-24s %9.2f us %.3f   legacy %9.2f us %.3f" % (k["kernel"], k["avg_us"], k["frac_of_measured_peak"], kb[k["kernel"]]["avg_us"], kb[k["kernel"]]["frac_of_measured_peak"]))
PY
      ;;
    mgpcg)
      # time to solution (relative residual 1e-6) of the stand-alone Poisson solve: multigrid-preconditioned CG against
      # the reference's Jacobi-PCG and against the stationary-system multigrid, random zero-mean rhs
      out=${O}_mgpcg.jsonl; : > $out
      run() { timeout 280 python tools/kernel_table.py "$@" 2>/dev/null | tail -1 >> $out; }
      for shp in 4096,4096,1 1024,1024,1 256,256,256 512,512,512; do
        for m in mgpcg pcg multigrid; do run --shape $shp --stepper poisson --method $m --tol 1e-6 --iters 20000 --steps 1 --warmup 1; done
      done
      python -c "
import json
for l in open('$out'):
    d=json.loads(l); print(d['shape'], d['method'], 'iterations', d['poisson_iterations'], 'ms', round(d['ms_per_step'],2))" ;;
    c4tables)
      out=${O}_c4_tables.jsonl; : > $out
      run() { timeout 150 python tools/kernel_table.py "$@" 2>/dev/null | tail -1 >> $out; }
      run --shape 4096,4096,1 --stepper projection --method pcg --iters 50 --energy
      run --shape 4096,4096,1 --stepper projection --method cg --iters 50
      run --shape 512,512,512 --stepper projection --method cg --iters 50
      run --shape 512,512,512 --stepper projection --method pcg --iters 50
      python -c "
import json
for l in open('$out'):
    d=json.loads(l); print(d['shape'], d['method'], round(d['ms_per_step'],3), [(k['kernel'],k['avg_us'],k['frac_of_measured_peak']) for k in d['kernels'][:3]])" ;;
    rbzc)
      out=${O}_rb_zchunk.jsonl; : > $out
      run() { timeout 150 python tools/kernel_table.py "$@" 2>/dev/null | tail -1 >> $out; }
      for shp in 256,256,129 256,256,256 512,512,257; do
        run --shape $shp --stepper projection --method rbsor --iters 50
        for zc in 16 19 22 26 32 43 64 127; do CFD_B200_ZCHUNK_RB=$zc run --shape $shp --stepper projection --method rbsor --iters 50; done
      done
      python -c "
import json
for l in open('$out'):
    d=json.loads(l); k=[k for k in d['kernels'] if 'rbsor' in k['kernel']][0]; print(d['shape'], d['env'].get('CFD_B200_ZCHUNK_RB','default'), k['avg_us'], k['frac_of_measured_peak'])" ;;
    rbtables)
      out=${O}_rb_tables.jsonl; : > $out
      run() { timeout 150 python tools/kernel_table.py "$@" 2>/dev/null | tail -1 >> $out; }
      for shp in 4096,4096,1 1024,1024,1 256,256,256; do
        run --shape $shp --stepper projection --method rbsor --iters 50
        CFD_B200_NO_RB_FUSED=1 run --shape $shp --stepper projection --method rbsor --iters 50
      done
      for yc in 16; do CFD_B200_YCHUNK_RB=$yc run --shape 4096,4096,1 --stepper projection --method rbsor --iters 50; done
      run --shape 4096,4096,1 --stepper projection --method rbsor --iters 50 --fast
      for zc in 29; do CFD_B200_ZCHUNK_RB=$zc run --shape 256,256,256 --stepper projection --method rbsor --iters 50; done
      run --shape 512,512,512 --stepper projection --method rbsor --iters 50
      python - $out <<'PY'
import json, sys
for l in open(sys.argv[1]):
    try: d = json.loads(l)
    except Exception: print(l[:200]); continue
    ks = {k["kernel"]: (k["launches"], k["avg_us"], k["frac_of_measured_peak"]) for k in d.get("kernels", []) if "rb" in k["kernel"] or "redblack" in k["kernel"] or "linf" in k["kernel"]}
    print(d.get("shape"), d.get("arithmetic"), d.get("env", ""), round(d.get("ms_per_step", 0), 3), ks)
PY
      ;;
    tables)
      out=${O}_kernel_tables.jsonl; : > $out
      run() { timeout 150 python tools/kernel_table.py "$@" 2>/dev/null | tail -1 >> $out; }
      for m in jacobi rbsor cg pcg bicgstab; do run --shape 512,512,512 --stepper projection --method $m --iters 50; done
      run --shape 512,512,512 --stepper rk2; run --shape 512,512,512 --stepper rk4; run --shape 512,512,512 --stepper euler
      run --shape 512,512,512 --stepper projection --method cg --iters 50 --energy
      run --shape 128,128,128 --stepper rk4 --steps 20; run --shape 256,256,256 --stepper projection --method rbsor --iters 100
      run --shape 4096,4096,1 --stepper projection --method pcg --iters 50 --energy
      run --shape 4096,4096,1 --stepper projection --method rbsor --iters 50
      run --shape 129,129,1 --stepper projection --method jacobi --iters 50
      run --shape 1024,1024,1024 --stepper projection --method rbsor --iters 20 --steps 1
      for st in "--stepper rk4" "--stepper euler" "--stepper rk2" "--stepper projection --method jacobi --iters 50" \
                "--stepper projection --method rbsor --iters 50" "--stepper projection --method cg --iters 50 --energy"; do
        run --shape 512,512,512 $st --fast; done
      run --shape 128,128,128 --stepper rk4 --steps 20 --fast; run --shape 128,128,128 --stepper rk4 --steps 200 --no-profile
      wc -l $out ;;
    launches)
      timeout 600 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > ${O}_launch_bench.json 2>/dev/null &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file ${O}_launches.csv \
          python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > ${O}_ncu_launches.log 2>&1
      python tools/launch_list_summary.py ${O}_launches.csv ${O}_launch_bench.json > ${O}_launch_list_summary.md; tail -20 ${O}_launch_list_summary.md ;;
    ncu)
      timeout 600 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu > /dev/null 2>&1 &&
      timeout 1200 ncu --set full --clock-control none --import-source on \
          -k regex:'bicg_dir_tma|bicg_s_tma|bicg_update|predictor_tma|divergence_tma|corrector_tma|krylov_init' \
          --launch-skip 0 -c 16 -o ${O}_ncu_headline -f python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu > ${O}_ncu_full.log 2>&1
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:'corrector_tma' -c 1 \
          -o ${O}_ncu_corrector -f python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu >> ${O}_ncu_full.log 2>&1
      box=$(python -c "import json;print(json.load(open('${O}_launch_bench.json'))['config']['grid_global'][0])" 2>/dev/null || echo 1024)
      python tools/ncu_summary.py "ncu --set full --clock-control none, python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu; box=$box; $(date -u +%F) $TAG" \
          ${O}_ncu_headline.csv ${O}_ncu_headline.ncu-rep ${O}_ncu_corrector.ncu-rep
      tail -3 ${O}_ncu_full.log | cut -c1-200
      for f in ${O}_ncu_headline.ncu-rep ${O}_ncu_corrector.ncu-rep; do if [ $(stat -c %s $f) -gt 25000000 ]; then rm $f; fi; done ;;
    ncu512)
      cat > /tmp/prof512.py <<'PY'
import subprocess, sys
for args in (["--stepper", "projection", "--method", "jacobi", "--iters", "3"],
             ["--stepper", "projection", "--method", "rbsor", "--iters", "3"],
             ["--stepper", "projection", "--method", "cg", "--iters", "3"],
             ["--stepper", "rk4"], ["--stepper", "euler"]):
    r = subprocess.run([sys.executable, "tools/kernel_table.py", "--shape", "512,512,512", "--steps", "1"] + args,
                       capture_output=True, text=True)
    if r.returncode != 0:
        print(r.stdout[-2000:], r.stderr[-2000:]); sys.exit(1)
PY
      timeout 300 python /tmp/prof512.py &&
      timeout 900 ncu --set full --clock-control none --import-source on --target-processes all \
          -k regex:'jacobi_tma|rbsor_fused|cg_dir_tma|cg_update|rk_stage_tma|predictor_tma|corrector_tma|euler_tma' \
          -c 40 -o ${O}_ncu_512 -f python /tmp/prof512.py > ${O}_ncu_512.log 2>&1
      python tools/ncu_summary.py "ncu --set full --clock-control none, tools/kernel_table.py --shape 512,512,512 (jacobi, rbsor, cg, rk4, euler); $(date -u +%F) $TAG" \
          ${O}_ncu_512.csv ${O}_ncu_512.ncu-rep
      if [ $(stat -c %s ${O}_ncu_512.ncu-rep) -gt 40000000 ]; then rm ${O}_ncu_512.ncu-rep; fi ;;
    ncuc4)
      cat > gpurun_out/prof_c4.py <<'PY'
import subprocess, sys
r = subprocess.run([sys.executable, "tools/kernel_table.py", "--shape", "4096,4096,1", "--steps", "1", "--warmup", "1",
                    "--stepper", "projection", "--method", "pcg", "--iters", "4", "--energy"], capture_output=True, text=True)
if r.returncode != 0:
    print(r.stdout[-2000:], r.stderr[-2000:]); sys.exit(1)
PY
      timeout 300 python gpurun_out/prof_c4.py &&
      timeout 900 ncu --set full --clock-control none --import-source on --target-processes all \
          -k regex:'cg_update|cg_dir_tma|predictor_tma|corrector_tma|krylov_init|k_copy|field_stats' --launch-skip 4 -c 14 -o ${O}_ncu_c4 -f python gpurun_out/prof_c4.py > ${O}_ncu_c4.log 2>&1
      python tools/ncu_summary.py "ncu --set full --clock-control none, tools/kernel_table.py --shape 4096,4096,1 --method pcg --energy (C4 kernels); $(date -u +%F) $TAG" \
          ${O}_ncu_c4.csv ${O}_ncu_c4.ncu-rep
      if [ $(stat -c %s ${O}_ncu_c4.ncu-rep) -gt 40000000 ]; then rm ${O}_ncu_c4.ncu-rep; fi ;;
    ncurb)
      cat > gpurun_out/prof_rb.py <<'PY'
import subprocess, sys
for shape in ("4096,4096,1", "256,256,256"):
    r = subprocess.run([sys.executable, "tools/kernel_table.py", "--shape", shape, "--steps", "1", "--warmup", "1",
                        "--stepper", "projection", "--method", "rbsor", "--iters", "4"], capture_output=True, text=True)
    if r.returncode != 0:
        print(r.stdout[-2000:], r.stderr[-2000:]); sys.exit(1)
PY
      timeout 300 python gpurun_out/prof_rb.py &&
      timeout 900 ncu --set full --clock-control none --import-source on --target-processes all \
          -k regex:'rbsor_fused' --launch-skip 2 -c 12 -o ${O}_ncu_rb -f python gpurun_out/prof_rb.py > ${O}_ncu_rb.log 2>&1
      python tools/ncu_summary.py "ncu --set full --clock-control none, tools/kernel_table.py --method rbsor at 4096^2 and 256^3; $(date -u +%F) $TAG" \
          ${O}_ncu_rb.csv ${O}_ncu_rb.ncu-rep
      if [ $(stat -c %s ${O}_ncu_rb.ncu-rep) -gt 40000000 ]; then rm ${O}_ncu_rb.ncu-rep; fi ;;
    c4waves)
      out=${O}_c4_pw_waves.jsonl; : > $out
      for w in 6 3 2 1; do for shape in 4096,4096,1 4096,2050,1 4096,1026,1; do
        CFD_B200_PW_WAVES=$w timeout 150 python tools/kernel_table.py --shape $shape --stepper projection --method pcg --iters 50 --energy 2>/dev/null | tail -1 |
          python -c "import json,sys; d=json.loads(sys.stdin.read()); print(json.dumps({'pw_waves': $w, 'shape': d['shape'], 'ms_per_step': round(d['ms_per_step'],3), 'kernels': {k['kernel']: [k['avg_us'], k['frac_of_measured_peak']] for k in d['kernels'][:3]}}))" >> $out
      done; done; cat $out ;;
    sass)
      python tools/sass_histogram.py > ${O}_sass_histogram.md; head -30 ${O}_sass_histogram.md ;;
    *) echo "unknown stage $stage" ;;
  esac
done
ls -la gpurun_out | tail -20
