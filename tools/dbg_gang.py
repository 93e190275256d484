import os, sys, time
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
os.environ["CFD_B200_VIRTUAL_RANKS"] = "1"
os.environ["CFD_B200_PEER_DEBUG"] = "1"
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, cases, oracle
from cfd_b200 import abi, api
orc = oracle.Oracle()
def run(G, shape, method, kind="projection", reps=1, **pkw):
    g, f = cases.random_field(*shape, seed=21, scale=0.05)
    p = cases.params(dt=1e-3, amp=(0.1, 0.05)); p.max_iter = 2
    pc = abi.poisson_config(method or "cg", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE, **pkw)
    api.set_poisson(pc); api.set_device_count(G)
    for rep in range(reps):
        ff = f.copy(); fo = f.copy()
        t = time.time(); rc = api.solve_method_gpu(kind, ff, g, p); dt = time.time() - t
        rco = orc.projection(fo, g, p, pc, 2)[0] if kind == "projection" else orc.rk(fo, g, p, 4, 2)
        same = all(np.array_equal(getattr(ff, k), getattr(fo, k)) for k in "uvwp")
        print("G=%d %s %s %s rep %d: rc=%d/%d bitwise=%s %.2fs %s" % (G, shape, kind, method, rep, rc, rco, same, dt, api.last_error() if rc else ""), flush=True)
    api.load().cfd_b200_release_cached_sessions()
import subprocess
mode = sys.argv[1] if len(sys.argv) > 1 else "driver"
if mode == "driver":
    for env, args in (({}, "3 3"), ({}, "3 30"), ({}, "4 30"), ({}, "2 31u"), ({"CUDA_MODULE_LOADING": "LAZY"}, "3 30")):
        e = dict(os.environ); e.update(env)
        r = subprocess.run([sys.executable, __file__, "case"] + args.split(), env=e, capture_output=True, text=True)
        print(env, args, "->", (r.stdout + r.stderr).strip()[-700:], flush=True)
else:
    G = int(sys.argv[2]); it = sys.argv[3]
    shape = (34, 22, 29) if it.endswith("u") else (34, 22, 26)
    run(G, shape, "jacobi", max_iterations=int(it.rstrip("u")))
