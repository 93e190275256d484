"""Build libcfd_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m cfd_b200.build [--force]

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# A/B builds: CFD_B200_NVCC_EXTRA="-DCFDB_TILE_I0=1" CFD_B200_LIB_OUT=libcfd_b200_legacy.so python -m cfd_b200.build
_OUT = os.environ.get("CFD_B200_LIB_OUT", "libcfd_b200.so")
OBJ = os.path.join(HERE, "_build" if _OUT == "libcfd_b200.so" else "_build_" + _OUT.replace(".so", ""))
LIB = os.path.join(HERE, _OUT)
SOURCES = ["poisson_kernels.cu", "poisson_mini.cu", "poisson_mg.cu", "poisson_mgpc.cu", "ns_kernels.cu", "ns_kernels2.cu", "session.cu", "session_io.cu", "xfer.cu", "tma_host.cu",
           "poisson_driver.cu", "step_driver.cu", "capi.cu", "capi_vtables.cu", "bc_api.cu", "gang.cu", "host_sync.cu", "peer.cu", "no_alloc.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-fmad=false", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC,-fvisibility=hidden", "-Xptxas", "-v"] + os.environ.get("CFD_B200_NVCC_EXTRA", "").split()


def _newest_dep():
    t = 0.0
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(root):
            t = max(t, os.path.getmtime(os.path.join(root, f)))
    return t


# kernel translation units that are also built with the "fast" arithmetic policy (csrc/common.cuh): FMA contraction on,
# quotients as reciprocal multiplies; their kernels / launchers land in cfdb::arith_fast
FAST_TWINS = ["poisson_kernels.cu", "ns_kernels.cu", "ns_kernels2.cu"]
FAST_FLAGS = [f for f in FLAGS if f != "-fmad=false"] + ["-fmad=true", "-DCFDB_FAST"]


def _compile(job):
    src, fast = job
    obj = os.path.join(OBJ, src.replace(".cu", "_fast.o" if fast else ".o"))
    r = subprocess.run([NVCC, *(FAST_FLAGS if fast else FLAGS), "-c", os.path.join(CSRC, src), "-o", obj],
                       capture_output=True, text=True)
    with open(obj + ".log", "w") as fh:
        fh.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed on %s%s:\n%s" % (src, " (fast)" if fast else "", (r.stdout + r.stderr)[-4000:]))
    return obj


def build(force=False, verbose=False):
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= _newest_dep():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        objs = list(ex.map(_compile, [(f, False) for f in SOURCES] + [(f, True) for f in FAST_TWINS]))
    r = subprocess.run([NVCC, "-shared", "-o", LIB, *objs, "-ldl", "-lpthread"],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    if verbose:
        print("built", LIB)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
