"""Build recipe: nvcc -> cellector_b200/libcellector_cuda.so (sm_100a only), g++ -> oracle, host CLI.

Run as `python -m cellector_b200.build` or through `__graft_entry__.build()`.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libcellector_cuda.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

CU_SOURCES = ["abi.cu", "load.cu", "em.cu", "comm.cu", "tiled.cu", "p2p.cu", "noalloc.cu", "soup.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "-fmad=false",                      # Rust never contracts a*b+c; parity needs the same (numerics.cuh)
    "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC,-ffp-contract=off,-Wall",
    "-Xptxas", "-v",
]


def _newer(src_list, out):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(s) > t for s in src_list)


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hs.append(os.path.join(ROOT, "include", "cellector_cuda.h"))
    return hs


def _compile(cu):
    src = os.path.join(CSRC, cu)
    obj = os.path.join(OBJ, cu.replace(".cu", ".o"))
    if _newer([src] + _headers(), obj):
        cmd = [NVCC, *NVCC_FLAGS, *os.environ.get("CLX_EXTRA_NVCC", "").split(), "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = os.path.join(OBJ, cu + ".log")
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {cu}:\n{r.stdout}\n{r.stderr}")
    return obj


def build_cuda(verbose=True):
    os.makedirs(OBJ, exist_ok=True)
    with cf.ThreadPoolExecutor(max_workers=4) as ex:
        objs = list(ex.map(_compile, CU_SOURCES))
    if _newer(objs, LIB):
        cmd = [NVCC, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-ldl"]
        subprocess.check_call(cmd)
    if verbose:
        print(f"built {LIB}")
    return LIB


def build_oracle():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])


def build_cli():
    mk = os.path.join(CSRC, "host", "Makefile")
    if os.path.exists(mk):
        subprocess.check_call(["make", "-C", os.path.dirname(mk), "-s"])


def build_all():
    build_cuda()
    build_oracle()
    build_cli()


if __name__ == "__main__":
    build_all()
    sys.exit(0)
