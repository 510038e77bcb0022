"""Builds gridss_b200/libcudasw.so in-tree with nvcc for sm_100a (no JIT cache; the .so travels with gpurun)."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
OUT = os.path.join(HERE, "libcudasw.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]
GROUPS = (1, 2, 4, 8, 16, 32)


def _deps():
    d = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    d.append(os.path.join(HERE, "..", "include", "cudasw.h"))
    return d


def up_to_date() -> bool:
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    return all(os.path.getmtime(d) <= t for d in _deps())


def build(force: bool = False, verbose: bool = False, check: bool = False) -> str:
    """check=True: the -DCUSW_BOUNDS_CHECK flavour (every data-dependent global index trapped when out of range) as
    libcudasw_check.so beside the product library; tests load it with CUDASW_LIB=<path> (gridss_b200/aligner.py)."""
    global OBJ, OUT
    if check:
        OBJ, OUT = os.path.join(HERE, "build_check"), os.path.join(HERE, "libcudasw_check.so")
    if not force and up_to_date():
        return OUT
    os.makedirs(OBJ, exist_ok=True)
    extra = (["-Xptxas", "-v"] if verbose else []) + (["-DCUSW_BOUNDS_CHECK"] if check else []) + os.environ.get("CUDASW_NVCC_EXTRA", "").split()
    jobs = [(os.path.join(OBJ, "cudasw.o"), [os.path.join(CSRC, "cudasw.cu")]),
            (os.path.join(OBJ, "variants.o"), [os.path.join(CSRC, "variants.cu")]),
            (os.path.join(OBJ, "multi.o"), [os.path.join(CSRC, "multi.cpp")]),
            (os.path.join(OBJ, "rev_variants.o"), [os.path.join(CSRC, "rev_variants.cu")]),
            (os.path.join(OBJ, "noalloc.o"), [os.path.join(CSRC, "noalloc.cu")])]
    for g in GROUPS:
        jobs.append((os.path.join(OBJ, f"score_g{g}.o"), [f"-DSCORE_G={g}", os.path.join(CSRC, "score_variants.cu")]))

    def cc(job):
        out, args = job
        subprocess.check_call([NVCC] + FLAGS + extra + ["-c", "-o", out] + args)
        return out

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(cc, jobs))
    subprocess.check_call([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "shared", "-o", OUT] + objs)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv, check="--check" in sys.argv)
    print(OUT)
