"""Build libhm_engine.so (all CUDA kernels + the C ABI) in-tree for sm_100a.

    python -m hivemind_b200.build [--force]

nvcc cross-compiles without a GPU.  The .so lands next to this file (git-ignored, shipped
to the GPU box by gpurun).
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libhm_engine.so")
OBJ = os.path.join(HERE, "_build")
SOURCES = ["hm_planes.cu", "hm_layers.cu", "hm_policy.cu", "hm_root.cu", "hm_xboard.cu", "hm_net.cu", "hm_abi.cu", "hm_chunk.cc"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
]
# bring-up builds only, e.g. HM_EXTRA_NVCC_FLAGS=-DHM_BRINGUP (timeline tracing in trunk_kernel, tools/trace_trunk.py);
# the default build carries no tracing or debug switches
FLAGS += os.environ.get("HM_EXTRA_NVCC_FLAGS", "").split()


def _deps():
    files = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    files.append(os.path.join(HERE, "..", "include", "hm_engine.h"))
    files.append(os.path.join(HERE, "..", "include", "hm_chunk.h"))
    return files


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(f) > t for f in _deps())


def _compile(src: str, force: bool = False) -> str:
    obj = os.path.join(OBJ, src + ".o")
    srcp = os.path.join(CSRC, src)
    if not force and os.path.exists(obj) and all(os.path.getmtime(f) <= os.path.getmtime(obj) for f in _deps()):
        return obj
    cmd = [NVCC, *FLAGS, "-c", srcp, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return OUT
    os.makedirs(OBJ, exist_ok=True)
    sources = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    with ThreadPoolExecutor(max_workers=len(sources)) as ex:
        objs = list(ex.map(lambda s: _compile(s, force), sources))
    # -Bsymbolic / --exclude-libs: keep our (possibly static) libstdc++ symbols private so
    # the library can be dlopen'ed into a process that already loaded another libstdc++.
    cmd = [NVCC, "-shared", "-o", OUT, *objs, "-cudart", "static",
           "-Xlinker", "-Bsymbolic", "-Xlinker", "--exclude-libs,ALL"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print(f"built {OUT}")
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
