"""Builds rwkv_cppy_b200/librwkv.so (the C-ABI library) from csrc/ with nvcc for sm_100a.

In-tree on purpose: the .so travels with the repo snapshot to the GPU box and is the file
the round-end driver sees loaded.  `python -m rwkv_cppy_b200.build` or
`__graft_entry__.build()`.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "librwkv.so")
# experiment builds (RWKV_B200_BUILD_NAME=trace RWKV_B200_NVCC_FLAGS=-DRWKV_B200_TRACE ...): objects and library go to
# build/variants/ and leave the product library alone
if os.environ.get("RWKV_B200_BUILD_NAME"):
    OBJ = os.path.join(HERE, "build", "variants", "obj_" + os.environ["RWKV_B200_BUILD_NAME"])
    LIB = os.path.join(HERE, "build", "variants", "librwkv_" + os.environ["RWKV_B200_BUILD_NAME"] + ".so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden,-Wall,-Wno-unused-function",
          "-DRWKV_BUILD", "-DRWKV_SHARED"]
COMMON += os.environ.get("RWKV_B200_NVCC_FLAGS", "").split()      # e.g. -DRWKV_B200_TRACE (tools/decode_trace.py)
PER_FILE = {"quantize.cpp": ["-Xcompiler", "-ffp-contract=off"]}


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _stamp() -> str:
    h = hashlib.sha256()
    for f in sorted(os.listdir(CSRC)) + ["../../include/rwkv.h", "../../include/rwkv_b200.h", "../build.py"]:
        p = os.path.join(CSRC, f)
        if os.path.isfile(p):
            h.update(f.encode())
            h.update(open(p, "rb").read())
    h.update(os.environ.get("RWKV_B200_NVCC_FLAGS", "").encode())
    return h.hexdigest()


def _compile(src: str) -> str:
    obj = os.path.join(OBJ, src + ".o")
    cmd = [NVCC, *ARCH, *COMMON, *PER_FILE.get(src, []), "-x", "cu", "-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    if r.stderr.strip():
        sys.stderr.write(r.stderr)
    return obj


def build_library(force: bool = False, verbose: bool = True) -> str:
    os.makedirs(OBJ, exist_ok=True)
    stamp_file = os.path.join(OBJ, "stamp")
    stamp = _stamp()
    if not force and os.path.exists(LIB) and os.path.exists(stamp_file) and open(stamp_file).read() == stamp:
        return LIB
    srcs = _sources()
    if verbose:
        print(f"[rwkv_cppy_b200] nvcc sm_100a: {len(srcs)} sources", flush=True)
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(_compile, srcs))
    cmd = [NVCC, *ARCH, "-shared", "-cudart", "static", "-o", LIB, *objs, "-lpthread", "-ldl", "-lrt"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    open(stamp_file, "w").write(stamp)
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv))
