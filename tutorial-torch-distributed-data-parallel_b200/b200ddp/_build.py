"""In-tree build of libb200ddp.so with nvcc for sm_100a (no JIT cache, the .so travels with the repo)."""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(os.path.dirname(PKG_DIR), "csrc")
REPO = os.path.dirname(os.path.dirname(PKG_DIR))
LIB_PATH = os.path.join(PKG_DIR, "libb200ddp.so")
OBJ_DIR = os.path.join(CSRC, "build")

SOURCES = ["abi.cu", "adam.cu", "reduce.cu", "comm.cu", "peer_allreduce.cu", "plan.cu", "linear.cu", "gemm_tf32.cu", "conv.cu", "conv1.cu", "conv_fp32.cu", "loss.cu",
           "pool.cu", "bn.cu", "eval.cu", "transform.cu", "engine.cu"]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]
# developer builds only, e.g. B200DDP_NVCC_DEFINES=-DB200_CONV_TRACE (clock64 trace of one conv CTA, scripts/conv_trace.py)
NVCC_FLAGS += [f for f in os.environ.get("B200DDP_NVCC_DEFINES", "").split() if f.startswith("-D")]


def _nccl_include() -> str:
    import importlib.util
    spec = importlib.util.find_spec("nvidia.nccl")
    if spec is not None and spec.submodule_search_locations:
        inc = os.path.join(list(spec.submodule_search_locations)[0], "include")
        if os.path.exists(os.path.join(inc, "nccl.h")):
            return inc
    for cand in ("/usr/include", "/usr/local/cuda/include"):
        if os.path.exists(os.path.join(cand, "nccl.h")):
            return cand
    raise RuntimeError("nccl.h not found (needed to compile comm.cu)")


def nccl_library_path() -> str | None:
    import importlib.util
    spec = importlib.util.find_spec("nvidia.nccl")
    if spec is not None and spec.submodule_search_locations:
        p = os.path.join(list(spec.submodule_search_locations)[0], "lib", "libnccl.so.2")
        if os.path.exists(p):
            return p
    return None


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _digest() -> str:
    h = hashlib.sha256()
    files = sorted(os.listdir(CSRC))
    for f in files:
        if f.endswith((".cu", ".cuh", ".h")):
            with open(os.path.join(CSRC, f), "rb") as fh:
                h.update(f.encode())
                h.update(fh.read())
    with open(os.path.join(REPO, "include", "b200ddp.h"), "rb") as fh:
        h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a and link libb200ddp.so next to this file."""
    stamp = LIB_PATH + ".sha256"
    dig = _digest()
    if not force and os.path.exists(LIB_PATH) and os.path.exists(stamp) and open(stamp).read().strip() == dig:
        return LIB_PATH
    os.makedirs(OBJ_DIR, exist_ok=True)
    nvcc = _nvcc()
    inc = ["-I", _nccl_include(), "-I", os.path.join(REPO, "include")]
    logs = {}

    def compile_one(src: str) -> str:
        obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, *inc, "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        logs[src] = r.stderr
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, "-shared", "-o", LIB_PATH, *objs, "-ldl", "-Xcompiler", "-fPIC",
           "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(os.path.join(OBJ_DIR, "ptxas.log"), "w") as fh:
        for k, v in logs.items():
            fh.write(f"==== {k}\n{v}\n")
    with open(stamp, "w") as fh:
        fh.write(dig)
    if verbose:
        for k, v in logs.items():
            sys.stderr.write(f"==== {k}\n{v}\n")
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
