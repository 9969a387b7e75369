"""Build libhf2d_gpu.so (the C-ABI shared library with the sm_100a kernels) in-tree with nvcc.

-fmad=false: the parity contract is the reference compiled for x86-64 without FMA contraction; IEEE division and
square root are nvcc's default for double.  -lineinfo so ncu's source page maps to the .cuh files."""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, os.environ.get("HF2D_SO_NAME", "libhf2d_gpu.so"))   # (HF2D_SO_NAME / HF2D_NVCC_EXTRA: tuning experiments)
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _newest_src() -> float:
    t = 0.0
    for root in (CSRC, os.path.join(os.path.dirname(HERE), "include")):
        for f in os.listdir(root):
            t = max(t, os.path.getmtime(os.path.join(root, f)))
    return t


def source_hash() -> str:
    """sha256 over the sources the library is built from (csrc/ and include/), written beside the library at build time."""
    h = hashlib.sha256()
    for root in (CSRC, os.path.join(os.path.dirname(HERE), "include")):
        for f in sorted(os.listdir(root)):
            h.update(f.encode())
            with open(os.path.join(root, f), "rb") as fh:
                h.update(fh.read())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and os.path.exists(SO) and os.path.getmtime(SO) >= _newest_src():
        return SO
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        # a box without a toolchain: the prebuilt library that travelled with the snapshot -- but only if it was built from
        # THESE sources (a run on an old library must not pass unnoticed)
        if os.path.exists(SO):
            want = source_hash()
            have = open(SO + ".srchash").read().strip() if os.path.exists(SO + ".srchash") else "(no hash recorded)"
            if have != want:
                raise RuntimeError(f"nvcc not found and the prebuilt {SO} was built from other sources (library {have[:12]}, tree {want[:12]})")
            print(f"openhyperflow2d_b200: no nvcc here, using the prebuilt {SO} (sources {want[:12]})", file=sys.stderr)
            return SO
        raise RuntimeError("nvcc not found and no prebuilt libhf2d_gpu.so")
    cmd = [nvcc] + NVCC_FLAGS + os.environ.get("HF2D_NVCC_EXTRA", "").split() + ["-o", SO, os.path.join(CSRC, "hf2d_gpu.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log = r.stdout + r.stderr
    with open(os.path.join(HERE, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if verbose:
        print(log)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + log[-4000:])
    with open(SO + ".srchash", "w") as f:
        f.write(source_hash() + "\n")
    return SO


if __name__ == "__main__":
    print(build(force=True, verbose=True))
