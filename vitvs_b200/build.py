"""Builds libvitvs_b200.so (hand-written sm_100a CUDA + the C ABI) in-tree with nvcc.

The .so is git-ignored but travels to the GPU box with the gpurun snapshot; on a box without
nvcc the prebuilt file is used as is.  `python -m vitvs_b200.build [--force]`.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libvitvs_b200.so")
SOURCES = ["api.cu", "gemm_tc.cu", "attention_ts.cu", "elementwise.cu", "match_ibvs.cu", "resize.cu", "gram_stencil.cu", "render.cu"]
HEADERS = ["ptx.cuh", "vvs_internal.h", os.path.join("..", "..", "include", "vitvs_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
]


def _nvcc() -> str | None:
    return shutil.which("nvcc") or (os.path.exists("/usr/local/cuda/bin/nvcc") and "/usr/local/cuda/bin/nvcc") or None


def source_id() -> str:
    """sha256 (first 16 hex digits) over the CUDA sources and headers the library is compiled from, in a fixed order.
    Compiled into the .so as vvs_build_id(); smoke() compares the two on the GPU box, so a binary built from other
    (e.g. uncommitted, since edited) sources cannot pass for the tree it travels with."""
    import hashlib
    h = hashlib.sha256()
    for f in SOURCES + HEADERS:
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(f.encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def binary_id(path: str = LIB) -> str | None:
    """The id compiled into an existing .so (read from the file, not by loading it)."""
    try:
        with open(path, "rb") as fh:
            blob = fh.read()
    except OSError:
        return None
    i = blob.find(b"VVSBUILD:")
    return blob[i + 9:i + 25].decode("ascii", "replace") if i >= 0 else None


def needs_build() -> bool:
    return not os.path.exists(LIB) or binary_id() != source_id()


def build_variant(name: str, defines: list[str]) -> str:
    """Experiment builds (tools/): libvitvs_b200_<name>.so with extra -D flags; load with VITVS_LIB=<path>."""
    nvcc = _nvcc()
    out = os.path.join(HERE, f"libvitvs_b200_{name}.so")
    srcs = [os.path.join(CSRC, f) for f in SOURCES]
    subprocess.check_call([nvcc, *NVCC_FLAGS, f'-DVVS_BUILD_ID="{source_id()}"', *[f"-D{d}" for d in defines], "-shared",
                           "-o", out, *srcs])
    return out


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    nvcc = _nvcc()
    if nvcc is None:
        if os.path.exists(LIB):
            return LIB
        raise RuntimeError("nvcc not found and no prebuilt libvitvs_b200.so present")
    objs = []
    build_dir = os.path.join(HERE, "build")
    os.makedirs(build_dir, exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(build_dir, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc, *NVCC_FLAGS, f'-DVVS_BUILD_ID="{source_id()}"', "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- nvcc {src}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed (see stderr)")
    tmp = LIB + ".tmp"
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp, *objs])
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
