"""Builds the engine's C-ABI shared library in-tree with nvcc for sm_100a.

    python -m readserver_b200.build [--force]

Output: readserver_b200/librs_b200.so (git-ignored; travels to the GPU box with the snapshot).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# RS_B200_LIB names another output / load path (tuning experiments build variants side by side)
LIB = os.environ.get("RS_B200_LIB") or os.path.join(HERE, "librs_b200.so")
SOURCES = ["rs_capi.cu", "rs_capi_locate.cu", "rs_kernels.cu", "rs_pipeline.cu", "rs_roofline.cu", "rs_peer.cu", "rs_tiles.cu", "rs_gt.cu", "rs_reads.cu", "rs_pop.cu", "rs_encode.cu", "rs_build.cpp"]
HEADERS = ["rs_internal.h", "rs_layout.h", "rs_swar.h", "rs_ops.h", "rs_kernels.h", "rs_build.h", "rs_pop.h", "rs_gt_ops.h", "rs_encode.h",
           os.path.join("..", "..", "include", "readserver_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-O2,-Wall", "--shared", "-cudart", "static"]


def _nvcc() -> str:
    nv = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nv):
        raise RuntimeError("nvcc not found: the engine needs the CUDA toolkit to build")
    return nv


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Each source is compiled to its own object (in parallel, only when it or a header is newer), then linked."""
    if not force and not is_stale():
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    extra = os.environ.get("RS_NVCC_EXTRA", "").split()  # e.g. -DRS_K1_BLOCKS=6 for tuning experiments
    tag = "".join(c if c.isalnum() else "_" for c in " ".join(extra))
    objdir = os.path.join(HERE, "build", tag or "default")
    os.makedirs(objdir, exist_ok=True)
    hdr_time = max(os.path.getmtime(os.path.join(CSRC, h)) for h in HEADERS)
    compile_flags = [f for f in NVCC_FLAGS if f not in ("--shared", "-cudart", "static")]

    def one(src):
        obj = os.path.join(objdir, src.rsplit(".", 1)[0] + ".o")
        path = os.path.join(CSRC, src)
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(hdr_time, os.path.getmtime(path)):
            return obj, ""
        cmd = [_nvcc(), *compile_flags, *extra, "-I", CSRC, "-c", "-o", obj, path]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n" + r.stdout + r.stderr)
        return obj, r.stdout + r.stderr

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        results = list(ex.map(one, SOURCES))
    if verbose:
        print("".join(log for _, log in results))
    cmd = [_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "--shared", "-cudart", "static", "-o", LIB] + [o for o, _ in results]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
