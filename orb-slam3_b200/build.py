"""In-tree build of liborbx.so (hand-written CUDA for sm_100a, C ABI in include/orbx.h)."""
from __future__ import annotations

import pathlib
import shutil
import subprocess

PKG = pathlib.Path(__file__).resolve().parent
CSRC = PKG / "csrc"
SOURCES = ["capi.cu", "geometry.cu", "prepare.cu", "pyramid.cu", "fast.cu", "octree.cu", "describe.cu", "stereo.cu", "match.cu", "search.cu", "bow.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared"]


def needs_build() -> bool:
    lib = PKG / "liborbx.so"
    if not lib.exists():
        return True
    t = lib.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.h")) + list(CSRC.glob("*.inc")) + list((PKG.parent / "include").glob("*"))
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> pathlib.Path:
    lib = PKG / "liborbx.so"
    if not force and not needs_build():
        return lib
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    import os
    extra = os.environ.get("ORBX_NVCC_EXTRA", "").split()
    cmd = [nvcc, *NVCC_FLAGS, *extra, "-o", str(lib), *[str(CSRC / s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return lib


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
