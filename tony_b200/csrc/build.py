"""Build the tony_b200 kernel library for sm_100a, in-tree.

    python -m tony_b200.csrc.build [--force] [--selftest]

Produces ``tony_b200/lib/libtony_b200_kernels.so`` (C ABI, see ``include/tony_b200_kernels.h``)
and, with ``--selftest``, the stand-alone ``tony_b200/lib/tony_selftest`` binary.  nvcc
cross-compiles without a GPU, so this runs on the CPU build box; the artefacts travel to the
GPU box with the repository snapshot.  The reference has no native code to build
(its build is Gradle/Java only: build.gradle:36-91), so this has no counterpart there.
"""
from __future__ import annotations

import argparse
import concurrent.futures
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

CSRC = Path(__file__).resolve().parent
ROOT = CSRC.parent.parent
LIB_DIR = CSRC.parent / "lib"
OBJ_DIR = LIB_DIR / "obj"
INCLUDE = ROOT / "include"
LIB_NAME = "libtony_b200_kernels.so"

SOURCES = [
    "runtime.cu",
    "gemm_tcgen05.cu",
    "layernorm.cu",
    "softmax_xent.cu",
    "colsum.cu",
    "optim.cu",
    "embedding.cu",
    "symm_vmm.cu",
    "attention.cu",
    "embbag.cu",
    "batchnorm.cu",
    "conv_tcgen05.cu",
]

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON_FLAGS = [
    "-O3",
    "-std=c++17",
    "-lineinfo",
    "--use_fast_math",
    "-Xcompiler",
    "-fPIC",
    "-Xcompiler",
    "-fvisibility=default",
    f"-I{INCLUDE}",
    f"-I{CSRC}",
]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; cannot build the sm_100a kernel library")


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(ARCH_FLAGS + COMMON_FLAGS).encode())
    return h.hexdigest()


def existing_sources():
    return [CSRC / s for s in SOURCES if (CSRC / s).exists()]


def _compile_one(nvcc: str, src: Path, verbose: bool) -> Path:
    obj = OBJ_DIR / (src.stem + ".o")
    cmd = [nvcc, *ARCH_FLAGS, *COMMON_FLAGS, "-c", str(src), "-o", str(obj)]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
    if verbose:
        sys.stderr.write(res.stderr)
    return obj


def build(force: bool = False, selftest: bool = False, verbose: bool = False) -> Path:
    """Compile every CUDA source for sm_100a and link the shared library (idempotent)."""
    LIB_DIR.mkdir(exist_ok=True)
    OBJ_DIR.mkdir(exist_ok=True)
    srcs = existing_sources()
    deps = srcs + [CSRC / "common.cuh", CSRC / "host_math.h", INCLUDE / "tony_b200_kernels.h"]
    stamp = LIB_DIR / ".build_stamp"
    digest = _digest(deps)
    lib = LIB_DIR / LIB_NAME
    up_to_date = lib.exists() and stamp.exists() and stamp.read_text() == digest
    if not up_to_date or force:
        nvcc = nvcc_path()
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as pool:
            objs = list(pool.map(lambda s: _compile_one(nvcc, s, verbose), srcs))
        cmd = [nvcc, *ARCH_FLAGS, "-shared", "-o", str(lib), *map(str, objs)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
        stamp.write_text(digest)
    if selftest:
        st_src = CSRC / "selftest.cu"
        st_bin = LIB_DIR / "tony_selftest"
        if force or not st_bin.exists() or st_bin.stat().st_mtime < max(
            st_src.stat().st_mtime, lib.stat().st_mtime
        ):
            nvcc = nvcc_path()
            cmd = [nvcc, *ARCH_FLAGS, "-O3", "-std=c++17", "-lineinfo", f"-I{INCLUDE}", f"-I{CSRC}",
                   str(st_src), "-o", str(st_bin), f"-L{LIB_DIR}", "-ltony_b200_kernels",
                   "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN"]
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                raise RuntimeError(f"selftest build failed:\n{res.stdout}\n{res.stderr}")
    return lib


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--selftest", action="store_true")
    ap.add_argument("--verbose", action="store_true", help="pass -Xptxas=-v (registers, spills, smem)")
    args = ap.parse_args(argv)
    lib = build(force=args.force, selftest=args.selftest, verbose=args.verbose)
    print(lib)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
