"""In-tree build of ``libb200dp.so`` (sm_100a only).

The reference ships no native code and no build system; this
framework owns its kernels, so it owns the build too.  Every translation unit
under ``csrc/`` is compiled with

    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17

into ``build/*.o`` and linked into
``cs744_pytorch_distributed_tutorial_b200/lib/libb200dp.so``.  nvcc
cross-compiles without a GPU, so this also is the "does it build" check that
``__graft_entry__.build()`` runs on the CPU-only dev box.  The ``.so`` is
git-ignored but travels with the working tree to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent.parent
REPO_ROOT = PKG_DIR.parent
CSRC = REPO_ROOT / "csrc"
INCLUDE = REPO_ROOT / "include"
BUILD_DIR = REPO_ROOT / "build"
LIB_PATH = PKG_DIR / "lib" / "libb200dp.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; libb200dp.so cannot be built")


def sources() -> list[Path]:
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cpp")))


def _fingerprint() -> str:
    h = hashlib.sha256()
    for f in sorted(list(CSRC.glob("*")) + list(INCLUDE.glob("*.h"))):
        if f.is_file():
            h.update(f.name.encode())
            h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_fresh() -> bool:
    stamp = LIB_PATH.with_suffix(".stamp")
    return LIB_PATH.exists() and stamp.exists() and stamp.read_text().strip() == _fingerprint()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile and link the library if any source changed.  Returns its path."""
    if not force and is_fresh():
        return LIB_PATH
    nvcc = _nvcc()
    BUILD_DIR.mkdir(exist_ok=True)
    LIB_PATH.parent.mkdir(exist_ok=True)
    srcs = sources()
    if not srcs:
        raise RuntimeError(f"no sources under {CSRC}")

    def compile_one(src: Path) -> Path:
        obj = BUILD_DIR / (src.name + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-c", str(src), "-o", str(obj)]
        if src.suffix == ".cpp":
            cmd.insert(1, "-x")
            cmd.insert(2, "cu")
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}:\n{r.stdout}\n{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    cmd = [nvcc, "-shared", "-o", str(LIB_PATH), *map(str, objs), "-lpthread", "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    LIB_PATH.with_suffix(".stamp").write_text(_fingerprint())
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
