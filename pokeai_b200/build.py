"""Build the sm_100a shared library in-tree (pokeai_b200/_lib/libpokeai_b200.so).

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the GPU box with
the repo snapshot.  ``python -m pokeai_b200.build`` or ``__graft_entry__.build()``.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB_DIR = PKG / "_lib"
LIB_PATH = LIB_DIR / "libpokeai_b200.so"
STAMP = LIB_DIR / "libpokeai_b200.stamp"
PACK_SRC = CSRC / "pkpack.c"
PACK_PATH = LIB_DIR / "_pkpack.so"          # CPython extension: the lock-step driver's request packer (host side only)
PACK_STAMP = LIB_DIR / "_pkpack.stamp"

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]


def _sources():
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.inc"))
                  + list((PKG.parent / "include").glob("*.h")))


def _digest() -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for p in _sources():
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def is_current() -> bool:
    return LIB_PATH.exists() and STAMP.exists() and STAMP.read_text().strip() == _digest()


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and is_current():
        return LIB_PATH
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    tmp = LIB_DIR / "libpokeai_b200.so.tmp"
    cmd = [find_nvcc(), *NVCC_FLAGS, "-o", str(tmp), str(CSRC / "pk_lib.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd), file=sys.stderr)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr, file=sys.stderr)
    os.replace(tmp, LIB_PATH)
    STAMP.write_text(_digest() + "\n")
    return LIB_PATH


def build_packer(force: bool = False):
    """gcc-compile csrc/pkpack.c into _lib/_pkpack.so for THIS interpreter.  Returns the path, or None when the
    Python headers or gcc are missing (protocol.pack_request_into then stays on its Python implementation)."""
    import sysconfig
    inc = sysconfig.get_paths().get("include")
    gcc = shutil.which("gcc") or shutil.which("cc")
    if not inc or not (Path(inc) / "Python.h").exists() or gcc is None:
        return None
    digest = hashlib.sha256(PACK_SRC.read_bytes() + sys.version.encode()).hexdigest()
    if not force and PACK_PATH.exists() and PACK_STAMP.exists() and PACK_STAMP.read_text().strip() == digest:
        return PACK_PATH
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    tmp = LIB_DIR / "_pkpack.so.tmp"
    res = subprocess.run([gcc, "-O2", "-shared", "-fPIC", "-I", inc, "-o", str(tmp), str(PACK_SRC)],
                         capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("gcc failed on pkpack.c:\n" + res.stdout + res.stderr)
    os.replace(tmp, PACK_PATH)
    PACK_STAMP.write_text(digest + "\n")
    return PACK_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_packer(force="--force" in sys.argv))
