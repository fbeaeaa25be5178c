"""Build the native sm_100a library in-tree (dedloc_b200/lib/libdedloc_b200.so).

One `nvcc -shared` per translation unit list; no PyTorch headers are involved, so a full rebuild is
seconds. nvcc cross-compiles without a GPU, which is what `__graft_entry__.build()` relies on.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIB_DIR = Path(__file__).resolve().parent / "lib"
LIB_PATH = LIB_DIR / "libdedloc_b200.so"
STAMP = LIB_DIR / ".build_stamp"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _fingerprint() -> str:
    h = hashlib.sha256()
    for f in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))):
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_fresh() -> bool:
    return LIB_PATH.exists() and STAMP.exists() and STAMP.read_text().strip() == _fingerprint()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every .cu under csrc/ for sm_100a into one shared library. Returns its path."""
    if not force and is_fresh():
        return LIB_PATH
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    objs = []
    obj_dir = ROOT / "build" / "obj"
    obj_dir.mkdir(parents=True, exist_ok=True)
    procs = []
    log_dir = ROOT / "build" / "log"
    log_dir.mkdir(parents=True, exist_ok=True)
    for src in sources():
        obj = obj_dir / (src.stem + ".o")
        objs.append(obj)
        cmd = [_nvcc(), *NVCC_FLAGS, "-I", str(INCLUDE), "-c", str(src), "-o", str(obj)]
        log = open(log_dir / (src.stem + ".log"), "w")
        procs.append((src, subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT), log))
    failed = []
    for src, proc, log in procs:
        rc = proc.wait()
        log.close()
        text = (log_dir / (src.stem + ".log")).read_text()
        if verbose or rc != 0:
            sys.stderr.write(f"--- nvcc {src.name} (rc={rc})\n{text}\n")
        if rc != 0:
            failed.append(src.name)
    if failed:
        raise RuntimeError(f"nvcc failed for: {failed}")
    link = [_nvcc(), "-shared", "-o", str(LIB_PATH), *map(str, objs)]
    subprocess.run(link, check=True)
    STAMP.write_text(_fingerprint())
    return LIB_PATH


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
