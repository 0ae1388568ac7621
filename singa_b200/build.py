"""In-tree build of the native library (``singa_b200/lib/libsinga_b200.so``).

Every ``csrc/*.cu`` is compiled for sm_100a only (``-gencode arch=compute_100a,code=sm_100a -lineinfo``)
and linked into one shared object with a C ABI (``include/singa_b200.h``); ``csrc/*.cc`` host-only
sources (shard file, etc.) are compiled with the same nvcc driver.  nvcc cross-compiles without a GPU, so
this also runs on the CPU-only development box.  Objects are cached by mtime under ``build/``.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
OBJ_DIR = ROOT / "build" / "obj"
LIB_DIR = Path(__file__).resolve().parent / "lib"
LIB_PATH = LIB_DIR / "libsinga_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def nvcc_path() -> str | None:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    return None


def _sources() -> list[Path]:
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cc")))


def _headers_mtime() -> float:
    hs = list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))
    return max((h.stat().st_mtime for h in hs), default=0.0)


def needs_build() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    return any(s.stat().st_mtime > t for s in _sources()) or _headers_mtime() > t


def build(verbose: bool = False, force: bool = False) -> Path:
    """Compile and link; returns the path of the shared library.  Raises on failure."""
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build libsinga_b200.so")
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    hdr_t = _headers_mtime()
    jobs = []
    objs = []
    for src in _sources():
        obj = OBJ_DIR / (src.name + ".o")
        objs.append(obj)
        if force or not obj.exists() or obj.stat().st_mtime < max(src.stat().st_mtime, hdr_t):
            cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-x", "cu", "-c", str(src), "-o", str(obj)]
            jobs.append((src, cmd))

    def run(job):
        src, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, r

    logs = []
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for src, r in ex.map(run, jobs):
            logs.append(f"== {src.name}\n{r.stderr}")
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    if jobs or not LIB_PATH.exists() or force:
        cmd = [nvcc, "-shared", "-o", str(LIB_PATH), *map(str, objs), "-lcudart", "-Xlinker", "-z,defs", "-ldl", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    (ROOT / "build" / "ptxas.log").write_text("\n".join(logs)) if logs else None
    if verbose:
        print("\n".join(logs))
        print(f"built {LIB_PATH}")
    return LIB_PATH


if __name__ == "__main__":
    build(verbose="-q" not in sys.argv, force="-f" in sys.argv)
