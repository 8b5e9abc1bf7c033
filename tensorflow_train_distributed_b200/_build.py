"""In-tree native build: nvcc for the sm_100a kernel library, g++ for the host-side I/O library.

The outputs land in ``tensorflow_train_distributed_b200/_lib/`` so they travel with a repository
snapshot (a JIT cache under ~/.cache would not). Nothing here needs a GPU: nvcc cross-compiles.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO_DIR = PKG_DIR.parent
CSRC = REPO_DIR / "csrc"
LIB_DIR = PKG_DIR / "_lib"

CUDA_LIB = LIB_DIR / "libttd_b200.so"
IO_LIB = LIB_DIR / "libttd_io.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-msse4.2", "-pthread"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; set NVCC or install the CUDA toolkit")


def _sources(pattern: str, sub: str = "") -> list[Path]:
    return sorted((CSRC / sub).glob(pattern))


def _stamp(files: list[Path], flags: list[str]) -> str:
    h = hashlib.sha256()
    for f in files:
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(flags).encode())
    return h.hexdigest()


def _up_to_date(target: Path, stamp: str) -> bool:
    sf = target.with_suffix(".stamp")
    return target.exists() and sf.exists() and sf.read_text() == stamp


def _run(cmd: list[str], log: Path | None = None) -> None:
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if log is not None:
        log.write_text(proc.stdout + proc.stderr)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout + proc.stderr)
        raise RuntimeError("native build failed: " + " ".join(cmd))


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    """Compile every csrc/*.cu into one shared library for sm_100a."""
    LIB_DIR.mkdir(exist_ok=True)
    cus = _sources("*.cu")
    deps = cus + _sources("*.cuh") + sorted((REPO_DIR / "include").glob("*.h"))
    stamp = _stamp(deps, NVCC_FLAGS)
    if not force and _up_to_date(CUDA_LIB, stamp):
        return CUDA_LIB
    objs = []
    obj_dir = LIB_DIR / "obj"
    obj_dir.mkdir(exist_ok=True)
    procs = []
    for cu in cus:
        obj = obj_dir / (cu.stem + ".o")
        objs.append(obj)
        cmd = [_nvcc(), *NVCC_FLAGS, "-c", str(cu), "-o", str(obj)]
        procs.append((cu, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    ptxas_log = []
    failed = False
    for cu, p in procs:
        out, _ = p.communicate()
        ptxas_log.append(f"==== {cu.name} ====\n{out}")
        if p.returncode != 0:
            failed = True
            sys.stderr.write(out)
    (LIB_DIR / "ptxas.log").write_text("\n".join(ptxas_log))
    if failed:
        raise RuntimeError("nvcc failed; see output above")
    _run([_nvcc(), "-shared", "-o", str(CUDA_LIB), *map(str, objs), "-cudart", "shared"])
    CUDA_LIB.with_suffix(".stamp").write_text(stamp)
    if verbose:
        print("\n".join(ptxas_log))
    return CUDA_LIB


def build_io(force: bool = False) -> Path:
    """Compile the host-only C++ library (TF1 tensor-bundle / events I/O, prefetch queue)."""
    LIB_DIR.mkdir(exist_ok=True)
    ccs = _sources("*.cc", "io")
    deps = ccs + _sources("*.h", "io")
    stamp = _stamp(deps, CXX_FLAGS)
    if not force and _up_to_date(IO_LIB, stamp):
        return IO_LIB
    _run(["g++", *CXX_FLAGS, "-shared", "-o", str(IO_LIB), *map(str, ccs)])
    IO_LIB.with_suffix(".stamp").write_text(stamp)
    return IO_LIB


def build_all(force: bool = False, verbose: bool = False) -> dict[str, Path]:
    out = {"cuda": build_cuda(force=force, verbose=verbose)}
    if _sources("*.cc", "io"):
        out["io"] = build_io(force=force)
    return out


if __name__ == "__main__":
    res = build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    for k, v in res.items():
        print(f"{k}: {v}")
