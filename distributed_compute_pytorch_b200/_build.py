"""In-tree build of the two native libraries.

* ``_lib/libdcp_host.so``  - g++: TCP store, CPU socket mesh, FD passing, reducer bookkeeping.
* ``_lib/libdcp_cuda.so``  - nvcc ``-gencode arch=compute_100a,code=sm_100a``: symmetric memory,
  P2P/NVLS collectives, compute kernels.  Links cudart statically and resolves the driver API at
  run time, so it loads (and cross-compiles) on a CPU-only box.

Both are plain C-ABI shared objects loaded with ctypes (see ``_native.py``); there is no torch
extension to JIT, so the built ``.so`` files travel with the source tree.
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
LIB = Path(__file__).resolve().parent / "_lib"
OBJ = ROOT / "build" / "obj"

HOST_SOURCES = ["host/store.cpp", "host/mesh.cpp", "ddp/reducer.cpp"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr",
]
# --use_fast_math (approximate expf / logf / division / rsqrt, flush-to-zero) only where the result feeds bf16 / tf32 tensor-core
# operands or a dropout mask.  The numerically sensitive kernels - losses, BatchNorm statistics, every optimizer update (stand-alone,
# fused with the collective, and the data-parallel tail of the flagship step) - are compiled with IEEE arithmetic so that long runs
# track torch.optim (tests/test_optimizer_drift_gpu.py).
PRECISE_SOURCES = {"ops/loss.cu", "ops/batchnorm.cu", "ops/optim.cu", "ops/convnet_dp.cu", "coll/fused_opt.cu"}


def _flags_for(src: str) -> list[str]:
    return NVCC_FLAGS + ([] if src in PRECISE_SOURCES else ["--use_fast_math"])


def _cuda_sources() -> list[str]:
    out = []
    for sub in ("symm", "coll", "ops", "ddp"):
        for p in sorted((CSRC / sub).glob("*.cu")):
            out.append(str(p.relative_to(CSRC)))
    return out


def _newer(target: Path, deps: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.exists() and d.stat().st_mtime > t for d in deps)


def _headers() -> list[Path]:
    return list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.cuh")) + list((ROOT / "include").glob("*.h"))


def _run(cmd: list[str], log: Path | None = None) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if log is not None:
        log.parent.mkdir(parents=True, exist_ok=True)
        log.write_text(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError(f"build failed: {' '.join(cmd)}")


def build_host(force: bool = False) -> Path:
    LIB.mkdir(parents=True, exist_ok=True)
    target = LIB / "libdcp_host.so"
    srcs = [CSRC / s for s in HOST_SOURCES]
    if force or _newer(target, srcs + _headers()):
        cxx = os.environ.get("CXX", "g++")
        _run([cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-Wall", "-I", str(ROOT / "include"),
              "-o", str(target)] + [str(s) for s in srcs])
    return target


def nvcc_path() -> str | None:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return cand if os.path.exists(cand) else None


def build_cuda(force: bool = False, verbose: bool = False) -> Path | None:
    """Compile every .cu under csrc/ for sm_100a and link libdcp_cuda.so.  Returns None without nvcc."""
    nvcc = nvcc_path()
    srcs = _cuda_sources()
    if nvcc is None or not srcs:
        return None
    LIB.mkdir(parents=True, exist_ok=True)
    OBJ.mkdir(parents=True, exist_ok=True)
    target = LIB / "libdcp_cuda.so"
    hdrs = _headers()
    objs: list[Path] = []
    jobs = []
    for s in srcs:
        src = CSRC / s
        obj = OBJ / (s.replace("/", "_") + ".o")
        objs.append(obj)
        if force or _newer(obj, [src] + hdrs):
            cmd = [nvcc] + _flags_for(s) + os.environ.get("DCP_NVCC_EXTRA", "").split() + ["-I", str(ROOT / "include"), "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
            jobs.append((cmd, OBJ / (obj.name + ".log")))
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(lambda j: _run(*j), jobs))
        if verbose:
            for _, log in jobs:
                sys.stderr.write(log.read_text())
    if force or jobs or _newer(target, objs):
        _run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static", "-o", str(target)]
             + [str(o) for o in objs] + ["-ldl", "-lpthread", "-lrt"])
    return target


def build_all(force: bool = False, verbose: bool = False) -> dict:
    return {"host": build_host(force), "cuda": build_cuda(force, verbose)}


if __name__ == "__main__":
    out = build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print({k: str(v) for k, v in out.items()})
