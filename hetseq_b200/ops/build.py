"""In-tree build of the native libraries.

* ``hetseq_b200/lib/libhetseq_b200.so``   -- every ``csrc/*.cu`` compiled by nvcc for sm_100a only
  (``-gencode arch=compute_100a,code=sm_100a -lineinfo``), C ABI in ``include/hetseq_b200.h``.
* ``hetseq_b200/lib/libhetseq_native.so`` -- host-side C++ (``hetseq_b200/native/*.cpp``): the
  batch planner that replaces the reference's Cython extension (data/data_utils_fast.pyx).

The libraries have no dependency on torch or Python headers (ctypes binds them), so a full rebuild
takes seconds and the artefacts travel with the source tree to the GPU box.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path
from typing import List

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
NATIVE_SRC = ROOT / "hetseq_b200" / "native"
LIB_DIR = ROOT / "hetseq_b200" / "lib"
OBJ_DIR = ROOT / "build" / "obj"

CUDA_LIB = LIB_DIR / "libhetseq_b200.so"
NATIVE_LIB = LIB_DIR / "libhetseq_native.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC or put it on PATH)")


def _digest(paths: List[Path], extra: str = "") -> str:
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _run(cmd: List[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))


def cuda_sources() -> List[Path]:
    return sorted(p for p in CSRC.glob("*.cu"))


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    srcs = cuda_sources()
    headers = sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.h")) + sorted((ROOT / "include").glob("*.h"))
    stamp = LIB_DIR / ".cuda.stamp"
    want = _digest(srcs + headers, " ".join(NVCC_FLAGS))
    if not force and CUDA_LIB.exists() and stamp.exists() and stamp.read_text() == want:
        return CUDA_LIB
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()
    hdr_digest = _digest(headers, " ".join(NVCC_FLAGS))

    def compile_one(src: Path) -> Path:
        obj = OBJ_DIR / (src.stem + ".o")
        ostamp = OBJ_DIR / (src.stem + ".stamp")
        key = _digest([src], hdr_digest)
        if not force and obj.exists() and ostamp.exists() and ostamp.read_text() == key:
            return obj
        if verbose:
            print(f"[build] nvcc {src.name}", file=sys.stderr, flush=True)
        _run([nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)])
        ostamp.write_text(key)
        return obj

    with cf.ThreadPoolExecutor(max_workers=min(8, max(1, len(srcs)))) as ex:
        objs = list(ex.map(compile_one, srcs))
    # Shared cudart on purpose: the library must see the same per-thread current device / primary
    # context as the host program (PyTorch's autograd threads call into it), which a private static
    # runtime would not. libcudart.so.12 resolves to the copy already loaded by torch; the rpath is the
    # fallback for stand-alone C/C++ hosts.
    cuda_lib_dir = str(Path(nvcc).resolve().parents[1] / "lib64")
    _run([nvcc, "-shared", "-o", str(CUDA_LIB), *map(str, objs), "-cudart", "shared",
          "-Xlinker", "-rpath", "-Xlinker", cuda_lib_dir])
    stamp.write_text(want)
    return CUDA_LIB


def build_native(force: bool = False, verbose: bool = False) -> Path:
    srcs = sorted(NATIVE_SRC.glob("*.cpp"))
    stamp = LIB_DIR / ".native.stamp"
    want = _digest(srcs)
    if not force and NATIVE_LIB.exists() and stamp.exists() and stamp.read_text() == want:
        return NATIVE_LIB
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    if verbose:
        print("[build] g++ native", file=sys.stderr, flush=True)
    _run([cxx, "-O3", "-std=c++17", "-fPIC", "-shared", "-o", str(NATIVE_LIB), *map(str, srcs)])
    stamp.write_text(want)
    return NATIVE_LIB


def build_all(force: bool = False, verbose: bool = True) -> None:
    build_native(force=force, verbose=verbose)
    build_cuda(force=force, verbose=verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
    print(CUDA_LIB)
    print(NATIVE_LIB)
