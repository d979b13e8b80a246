"""In-tree build of the two native libraries.

* ``lib/libdtf_b200_cuda.so`` — hand-written sm_100a kernels + peer-memory runtime (nvcc,
  ``-gencode arch=compute_100a,code=sm_100a -lineinfo``). No PyTorch headers: the ABI is plain C
  (``include/dtf_b200.h``) and is bound with ctypes, so a rebuild takes seconds.
* ``lib/libdtf_b200_host.so`` — host C++: crc32c, TFRecord/event writer, LevelDB-format SSTable,
  TF1 tensor-bundle checkpoint reader/writer (``include/dtf_b200_io.h``).

The reference has no native code of its own; these libraries replace what it
gets from inside the TensorFlow wheel.
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
LIB_DIR = Path(__file__).resolve().parent / "lib"
CUDA_LIB = LIB_DIR / "libdtf_b200_cuda.so"
HOST_LIB = LIB_DIR / "libdtf_b200_host.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=default",
    "--expt-relaxed-constexpr",
]
# --use_fast_math only where it is harmless (GEMM / conv epilogues). The optimizer rules and the
# losses are built with IEEE division, sqrt and expf and without flush-to-zero: they decide what goes
# into a checkpoint, they are bandwidth- or latency-bound, and approximate math buys them nothing.
PRECISE_SOURCES = {"apply.cu", "mailbox.cu", "sync_step.cu", "softmax_xent.cu", "head_fused.cu"}


def _precise() -> set:
    # DTF_B200_FAST_MATH_ALL=1 builds everything with --use_fast_math (A/B measurements only)
    return set() if os.environ.get("DTF_B200_FAST_MATH_ALL") == "1" else PRECISE_SOURCES


def nvcc_flags_for(src: Path) -> list:
    if src.name in _precise():
        extra = os.environ.get("DTF_B200_PRECISE_EXTRA", "")
        return NVCC_FLAGS + (extra.split() if extra else [])
    return NVCC_FLAGS + ["--use_fast_math"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def cuda_sources() -> list[Path]:
    return sorted((CSRC / "kernels").glob("*.cu")) + sorted((CSRC / "runtime").glob("*.cu"))


def host_sources() -> list[Path]:
    return sorted((CSRC / "io").glob("*.cc"))


def _digest(paths: list[Path], extra: str) -> str:
    h = hashlib.sha256(extra.encode())
    deps = list(paths) + sorted((CSRC / "kernels").glob("*.cuh")) + sorted((ROOT / "include").glob("*.h"))
    for p in deps:
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _stale(lib: Path, digest: str) -> bool:
    stamp = lib.with_suffix(".stamp")
    return not lib.exists() or not stamp.exists() or stamp.read_text().strip() != digest


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    srcs = cuda_sources()
    digest = _digest(srcs, " ".join(NVCC_FLAGS) + "|precise:" + ",".join(sorted(_precise())) + os.environ.get("DTF_B200_PRECISE_EXTRA", ""))
    if not force and not _stale(CUDA_LIB, digest):
        return CUDA_LIB
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    obj_dir = ROOT / "build" / "cuda"
    obj_dir.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()
    procs = []
    objs = []
    for s in srcs:
        o = obj_dir / (s.stem + ".o")
        objs.append(o)
        cmd = [nvcc, *nvcc_flags_for(s), "-c", str(s), "-o", str(o)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- nvcc {s.name} ---\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed (see stderr)")
    cmd = [nvcc, "-shared", "-o", str(CUDA_LIB), *map(str, objs), "-cudart", "static"]
    subprocess.run(cmd, check=True)
    CUDA_LIB.with_suffix(".stamp").write_text(digest)
    return CUDA_LIB


def build_host(force: bool = False) -> Path:
    srcs = host_sources()
    if not srcs:
        return HOST_LIB
    flags = ["-O2", "-std=c++17", "-fPIC", "-shared", "-msse4.2", "-Wall"]
    digest = _digest(srcs, " ".join(flags))
    if not force and not _stale(HOST_LIB, digest):
        return HOST_LIB
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    cmd = ["g++", *flags, "-o", str(HOST_LIB), *map(str, srcs)]
    subprocess.run(cmd, check=True)
    HOST_LIB.with_suffix(".stamp").write_text(digest)
    return HOST_LIB


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_host(force)
    build_cuda(force, verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print("built", CUDA_LIB, HOST_LIB)
