"""In-tree native build: two shared objects with a plain C ABI, loaded through ctypes.

* `garfield_b200/lib/libgarfield_b200_cpu.so`  — C++17 host runtime (thread pool, combinations,
  CPU GARs, multi-replica register); g++ only.
* `garfield_b200/lib/libgarfield_b200_cuda.so` — hand-written sm_100a kernels + CUDA host
  runtime (peer-memory arena, IPC); nvcc `-gencode arch=compute_100a,code=sm_100a -lineinfo`.

Counterpart of the reference's lazy `cpp_extension.load` build
(/root/reference/pytorch_impl/libs/native/__init__.py:19-156), with three differences: the
arch is pinned to sm_100a, the objects are built *in tree* (so they travel with a repo
snapshot) and nothing includes torch headers, which keeps a full rebuild under a minute.

Env knobs (same spirit as NATIVE_OPT/NATIVE_QUIET): GARFIELD_DEBUG=1 → -O0 -g -G,
GARFIELD_QUIET=1 → no build chatter, GARFIELD_NO_BUILD=1 → never build, only load.
"""

from __future__ import annotations

import concurrent.futures
import ctypes
import hashlib
import os
import pathlib
import shutil
import subprocess
import sys
import threading

ROOT = pathlib.Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIBDIR = ROOT / "garfield_b200" / "lib"
OBJDIR = ROOT / "build"

CPU_LIB = LIBDIR / "libgarfield_b200_cpu.so"
CUDA_LIB = LIBDIR / "libgarfield_b200_cuda.so"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]

_lock = threading.Lock()
_handles = {}


def _say(*args):
    if not os.environ.get("GARFIELD_QUIET"):
        print("[garfield_b200 build]", *args, file=sys.stderr, flush=True)


def _debug():
    return bool(os.environ.get("GARFIELD_DEBUG"))


def nvcc_path():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def _headers():
    found = list(INCLUDE.glob("*.h")) + list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.hpp")) + list(CSRC.rglob("*.h"))
    return found


def _stamp(paths, flags):
    h = hashlib.sha1()
    for p in sorted(map(str, paths)):
        h.update(os.path.basename(p).encode())
        with open(p, "rb") as fh:  # content hash: survives copies that reset mtimes (gpurun snapshots)
            h.update(fh.read())
    h.update(" ".join(flags).encode())
    return h.hexdigest()


def _compile_one(cmd, obj, stamp_file, stamp):
    obj.parent.mkdir(parents=True, exist_ok=True)
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("native compile failed:\n$ %s\n%s\n%s" % (" ".join(cmd), proc.stdout, proc.stderr))
    if proc.stderr.strip() and not os.environ.get("GARFIELD_QUIET"):
        sys.stderr.write(proc.stderr)
    stamp_file.write_text(stamp)
    return obj


def _build(kind, sources, compiler, cflags, ldflags, out):
    """Compile stale translation units in parallel, then link when anything changed."""
    headers = _headers()
    jobs, objs = [], []
    for src in sources:
        obj = OBJDIR / kind / (src.stem + ".o")
        stamp_file = obj.with_suffix(".stamp")
        stamp = _stamp([src] + headers, cflags)
        objs.append(obj)
        if obj.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
            continue
        cmd = [compiler] + cflags + ["-I", str(INCLUDE), "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
        jobs.append((cmd, obj, stamp_file, stamp))
    if jobs:
        _say("compiling %d %s source(s): %s" % (len(jobs), kind, ", ".join(j[1].stem for j in jobs)))
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as pool:
            list(pool.map(lambda job: _compile_one(*job), jobs))
    if jobs or not out.exists():
        out.parent.mkdir(parents=True, exist_ok=True)
        tmp = out.with_suffix(".so.tmp%d" % os.getpid())
        cmd = [compiler] + ldflags + [str(o) for o in objs] + ["-o", str(tmp)]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError("native link failed:\n$ %s\n%s\n%s" % (" ".join(cmd), proc.stdout, proc.stderr))
        os.replace(tmp, out)
        _say("linked", out.relative_to(ROOT))
    return out


def build_cpu(force=False):
    """Build the host library (g++). Returns its path."""
    sources = sorted((CSRC / "cpu").glob("*.cpp"))
    opt = ["-O0", "-g"] if _debug() else ["-O3", "-DNDEBUG"]
    cflags = ["-std=c++17", "-fPIC", "-pthread", "-Wall", "-Wextra", "-fvisibility=hidden", "-march=x86-64-v2"] + opt
    ldflags = ["-shared", "-pthread"]
    if force:
        shutil.rmtree(OBJDIR / "cpu", ignore_errors=True)
    return _build("cpu", sources, os.environ.get("CXX", "g++"), cflags, ldflags, CPU_LIB)


def build_cuda(force=False):
    """Cross-compile every .cu under csrc/cuda for sm_100a (works without a GPU)."""
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found; cannot build the sm_100a library")
    sources = sorted((CSRC / "cuda").glob("*.cu"))
    opt = ["-O0", "-g", "-G"] if _debug() else ["-O3", "-DNDEBUG"]
    cflags = ARCH_FLAGS + ["-std=c++17", "-lineinfo", "--expt-relaxed-constexpr", "--extended-lambda", "-Xcompiler", "-fPIC,-fvisibility=hidden",
                           "-Xptxas", "-v" if os.environ.get("GARFIELD_PTXAS_V") else "-O3"] + opt
    ldflags = ARCH_FLAGS + ["-shared", "-Xcompiler", "-fPIC", "-cudart", "shared"]
    if force:
        shutil.rmtree(OBJDIR / "cuda", ignore_errors=True)
    return _build("cuda", sources, nvcc, cflags, ldflags, CUDA_LIB)


def build_all(force=False):
    return build_cpu(force), build_cuda(force)


def _locked_build(builder):
    """Run `builder` under an inter-process file lock: under torchrun every rank imports the package at
    the same time and only one of them may compile/link (the others then find everything up to date)."""
    import fcntl
    LIBDIR.mkdir(parents=True, exist_ok=True)
    with open(LIBDIR / ".build.lock", "w") as lock_file:
        fcntl.flock(lock_file, fcntl.LOCK_EX)
        try:
            return builder()
        finally:
            fcntl.flock(lock_file, fcntl.LOCK_UN)


def _load(key, path, builder):
    with _lock:
        if key in _handles:
            return _handles[key]
        if not os.environ.get("GARFIELD_NO_BUILD"):
            try:
                _locked_build(builder)
            except Exception as err:
                if not path.exists():
                    raise
                _say("rebuild failed (%s); using existing %s" % (err, path.name))
        handle = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
        _handles[key] = handle
        return handle


def load_cpu():
    """ctypes handle of the host library (built on first use)."""
    return _load("cpu", CPU_LIB, build_cpu)


def load_cuda():
    """ctypes handle of the sm_100a library (built on first use). Loading needs libcudart, which
    the CUDA toolkit in this image provides even on GPU-less machines."""
    return _load("cuda", CUDA_LIB, build_cuda)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
    print(CPU_LIB, CUDA_LIB)
