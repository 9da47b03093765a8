#!/usr/bin/env python
"""Build the reference's native GAR modules (py_krum / py_bulyan / py_median / py_brute + so_threadpool) UNMODIFIED from
baseline/_ref, CUDA sources included, for sm_100 — on the CPU build machine, so that the GPU box only has to load them.

The sources do not compile as they are on this toolchain (g++ 13 / torch 2.11): they use ::free and std::overflow_error
without including <cstdlib> / <stdexcept>, and their default -std=c++14 is too old for the torch headers. Both are fixed from
the OUTSIDE: NATIVE_STD=c++17 (an environment knob of the reference's own build script) and a compiler-side forced include
(`-include cstdlib -include stdexcept` through CXX / NVCC_PREPEND_FLAGS). No reference file is edited.
`torch.cuda.is_available()` is patched to True for the duration of the build only (the reference compiles its .cu files only
when a GPU is visible; nvcc cross-compiles here)."""
import os
import pathlib
import sys

HERE = pathlib.Path(__file__).resolve().parent
LIBS = HERE / "_ref" / "pytorch_impl" / "libs"


def main():
    os.environ.setdefault("NATIVE_STD", "c++17")
    os.environ.setdefault("NATIVE_OPT", "1")
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ["CXX"] = "g++ -include cstdlib -include stdexcept"
    os.environ["NVCC_PREPEND_FLAGS"] = "-include cstdlib -include stdexcept " + os.environ.get("NVCC_PREPEND_FLAGS", "")
    os.environ.setdefault("MAX_JOBS", "8")
    sys.path.insert(0, str(LIBS))
    import torch
    real = torch.cuda.is_available
    torch.cuda.is_available = lambda: True
    try:
        import native  # noqa: F401  (the reference's lazy build runs at import)
    finally:
        torch.cuda.is_available = real
    built = sorted(str(p.relative_to(LIBS)) for p in (LIBS / "native").rglob("*.so"))
    print("built:", built)
    return 0 if len(built) >= 5 else 1


if __name__ == "__main__":
    sys.exit(main())
