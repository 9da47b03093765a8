#!/usr/bin/env python
"""GAR latency of the UNMODIFIED reference rules (baseline/_ref/pytorch_impl/libs/aggregators) on this
box, with the synchronisation the reference's own gar_bench.py lacks. Prints one JSON line per point.

    python baseline/ref_gar_bench.py [--device cuda] [--n 8 --f 1 --d 23528522] [--gars krum,median,bulyan]
"""
import argparse
import json
import pathlib
import sys
import time

HERE = pathlib.Path(__file__).resolve().parent
LIBS = HERE / "_ref" / "pytorch_impl" / "libs"
sys.path = [str(LIBS)] + [p for p in sys.path if p and pathlib.Path(p).resolve() not in (HERE.parent, HERE)]
import torch  # noqa: E402


def load_prebuilt_natives():
    """The reference's native modules as built UNMODIFIED by baseline/build_ref_natives.py (forced includes from the compiler
    side, NATIVE_STD=c++17): a stand-in `native` package whose krum / bulyan / median / brute attributes are the reference's
    pybind11 modules, so that its registry offers native-krum / native-bulyan / native-median / native-brute (CPU and CUDA
    paths, the kernels of libs/native/py_*/ *.cu). None when they were not built."""
    import ctypes
    import importlib.util
    import types
    base = LIBS / "native"
    pool = base / "so_threadpool" / "so_threadpool.so"
    if not pool.exists():
        return None
    ctypes.CDLL(str(pool), mode=ctypes.RTLD_GLOBAL)
    package = types.ModuleType("native")
    loaded = []
    for name in ("krum", "bulyan", "median", "brute"):
        so = base / ("py_" + name) / ("py_%s.so" % name)
        if not so.exists():
            continue
        try:
            spec = importlib.util.spec_from_file_location("py_" + name, so)
            module = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(module)
            setattr(package, name, module)
            loaded.append(name)
        except Exception as err:       # e.g. built against another torch
            print(json.dumps({"impl": "reference", "native_load_error": "%s: %s" % (name, err)}), flush=True)
    return package if loaded else None


NATIVE = load_prebuilt_natives()
sys.modules["native"] = NATIVE      # None -> ImportError inside the reference -> pure-torch rules only


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--n", type=int, default=8)
    ap.add_argument("--f", type=int, default=1)
    ap.add_argument("--d", type=int, default=23528522)
    ap.add_argument("--gars", default="average,median,krum,bulyan,brute,aksel,native-median,native-krum,native-bulyan,native-brute")
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    stdout = sys.stdout
    import aggregators  # the reference's registry (wraps stdout on import)
    sys.stdout = stdout
    assert str(LIBS) in aggregators.__file__
    dev = torch.device(args.device)
    torch.manual_seed(1234)
    grads = [torch.randn(args.d, device=dev) for _ in range(args.n)]
    print(json.dumps({"impl": "reference", "natives": sorted(k for k in ("krum", "bulyan", "median", "brute") if NATIVE is not None and hasattr(NATIVE, k))}),
          flush=True)
    for name in args.gars.split(","):
        rule = aggregators.gars.get(name)
        if rule is None:
            continue
        if rule.check(gradients=grads, f=args.f) is not None:
            continue
        times = []
        for it in range(args.reps + 1):
            if dev.type == "cuda":
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            rule(gradients=grads, f=args.f)
            if dev.type == "cuda":
                torch.cuda.synchronize()
            if it:
                times.append((time.perf_counter() - t0) * 1e3)
        times.sort()
        print(json.dumps({"impl": "reference", "gar": name, "n": args.n, "f": args.f, "d": args.d, "device": str(dev),
                          "ms": times[len(times) // 2]}), flush=True)


if __name__ == "__main__":
    main()
