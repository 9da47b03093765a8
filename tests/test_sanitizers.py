"""Race / memory checking of the C++ host runtime: the sources under csrc/cpu are rebuilt with
ThreadSanitizer and with AddressSanitizer + UBSan together with tests/native/sanitize_driver.cpp and must run clean."""
import pathlib
import shutil
import subprocess

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent


def _build_and_run(tmp_path, name, flags, env=None):
    sources = sorted(str(p) for p in (ROOT / "csrc" / "cpu").glob("*.cpp")) + [str(ROOT / "tests" / "native" / "sanitize_driver.cpp")]
    exe = tmp_path / name
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fno-omit-frame-pointer", *flags, "-I", str(ROOT / "include"), "-I", str(ROOT / "csrc" / "cpu"),
           *sources, "-o", str(exe), "-lpthread"]
    build = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if build.returncode != 0 and ("cannot find -ltsan" in build.stderr or "cannot find -lasan" in build.stderr or "unrecognized" in build.stderr):
        pytest.skip("sanitizer runtime not installed: " + build.stderr[-200:])
    assert build.returncode == 0, build.stderr[-3000:]
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600, env=env)
    return run


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_host_runtime_is_clean_under_thread_sanitizer(tmp_path):
    run = _build_and_run(tmp_path, "tsan_driver", ["-fsanitize=thread"])
    assert "ThreadSanitizer" not in run.stderr, run.stderr[-3000:]
    assert run.returncode == 0, run.stdout[-1000:] + run.stderr[-2000:]
    assert "0 failure(s)" in run.stdout


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_host_runtime_is_clean_under_address_and_ub_sanitizers(tmp_path):
    run = _build_and_run(tmp_path, "asan_driver", ["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"])
    assert "AddressSanitizer" not in run.stderr and "runtime error" not in run.stderr, run.stderr[-3000:]
    assert run.returncode == 0, run.stdout[-1000:] + run.stderr[-2000:]
    assert "0 failure(s)" in run.stdout
