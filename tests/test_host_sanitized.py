"""Host C++ library under AddressSanitizer + UBSan (the reference has no sanitizer
set-up; the native I/O code that replaces TF's C++ EventsWriter / Saver gets one here)."""
import glob
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_host_io_selftest_under_asan_ubsan(tmp_path):
    srcs = sorted(glob.glob(os.path.join(ROOT, "csrc", "io", "*.cc")))
    srcs.append(os.path.join(ROOT, "csrc", "io", "tests", "selftest.cc"))
    exe = tmp_path / "dtf_io_selftest"
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fno-omit-frame-pointer", "-msse4.2",
           "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
           "-I", os.path.join(ROOT, "include"), *srcs, "-o", str(exe)]
    build = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if build.returncode != 0 and "cannot find -lasan" in build.stderr + build.stdout:
        pytest.skip("libasan is not installed")
    assert build.returncode == 0, build.stderr[-4000:]
    scratch = tmp_path / "scratch"
    scratch.mkdir()
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")
    run = subprocess.run([str(exe), str(scratch)], capture_output=True, text=True, timeout=300, env=env)
    assert run.returncode == 0, (run.stdout + run.stderr)[-6000:]
    assert "all checks passed" in run.stdout
