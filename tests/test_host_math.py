"""Builds tests/host/test_host_math.cpp with AddressSanitizer + UndefinedBehaviorSanitizer and runs it: the integer
helpers the launch code and the kernels share (division by multiply + shift, two-level fold sizing, GEMM tiling
choice) checked exhaustively on the CPU (sanitizer build of the host-side C++)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs a C++ compiler")
def test_host_math_under_asan_ubsan(tmp_path):
    exe = str(tmp_path / "test_host_math")
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
           "-I", os.path.join(ROOT, "tony_b200", "csrc"), os.path.join(ROOT, "tests", "host", "test_host_math.cpp"),
           "-o", exe]
    build = subprocess.run(cmd, capture_output=True, text=True)
    if build.returncode != 0 and "sanitize" in build.stderr:
        pytest.skip("this toolchain has no sanitizer runtime: " + build.stderr[-200:])
    assert build.returncode == 0, build.stderr[-2000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, (run.stdout + run.stderr)[-2000:]
    assert "0 failure(s)" in run.stdout
