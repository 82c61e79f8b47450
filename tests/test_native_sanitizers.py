"""The host-side native code (C++ HDF5 writer + reader) built and run under AddressSanitizer + UndefinedBehaviorSanitizer
(SURVEY 5.2).  The CUDA side cannot be sanitized on this GPU pool (compute-sanitizer is closed); its stand-ins are the canary-band
build (tests/test_gpu_engine.py) and the bounded device-side waits."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_hdf5_writer_and_reader_are_clean_under_asan_and_ubsan(tmp_path):
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "h5_roundtrip")
    cmd = ["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-fno-omit-frame-pointer",
           os.path.join(ROOT, "csrc", "h5writer.cpp"), os.path.join(ROOT, "csrc", "h5reader.cpp"),
           os.path.join(ROOT, "tests", "native", "h5_roundtrip.cpp"), "-o", exe]
    b = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    if b.returncode != 0 and ("libasan" in b.stderr or "cannot find -lasan" in b.stderr or "libubsan" in b.stderr):
        pytest.skip("sanitizer runtimes are not installed: " + b.stderr[-200:])
    assert b.returncode == 0, b.stderr[-3000:]
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")
    r = subprocess.run([exe, str(tmp_path / "rt.h5")], capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0 and "h5 round trip ok" in r.stdout, (r.stdout[-1000:], r.stderr[-3000:])
    assert "ERROR: AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr, r.stderr[-3000:]
