"""ASan + UBSan build of the host runtime (ring, producers, generators, JSON scanner)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_host_library_under_asan_ubsan(tmp_path):
    exe = tmp_path / "ring_test"
    srcs = [os.path.join(ROOT, "csrc", "host", f) for f in ("ring.cpp", "producer.cpp", "json_records.cpp", "ring_test.cpp")]
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-pthread",
           "-I", os.path.join(ROOT, "include"), *srcs, "-o", str(exe)]
    b = subprocess.run(cmd, capture_output=True, text=True)
    if b.returncode != 0 and "asan" in (b.stderr or "").lower():
        pytest.skip("sanitizer runtime not installed")
    assert b.returncode == 0, b.stderr[-2000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120,
                       env=dict(os.environ, ASAN_OPTIONS="detect_leaks=1", UBSAN_OPTIONS="halt_on_error=1"))
    assert r.returncode == 0 and "ring_test ok" in r.stdout, r.stdout[-1000:] + r.stderr[-3000:]
