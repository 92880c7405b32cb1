"""ASan + UBSan run of the native host runtime through csrc/host_selftest (the reference advertises
sanitizer builds but never enables them)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_host_runtime_under_asan_ubsan(tmp_path):
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "sanitize_host.sh"), str(tmp_path)], capture_output=True,
                       text=True, timeout=600)
    if r.returncode != 0 and ("cannot find -lasan" in r.stderr + r.stdout or "cannot find -ltsan" in r.stderr + r.stdout):
        pytest.skip("sanitizer runtime libraries not installed")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-6000:]
    assert "all checks passed" in r.stdout
    assert "ERROR: AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr
    # second half of the script: the TCP transport alone under ThreadSanitizer
    assert "tcp selftest: all checks passed" in r.stdout
    assert "WARNING: ThreadSanitizer" not in r.stderr
