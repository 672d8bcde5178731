"""AddressSanitizer + UBSan over the native host code (oracle library through the native V1 binary, the MPI shim
through its self-test and the reference's HW1), warnings as errors: scripts/sanitize_host.sh must come back clean.
The device-side counterpart (compute-sanitizer) is closed on the GPU pool; see DESIGN.md 4b."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="no host compiler")
def test_host_code_is_clean_under_asan_ubsan(tmp_path):
    p = subprocess.run(["bash", os.path.join(ROOT, "scripts", "sanitize_host.sh"), str(tmp_path)],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    log = open(tmp_path / "sanitize_host.log").read()
    assert p.returncode == 0 and "sanitize_host: clean" in p.stdout, p.stdout + log[-3000:]
    assert "ERROR: AddressSanitizer" not in log and "runtime error:" not in log
