"""nfb::advance() (nerfacc_b200/csrc/grid.cuh) — the closed-form evaluation of the reference's
marching recurrence — compiled for the HOST and checked against the serial loop it replaces
(reference csrc/grid.cu:157-161, 199-203, 208-262) on random (t, dt, target, cap) cases that
include binade crossings, exact ties, negative / zero starts and step caps.  No GPU needed."""
import os
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"),
                    reason="needs nvcc (host compile of the .cuh)")
def test_advance_matches_serial_loop(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    exe = str(tmp_path / "advance_check")
    cmd = [nvcc, "-O2", "-std=c++17", "--expt-extended-lambda", "--expt-relaxed-constexpr",
           "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-ffp-contract=off",
           "-I", os.path.join(HERE, "..", "include"), "-o", exe,
           os.path.join(HERE, "host", "advance_check.cu")]
    subprocess.run(cmd, check=True, capture_output=True)
    r = subprocess.run([exe, "400000"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("ok 400000")
