"""CPU: the exact integer-scan evaluation of sequential fp64 sums (velociraptor-stf_b200/csrc/seqsum.h, used by the 6DFOF
dispersion kernels) against the plain left-to-right loop, bit for bit, on adversarial sequences.  The header is shared
between the CUDA kernels and this host harness."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_seqsum_bit_identical(tmp_path):
    exe = str(tmp_path / "seqsum_test")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "velociraptor-stf_b200", "csrc"),
                    os.path.join(ROOT, "tests", "host", "seqsum_test.cpp"), "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    sys.stdout.write(r.stdout)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "all sequences bit-identical" in r.stdout
