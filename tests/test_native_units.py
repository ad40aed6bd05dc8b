"""Native unit checks compiled on the fly with g++ (host build of the kernel headers, -DYB_HOSTEMU)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_halton2_closed_form_equals_incremental(tmp_path):
    exe = str(tmp_path / "halton_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", os.path.join(ROOT, "tests", "native", "halton_check.cc"), "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout
    assert "0 mismatches" in r.stdout
