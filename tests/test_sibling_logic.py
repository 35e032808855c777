"""The sibling-speculation bookkeeping of the resident-server host path (csrc/lamm_sibling.h) on a CPU: which followers
get their jobs posted behind a weight's job, what counts as a hit, what a misprediction leaves behind, and that a link
learned before a weight may have gone away is never followed.  The GPU side of the same feature is
tests/test_gpu_parity.py::test_operator_sibling_speculation."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sibling_logic(tmp_path):
    exe = str(tmp_path / "sibling_logic_test")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-fsanitize=address,undefined",
                           "-I", os.path.join(ROOT, "la-llama.cpp_b200", "csrc"),
                           os.path.join(ROOT, "tests", "sibling_logic_test.cpp"), "-o", exe])
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "sibling logic: ok" in r.stdout
