"""The C++ host API (include/hyteg_b200/hyteg.hpp: HyTeG's class names and signatures over the C-ABI) used from a C++ program
written like a HyTeG application: compiled with g++ against libhyteg_b200.so and run on the GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_cpp_gmg_program(tmp_path):
    exe = str(tmp_path / "p1_gmg_dropin")
    libdir = os.path.join(ROOT, "hyteg_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "p1_gmg_dropin.cpp"), "-L" + libdir, "-lhyteg_b200",
                           "-Wl,-rpath," + libdir, "-o", exe])
    res = subprocess.run([exe, "5"], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "OK" in res.stdout
