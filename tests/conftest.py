import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); runs the product library")


@pytest.fixture(scope="session")
def emul_lib():
    """TEST-ONLY host emulation build of the kernel sources (tests/emul/); exercises host logic on CPU."""
    from hyteg_b200 import capi
    so = os.path.join(ROOT, "tests", "emul", "libhyteg_b200_emul.so")
    srcs = [os.path.join(ROOT, "hyteg_b200", "csrc", f) for f in os.listdir(os.path.join(ROOT, "hyteg_b200", "csrc"))]
    srcs += [os.path.join(ROOT, "include", "hyteg_b200.h"), os.path.join(ROOT, "include", "hyteg_b200", "hyteg.hpp"),
             os.path.join(ROOT, "include", "hyteg_b200", "hyteg_solvers.hpp")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call([os.path.join(ROOT, "tests", "emul", "build_emul.sh")])
    return capi.Lib(so)


@pytest.fixture(scope="session")
def cuda_lib():
    """The product library on a real device."""
    import hyteg_b200
    lib = hyteg_b200.get_lib()
    lib.check(lib.dll.hb_device_init(0))
    return lib
