"""The reference's OWN tests, unmodified, with hyteg::P1ConstantLaplaceOperator swapped for the bridge class of
integration/hyteg_bridge/ (oracle/build_ref.py compiles /root/reference/tests/hyteg/... with a forced include; nothing of the test source
is edited or copied).  A binary exits 0 only if every WALBERLA_CHECK of the reference test held, e.g. the per-cycle residual reduction
< 3.2e-2 of tests/hyteg/convergence/P1GMG3DConvergenceTest.cpp:151.

Also built this way: convergence/P1GSConvergenceTest.cpp and convergence/P1JacobiConvergenceTest.cpp, which loop over 2D and 3D meshes.

  <Test>_ref    the test as the reference builds it (its CPU operator)                      -- the expected log
  <Test>_emul   bridge + tests/emul/libhyteg_b200_emul.so (kernels run on the host)          -- CPU suite
  <Test>_b200   bridge + hyteg_b200/libhyteg_b200.so (the CUDA kernels)                      -- `-m gpu`
"""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
TESTS = ["P1GMG3DConvergenceTest", "P1CG3DConvergenceTest", "P1LaplaceOperator3DTest", "P1GSConvergenceTest", "P1JacobiConvergenceTest", "P1ElementwiseOperatorTest"]
NUM = r"[-+]?\d\.?\d*(?:e[-+]?\d+)?"


def run(name, tag, tmp_path):
    exe = os.path.join(REF, "cpu_only" if tag == "emul" else "", name + "_" + tag)
    if not os.path.exists(exe):
        pytest.skip("%s not built (oracle/build_ref.py needs the reference tree)" % os.path.basename(exe))
    # the tests write their partitioning VTK files to ../../output
    cwd = tmp_path / "a" / "b"
    cwd.mkdir(parents=True)
    (tmp_path / "output").mkdir()
    r = subprocess.run([exe], cwd=cwd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def gmg_history(log):
    rows = re.findall(r"After\s+(\d+) VCycles: Residual: (%s) \| convRate: (%s) \| Error L2: (%s)" % (NUM, NUM, NUM), log)
    assert len(rows) == 4, log[-2000:]
    return [(float(r), float(c), float(e)) for _, r, c, e in rows]


def compare(name, log, ref):
    if name == "P1GMG3DConvergenceTest":
        h, hr = gmg_history(log), gmg_history(ref)
        for (res, rate, err), (_, rrate, rerr) in zip(h, hr):
            assert rate < 3.2e-2
            # multicolour instead of lexicographic Gauss-Seidel: another iteration with the same fixed point;
            # the squared-residual reduction per cycle stays within a factor 2.5 of the reference's
            assert rate < 2.5 * rrate and rate > rrate / 2.5
        # the discrete solution both converge to: the error against the analytic solution agrees to 4 digits after 4 cycles
        assert abs(h[-1][2] - hr[-1][2]) < 1e-3 * hr[-1][2]
    elif name == "P1CG3DConvergenceTest":
        a = re.findall(r"Residual L2 on level (\d): (%s) \| Error L2: (%s)" % (NUM, NUM), log)
        b = re.findall(r"Residual L2 on level (\d): (%s) \| Error L2: (%s)" % (NUM, NUM), ref)
        assert len(a) == 2 and len(b) == 2
        for (la, ra, ea), (lb, rb, eb) in zip(a, b):
            assert la == lb and float(ra) < 1e-16
            assert abs(float(ea) - float(eb)) < 1e-6 * float(eb)  # CG solves the same system to round-off: same discretisation error
    elif name == "P1ElementwiseOperatorTest":
        # the reference's on-the-fly elementwise operator A against the constant-stencil operator B (here: the device) on a 2D and a 3D mesh
        a = [float(x) for x in re.findall(r"\|\|\(A - B\)u\|\|_inf = (%s)" % NUM, log)]
        assert len(a) >= 2 and max(a) < 1e-14, a
    elif name in ("P1GSConvergenceTest", "P1JacobiConvergenceTest"):
        # 20 sweeps on 2D meshes (the bridge keeps the reference's operator there) and on 3D/tet_1el, 3D/regular_octahedron_8el (device);
        # the test itself checks that the residual falls in every sweep
        a = [float(x) for x in re.findall(r"discrete L2 error = (%s)" % NUM, log)]
        b = [float(x) for x in re.findall(r"discrete L2 error = (%s)" % NUM, ref)]
        assert len(a) == len(b) >= 3
        for ea, eb in zip(a, b):
            if name == "P1JacobiConvergenceTest":
                assert abs(ea - eb) <= 1e-5 * eb  # the same iteration
            else:
                assert abs(ea - eb) <= 0.25 * eb  # multicolour against lexicographic ordering after 20 sweeps (2D cases: identical)
    else:
        a = [float(x) for x in re.findall(r"L2 error: (%s)" % NUM, log)]
        b = [float(x) for x in re.findall(r"L2 error: (%s)" % NUM, ref)]
        assert len(a) == len(b) == 28  # 7 mesh / level cases x 4 functions
        assert max(a) < 1e-12


@pytest.mark.parametrize("name", TESTS)
def test_reference_test_through_bridge_emulated(name, tmp_path):
    ref = run(name, "ref", tmp_path / "ref")
    log = run(name, "emul", tmp_path / "dev")
    compare(name, log, ref)


@pytest.mark.gpu
@pytest.mark.parametrize("name", TESTS)
def test_reference_test_through_bridge_gpu(name, tmp_path):
    ref = run(name, "ref", tmp_path / "ref")
    log = run(name, "b200", tmp_path / "dev")
    compare(name, log, ref)
