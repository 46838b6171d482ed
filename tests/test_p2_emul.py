"""CPU: the P2 host logic (edge-DoF layout, interface groups, shell-map classification) and the arithmetic of the P2 kernels through
the TEST-ONLY emulation build, against outputs of the reference itself (tests/golden/ref_p2.npz)."""
import pytest

import p2_cases as pc


def test_shell_map(emul_lib):
    pc.case_shell_map(emul_lib)


@pytest.mark.parametrize("kind,L", [("shell", 2), ("cube", 2), ("cube", 3)])
def test_divkgrad_matches_reference(emul_lib, kind, L):
    pc.case_divkgrad(emul_lib, kind, L)


def test_divkgrad_p2_coefficient(emul_lib):
    pc.case_divkgrad(emul_lib, "shell", 2, mode="p2")


@pytest.mark.parametrize("i,j", [(0, 0), (0, 1), (2, 0)])
def test_epsilon_blocks(emul_lib, i, j):
    pc.case_epsilon(emul_lib, i, j)


def test_p2_blas(emul_lib):
    pc.case_blas(emul_lib)


@pytest.mark.parametrize("kind,Lc", [("cube", 2), ("shell", 1)])
def test_quadratic_transfer(emul_lib, kind, Lc):
    pc.case_transfer(emul_lib, kind, Lc)


@pytest.mark.parametrize("comp", [0, 2])
def test_div_blocks(emul_lib, comp):
    pc.case_div_blocks(emul_lib, comp)


@pytest.mark.parametrize("fused", [True, False])
def test_stokes_composite(emul_lib, fused):
    pc.case_stokes(emul_lib, fused)


def test_chebyshev_gmg(emul_lib):
    pc.case_chebyshev_gmg(emul_lib, cycles=2)
