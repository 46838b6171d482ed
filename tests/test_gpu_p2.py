"""-m gpu: the P2 / blending / fused elementwise path of the product CUDA library (through the C-ABI) against OUTPUTS OF THE REFERENCE
ITSELF (tests/golden/ref_p2.npz: P2ElementwiseOperator< p2_div_k_grad_blending_q4 >, < p2_epsilonvar_i_j_blending_q2 > on
meshSphericalShell(3,2,0.5,1.0) with IcosahedralShellMap, the mesh of P2P1ElementwiseStokesOperatorTest.cpp:103-155) at 1e-12 relative."""
import numpy as np
import pytest

import p2_cases as pc

pytestmark = pytest.mark.gpu


def test_shell_map(cuda_lib):
    pc.case_shell_map(cuda_lib)


@pytest.mark.parametrize("kind,L", [("shell", 2), ("shell", 3), ("cube", 2), ("cube", 3)])
def test_divkgrad_matches_reference(cuda_lib, kind, L):
    pc.case_divkgrad(cuda_lib, kind, L)


@pytest.mark.parametrize("L", [2, 3])
def test_divkgrad_p2_coefficient(cuda_lib, L):
    pc.case_divkgrad(cuda_lib, "shell", L, mode="p2")


@pytest.mark.parametrize("i,j", [(0, 0), (0, 1), (1, 2), (2, 0), (2, 2)])
def test_epsilon_blocks(cuda_lib, i, j):
    pc.case_epsilon(cuda_lib, i, j)


def test_p2_blas(cuda_lib):
    pc.case_blas(cuda_lib)


@pytest.mark.parametrize("kind,Lc", [("cube", 2), ("shell", 1)])
def test_quadratic_transfer(cuda_lib, kind, Lc):
    pc.case_transfer(cuda_lib, kind, Lc)


@pytest.mark.parametrize("comp", [0, 1, 2])
def test_div_blocks(cuda_lib, comp):
    pc.case_div_blocks(cuda_lib, comp)


@pytest.mark.parametrize("fused", [True, False])
def test_stokes_composite(cuda_lib, fused):
    pc.case_stokes(cuda_lib, fused)


def test_chebyshev_gmg_matches_reference(cuda_lib):
    pc.case_chebyshev_gmg(cuda_lib)
