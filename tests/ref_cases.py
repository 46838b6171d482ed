"""Shared by the CPU pinning tests and the -m gpu tests: access to tests/golden/ref_p1.npz, the outputs of the REFERENCE ITSELF
(HyTeG built from /root/reference by oracle/build_ref.py, driven by oracle/ref_dump.cpp, written by
oracle/make_ref_fixtures.py).  Vectors are stored over the oracle's unique-DoF numbering; inputs are regenerated from seeds."""
import os

import numpy as np

from hyteg_b200 import meshes
from oracle import p1_oracle as po

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_p1.npz")
CASES = (("cube_6el", (3, 4, 5)), ("three_tets_with_two_joint_faces", (3, 4)), ("regular_octahedron_8el", (3, 4)))
_cache = {}


def gold():
    if "g" not in _cache:
        _cache["g"] = np.load(GOLD)
    return _cache["g"]


def oracle(name, max_level, min_level=0):
    key = (name, min_level, max_level)
    if key not in _cache:
        v, c = meshes.golden(name)
        _cache[key] = po.P1Oracle(v, c, min_level, max_level)
    return _cache[key]


def inputs(o, L, seed):
    """same generator as oracle/make_ref_fixtures.py: inputs()"""
    rng = np.random.default_rng(seed)
    lv = o.levels[L]
    x = rng.random(lv.n)
    b = rng.random(lv.n) - 0.5
    return x, b


def gmg_start(o, L):
    """initial guess of the V-cycle fixtures (P1GMG3DConvergenceTest.cpp:96-101): random inner values, smooth boundary values"""
    lv = o.levels[L]
    x, _ = inputs(o, L, 100 + L)
    m = o.mask(L, po.INNER)
    u0 = np.sin(lv.xyz[:, 0]) * np.sinh(lv.xyz[:, 1]) * lv.xyz[:, 2]
    uex = u0.copy()
    u0[m] = x[m]
    return u0, uex


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


# ---- P2 elementwise path (tests/golden/ref_p2.npz, oracle/make_ref_fixtures_p2.py) ---------------------------------------------------
GOLD_P2 = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_p2.npz")
SHELL = (3, 2, 0.5, 1.0)
BARY = np.array([[0.25, 0.25, 0.25, 0.25], [0.7, 0.1, 0.1, 0.1], [0.1, 0.6, 0.2, 0.1], [0.05, 0.15, 0.3, 0.5], [1, 0, 0, 0]])


def gold_p2():
    if "g2" not in _cache:
        _cache["g2"] = np.load(GOLD_P2)
    return _cache["g2"]


def p2_oracle(kind, L):
    from oracle import p2_oracle as p2
    key = ("p2", kind, L)
    if key not in _cache:
        v, c = meshes.spherical_shell(*SHELL) if kind == "shell" else meshes.cube_6el()
        _cache[key] = p2.P2Oracle(v, c, L, shell=(kind == "shell"))
    return _cache[key]


def p2_inputs(n, seed):
    """same generator as oracle/make_ref_fixtures_p2.py: p2_inputs()"""
    rng = np.random.default_rng(seed)
    return rng.random(n) - 0.3, rng.random(n)
