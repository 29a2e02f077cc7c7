"""geom.cuh (the float64 NeRF / Kabsch device functions) compiled for the host with g++ and compared
with the oracle's numpy restatement of nerf.py / algo.py.  Tolerance 1e-9 A (north_star: 1e-4 A)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import geobpe_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("geom") / "libgeom_host.so")
    subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", "-I", os.path.join(ROOT, "pt-bpe_b200/geobpe_b200/csrc"),
                           os.path.join(ROOT, "tests/csrc/geom_host.cpp"), "-o", out])
    l = C.CDLL(out)
    l.t_rmsd.restype = C.c_double
    l.t_rmsd.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    l.t_first_residue.argtypes = [C.c_double] * 3 + [C.c_void_p]
    l.t_place.argtypes = [C.c_void_p] + [C.c_double] * 3 + [C.c_void_p]
    return l


def test_first_residue_and_place(lib):
    rng = np.random.default_rng(0)
    for _ in range(200):
        nca, cac, tau = 1.458 + 0.02 * rng.standard_normal(), 1.525 + 0.02 * rng.standard_normal(), 1.94 + 0.1 * rng.standard_normal()
        out = np.zeros(9)
        lib.t_first_residue(nca, cac, tau, out.ctypes.data)
        ref = np.concatenate(O.update_backbone_positions(O.N_INIT, O.CA_INIT, O.C_INIT, cac, nca, tau))
        np.testing.assert_allclose(out, ref, rtol=0, atol=1e-12)
        abc = out.copy()
        ang, ln, tor = rng.uniform(1.8, 2.3), rng.uniform(1.3, 1.6), rng.uniform(-np.pi, np.pi)
        d = np.zeros(3)
        lib.t_place(abc.ctypes.data, ang, ln, tor, d.ctypes.data)
        ref = O.place_dihedral(abc[0:3], abc[3:6], abc[6:9], ang, ln, tor)
        np.testing.assert_allclose(d, ref, rtol=0, atol=1e-12)


@pytest.mark.parametrize("A", [3, 4, 7, 10, 25])
def test_kabsch_rmsd(lib, A):
    rng = np.random.default_rng(A)
    worst = 0.0
    for it in range(300):
        P = rng.standard_normal((A, 3)) * 2
        if it % 3 == 0:   # near-identical pair (small RMSD)
            Q = P + 1e-3 * rng.standard_normal((A, 3))
        elif it % 3 == 1:  # reflected -> det < 0 branch
            Q = P * np.array([1, 1, -1.0]) + 0.1 * rng.standard_normal((A, 3))
        else:
            Q = rng.standard_normal((A, 3)) * 2
        R = np.linalg.qr(rng.standard_normal((3, 3)))[0]
        Q = Q @ R + rng.standard_normal(3)
        got = lib.t_rmsd(np.ascontiguousarray(P).ctypes.data, np.ascontiguousarray(Q).ctypes.data, A)
        ref = O.compute_rmsd(P, Q)
        worst = max(worst, abs(got - ref))
    assert worst < 1e-9, worst
    P = rng.standard_normal((A, 3))
    assert lib.t_rmsd(P.ctypes.data, P.ctypes.data, A) < 1e-7
