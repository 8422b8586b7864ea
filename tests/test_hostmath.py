"""The product's per-particle device math (csrc/svd3.cuh, csrc/constitutive.cuh), compiled as plain C++
by tests/hostmath/host_math.cpp, against the CPU oracle on random deformation gradients.  This is the
no-GPU check of the return mappings / stresses / SVD / polar rotation; the same comparison runs on the
device in tests/test_gpu_parity.py::test_constitutive_against_oracle."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import mpm_oracle as O

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hostmath")
_fp = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


@pytest.fixture(scope="module")
def hm():
    so = os.path.join(HERE, "libhostmath.so")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-fvisibility=hidden", "-o", so,
                           os.path.join(HERE, "host_math.cpp")])
    L = C.CDLL(so)
    L.hm_constitutive.argtypes = [C.c_int, _fp, _ip, _fp, _fp, _fp, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float,
                                  C.c_int, C.c_uint, _fp, _fp]
    L.hm_svd3.argtypes = [C.c_int, _fp, _fp, _fp, _fp]
    L.hm_polar3.argtypes = [C.c_int, _fp, _fp]
    L.hm_polar3.restype = C.c_int
    return L


def _random_F(n, seed, scale):
    rng = np.random.default_rng(seed)
    return (np.eye(3)[None] + scale * rng.standard_normal((n, 3, 3))).astype(np.float32)


@pytest.mark.parametrize("scale", [1e-4, 1e-2, 0.15, 0.6])
def test_svd3_contract(hm, scale):
    n = 20000
    A = _random_F(n, 1, scale)
    U = np.empty_like(A); V = np.empty_like(A); s = np.empty((n, 3), np.float32)
    hm.hm_svd3(n, A.reshape(n, 9), U.reshape(n, 9), s, V.reshape(n, 9))
    rec = np.einsum("nij,nj,nkj->nik", U.astype(np.float64), s.astype(np.float64), V.astype(np.float64))
    assert np.abs(rec - A).max() < 5e-6 * max(1.0, np.abs(A).max())
    assert np.abs(np.linalg.det(U.astype(np.float64)) - 1).max() < 1e-5
    assert np.abs(np.linalg.det(V.astype(np.float64)) - 1).max() < 1e-5
    assert np.all(np.abs(s[:, 0]) >= np.abs(s[:, 1]) - 1e-6) and np.all(np.abs(s[:, 1]) >= np.abs(s[:, 2]) - 1e-6)
    sv = np.linalg.svd(A.astype(np.float64), compute_uv=False)
    assert np.abs(np.abs(s) - sv).max() < 5e-6 * sv.max()
    assert np.array_equal(np.sign(s[:, 2]), np.sign(np.linalg.det(A.astype(np.float64))).astype(np.float32))


@pytest.mark.parametrize("scale", [1e-4, 1e-2, 0.15, 0.4])
def test_polar3_is_the_rotation_factor(hm, scale):
    n = 20000
    A = _random_F(n, 2, scale)
    det = np.linalg.det(A.astype(np.float64))
    A = np.ascontiguousarray(A[det > 1e-3])
    n = len(A)
    R = np.empty_like(A)
    bad = hm.hm_polar3(n, A.reshape(n, 9), R.reshape(n, 9))
    assert bad == 0
    U, s, Vt = np.linalg.svd(A.astype(np.float64))
    want = U @ Vt
    # the rotation factor's condition number is 2 / (s_2 + s_3)
    cond = 2.0 / (s[:, 1] + s[:, 2])
    assert (np.abs(R - want).reshape(n, -1).max(1) / cond).max() < 2e-6


MATS = ["jelly", "metal", "sand", "foam", "fluid", "plasticine"]


@pytest.mark.parametrize("material", MATS)
@pytest.mark.parametrize("strict", [True, False])
def test_constitutive_matches_oracle(hm, material, strict):
    n = 6000
    F = _random_F(n, 33, 0.15)
    F[: n // 4] = _random_F(n // 4, 34, 1e-3)
    mat_id = O.MATERIAL_TYPES[material]
    E, nu = 2e5, 0.3
    mu0 = np.full(n, E / (2 * (1 + nu)), np.float32)
    lam0 = np.full(n, E * nu / ((1 + nu) * (1 - 2 * nu)), np.float32)
    ys0 = np.full(n, 5e3, np.float32)
    flags = (O.STRICT_Q1 | O.STRICT_Q3) if strict else 0
    P = O.OrcParams(alpha=0.26776504, plastic_viscosity=0.008, xi=1.0, softening=1.0, hardening=1, flags=flags)
    dt = 1e-4
    # oracle
    mu_o, lam_o, ys_o = mu0.copy(), lam0.copy(), ys0.copy()
    Fo = np.zeros((n, 9), np.float32); So = np.zeros((n, 9), np.float32)
    O.lib().orc_compute_stress_from_F_trial(n, np.full(n, mat_id, np.float32), mu_o, lam_o, ys_o, F.reshape(n, 9), Fo, So,
                                            None, dt, C.byref(P))
    # device math on the host
    mu_h, lam_h, ys_h = mu0.copy(), lam0.copy(), ys0.copy()
    Fh = np.zeros((n, 9), np.float32); s6 = np.zeros((n, 6), np.float32)
    hm.hm_constitutive(n, F.reshape(n, 9), np.full(n, mat_id, np.int32), mu_h, lam_h, ys_h, dt, P.alpha, P.plastic_viscosity,
                       P.xi, P.softening, P.hardening, flags, Fh, s6)
    Sh = s6[:, [0, 1, 2, 1, 3, 4, 2, 4, 5]]
    ok = np.isfinite(So).all(1) & np.isfinite(Fo).all(1)
    assert ok.mean() > 0.7
    assert np.array_equal(np.isfinite(Sh).all(1), np.isfinite(So).all(1))
    if material == "foam" and strict:
        s = np.linalg.svd(F.astype(np.float64), compute_uv=False)
        ok &= (np.minimum(s[:, 0] - s[:, 1], s[:, 1] - s[:, 2]) > 0.05)
    rel = np.linalg.norm(Fh[ok].astype(np.float64) - Fo[ok]) / np.linalg.norm(Fo[ok].astype(np.float64))
    assert rel < 2e-5, rel
    scale = float(np.abs(So[ok]).max()) if np.abs(So[ok]).max() > 0 else 1.0
    assert np.abs(Sh[ok] - So[ok]).max() <= 2e-4 * scale
    assert np.allclose(ys_h[ok], ys_o[ok], rtol=1e-4, atol=1e-3)
    assert np.allclose(mu_h[ok], mu_o[ok], rtol=1e-5)
