"""CPU: the hand-written per-element NB/ZINB math (scvi_tools_b200/csrc/fastmath.cuh, compiled for the
host) against float64 evaluations of the oracle's formulas and their autograd derivatives."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
import torch

from oracle import scvi_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host", "fastmath_host.cpp")
SO = os.path.join(HERE, "host", "_fastmath_host.so")
EPS = 1e-8


@pytest.fixture(scope="module")
def lib():
    hdr = os.path.join(HERE, "..", "scvi_tools_b200", "csrc", "fastmath.cuh")
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(hdr)):
        subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-x", "c++", SRC, "-o", SO], check=True)
    return C.CDLL(SO)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _inputs(n, seed):
    rng = np.random.default_rng(seed)
    x = (rng.negative_binomial(5, 0.3, size=n) * rng.binomial(1, 0.7, size=n)).astype(np.float32)
    x[: n // 50] = rng.integers(200, 5000, size=n // 50)  # large counts (series path of lgamma(x+1))
    s = rng.normal(0, 2, size=n).astype(np.float32)
    pi = rng.normal(0, 3, size=n).astype(np.float32)
    pi[: n // 20] *= 10  # saturated dropout logits
    krow = rng.normal(3, 3, size=n).astype(np.float32)  # log2 of the row scale
    theta = np.exp(rng.normal(0, 1.5, size=n)).astype(np.float32)
    return x, s, pi, krow, theta


def _reference(zinb, x, s, pi, krow, theta):
    t = lambda a: torch.tensor(a, dtype=torch.float64)  # noqa: E731
    xs, ss, pis, ths = t(x), t(s).requires_grad_(), t(pi).requires_grad_(), t(theta).requires_grad_()
    mu = torch.exp2(ss * 1.4426950408889634 + t(krow))
    ll = O.log_zinb_positive(xs, mu, ths, pis, EPS) if zinb else O.log_nb_positive(xs, mu, ths, EPS)
    gs, gpi, gth = torch.autograd.grad(ll.sum(), [ss, pis, ths], allow_unused=True)
    gpi = torch.zeros_like(gs) if gpi is None else gpi
    # d ll / d s = mu * dll/dmu  (mu = exp(s + const))  => g1 = -that
    return (ll.detach().numpy().reshape(-1), -gs.numpy().reshape(-1), gpi.numpy().reshape(-1), gth.numpy().reshape(-1),
            mu.detach().numpy().reshape(-1))


@pytest.mark.parametrize("zinb", [0, 1])
def test_elem_fwd_bwd_vs_float64(lib, zinb):
    n = 200000
    x, s, pi, krow, theta = _inputs(n, seed=zinb)
    ll, g1, dpi, dth, mu = _reference(zinb, x, s, pi, krow, theta)
    out = np.empty((n, 4), dtype=np.float32)
    lib.fm_elem_fwd(zinb, n, _p(x), _p(s), _p(pi), _p(krow), _p(theta), C.c_float(EPS), _p(out))
    # lgamma(x+theta) - lgamma(x+1) cancels two O(x ln x) float32 numbers (in torch's float32 evaluation as
    # well): allow their rounding on top of the relative term
    z = x.astype(np.float64) + theta
    scale = 1.0 + np.abs(ll) + 0.1 * z * np.log1p(z)
    # the float32 reference itself carries ~1e-6 relative error in mu (exp of a float32 argument)
    assert np.max(np.abs(out[:, 3] - mu) / mu) < 2e-6
    assert np.max(np.abs(out[:, 0] - ll) / scale) < 5e-6
    small = x < 200
    assert np.mean(np.abs(out[small, 0] - ll[small])) < 5e-6
    gsc = 1.0 + np.abs(g1) + x
    assert np.max(np.abs(out[:, 1] - g1) / gsc) < 5e-6
    assert np.max(np.abs(out[:, 2] - dpi)) < 1e-5
    # table-driven branch-free variant of the fused training kernel: same bars
    ot = np.empty((n, 4), dtype=np.float32)
    lib.fm_elem_fwd_tab(zinb, n, _p(x), _p(s), _p(pi), _p(krow), _p(theta), C.c_float(EPS), _p(ot))
    assert np.max(np.abs(ot[:, 3] - mu) / mu) < 2e-6
    assert np.max(np.abs(ot[:, 0] - ll) / scale) < 5e-6
    assert np.mean(np.abs(ot[small, 0] - ll[small])) < 5e-6
    assert np.max(np.abs(ot[:, 1] - g1) / gsc) < 5e-6
    assert np.max(np.abs(ot[:, 2] - dpi)) < 1e-5
    ob = np.empty((n, 4), dtype=np.float32)
    lib.fm_elem_bwd(zinb, n, _p(x), _p(s), _p(pi), _p(krow), _p(theta), C.c_float(EPS), _p(ob))
    assert np.max(np.abs(ob[:, 0] - g1) / gsc) < 5e-6
    assert np.max(np.abs(ob[:, 1] - dpi)) < 1e-5
    tsc = 1.0 + np.abs(dth) + x / theta
    assert np.max(np.abs(ob[:, 2] - dth) / tsc) < 1e-5
    # table-driven branch-free variant of the dW kernel: same bars
    obt = np.empty((n, 4), dtype=np.float32)
    lib.fm_elem_bwd_tab(zinb, n, _p(x), _p(s), _p(pi), _p(krow), _p(theta), C.c_float(EPS), _p(obt))
    assert np.max(np.abs(obt[:, 0] - g1) / gsc) < 5e-6
    assert np.max(np.abs(obt[:, 1] - dpi)) < 1e-5
    assert np.max(np.abs(obt[:, 2] - dth) / tsc) < 1e-5


@pytest.mark.parametrize("zinb", [0, 1])
def test_lean_elements_vs_float64(lib, zinb):
    """The straight-line variants of the fused training kernels (integer counts below the table length;
    tiny / huge mu and saturated logits included)."""
    n = 200000
    x, s, pi, krow, theta = _inputs(n, seed=10 + zinb)
    x = np.minimum(x, 47.0).astype(np.float32)  # below both table lengths (LGT 64, DGT 48)
    krow[: n // 10] -= 30.0  # mu down to ~1e-9: the eps terms matter
    ll, g1, dpi, dth, mu = _reference(zinb, x, s, pi, krow, theta)
    out = np.empty((n, 4), dtype=np.float32)
    lib.fm_elem_fwd_lean(zinb, n, _p(x), _p(s), _p(pi), _p(krow), _p(theta), C.c_float(EPS), _p(out))
    z = x.astype(np.float64) + theta
    scale = 1.0 + np.abs(ll) + 0.1 * z * np.log1p(z)
    ok = mu > 1e-30
    assert np.max(np.abs(out[ok, 3] - mu[ok]) / mu[ok]) < 2e-5  # exp of a float32 argument of magnitude ~60
    assert np.max(np.abs(out[:, 0] - ll) / scale) < 5e-6
    assert np.mean(np.abs(out[:, 0] - ll)) < 5e-6
    gsc = 1.0 + np.abs(g1) + x
    assert np.max(np.abs(out[:, 1] - g1) / gsc) < 5e-6
    assert np.max(np.abs(out[:, 2] - dpi)) < 1e-5
    ob = np.empty((n, 4), dtype=np.float32)
    lib.fm_elem_bwd_lean(zinb, n, _p(x), _p(s), _p(pi), _p(krow), _p(theta), C.c_float(EPS), _p(ob))
    assert np.max(np.abs(ob[:, 0] - g1) / gsc) < 5e-6
    assert np.max(np.abs(ob[:, 1] - dpi)) < 1e-5
    tsc = 1.0 + np.abs(dth) + x / theta
    assert np.max(np.abs(ob[:, 2] - dth) / tsc) < 1e-5
    # the lean and the table-driven generic variants are interchangeable element by element
    ot = np.empty((n, 4), dtype=np.float32)
    lib.fm_elem_fwd_tab(zinb, n, _p(x), _p(s), _p(pi), _p(krow), _p(theta), C.c_float(EPS), _p(ot))
    assert np.max(np.abs(ot[:, 0] - out[:, 0]) / scale) < 2e-6
    assert np.max(np.abs(ot[:, 1] - out[:, 1]) / gsc) < 2e-6


def test_lgamma_digamma_series(lib):
    from scipy.special import digamma, gammaln

    z = np.concatenate([np.exp(np.linspace(np.log(1e-3), np.log(1e5), 20000)), np.arange(1, 300)]).astype(np.float32)
    out = np.empty_like(z)
    lib.fm_lgamma(z.size, _p(z), _p(out))
    ref = gammaln(z.astype(np.float64))
    assert np.max(np.abs(out - ref) / (1.0 + np.abs(ref))) < 2e-6
    lib.fm_digamma(z.size, _p(z), _p(out))
    ref = digamma(z.astype(np.float64))
    assert np.max(np.abs(out - ref) / (1.0 + np.abs(ref))) < 2e-6


def test_ex2_on_the_fma_pipe(lib):
    """`ex2_fma` (2^x for x <= 0 without an SFU op): relative error below 3e-7 (ex2.approx's own accuracy) over the
    whole range the likelihood uses, exact 1 at 0, flushes to (nearly) zero below -125."""
    x = np.concatenate([-np.linspace(0.0, 130.0, 200001), -np.logspace(-8, 2, 20001)]).astype(np.float32)
    y = np.empty_like(x)
    lib.fm_ex2_fma(len(x), x.ctypes.data_as(C.POINTER(C.c_float)), y.ctypes.data_as(C.POINTER(C.c_float)))
    ref = np.exp2(np.maximum(x.astype(np.float64), -125.0))
    assert np.max(np.abs(y / ref - 1.0)) < 3e-7
    assert y[0] == 1.0 or abs(y[0] - 1.0) < 1.2e-7
    assert np.all(y > 0) and np.all(np.isfinite(y))
