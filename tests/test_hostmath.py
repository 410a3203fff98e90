"""not-gpu tier: the PRODUCT's device math header (combat-seq_b200/csrc/nbmath.cuh) compiled for the host
(tests/hostmath/hostmath.cu) — the exact pmf / continued-fraction cdf / quantile-search / match_quantiles routines the
kernels run — against the golden known answers and the oracle.  No GPU, no product library call."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
f64p = C.POINTER(C.c_double)


@pytest.fixture(scope="module")
def hm(hostmath_lib):
    lib = hostmath_lib

    def call(fn, *arrs, extra=()):
        arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in np.broadcast_arrays(*arrs)]
        res = np.empty(arrs[0].size)
        getattr(lib, fn)(C.c_int64(res.size), *[a.ctypes.data_as(f64p) for a in arrs], res.ctypes.data_as(f64p), *extra)
        return res
    return call


@pytest.fixture(scope="module")
def kat():
    return np.load(os.path.join(GOLD, "nb_kat.npz"))


def test_device_pmf_cdf_against_mpmath(hm, kat):
    x, size, mu = kat["x_size_mu"].T
    m = kat["logpmf"] > -700
    got = hm("hm_nb_logpmf", x, size, mu)
    assert np.all(np.abs(got[m] - kat["logpmf"][m]) <= 2e-13 * np.maximum(1, np.abs(kat["logpmf"][m])))
    mc = np.isfinite(kat["cdf"]) & (kat["cdf"] > 1e-300)
    for fn in ("hm_nb_cdf", "hm_nb_cdf_fast"):
        c = hm(fn, x[mc], size[mc], mu[mc])
        assert np.max(np.abs(c - kat["cdf"][mc]) / kat["cdf"][mc]) < 5e-13, fn


def test_device_quantile_against_mpmath(hm, kat):
    p, size, mu = kat["q_p_size_mu"].T
    for fn in ("hm_nb_quantile", "hm_nb_quantile_fast"):
        assert np.array_equal(hm(fn, p, size, mu), kat["quantile"]), fn


def test_device_math_against_oracle_random(hm, oracle):
    rng = np.random.default_rng(21)
    n = 20000
    size = 1 / np.exp(rng.uniform(np.log(2e-3), np.log(20), n))
    mu = np.exp(rng.uniform(np.log(0.05), np.log(5e4), n))
    x = np.floor(rng.gamma(size, mu / size) * rng.uniform(0.5, 1.5, n))
    a, b = hm("hm_nb_cdf_fast", x, size, mu), oracle.nb_cdf(x, size, mu)
    m = b > 1e-250
    assert np.max(np.abs(a[m] - b[m]) / b[m]) < 2e-12
    pm, po = hm("hm_nb_pmf_fast", x, size, mu), oracle.nb_pmf(x, size, mu)
    m = po > 1e-250
    assert np.max(np.abs(pm[m] - po[m]) / po[m]) < 2e-12
    p = rng.uniform(1e-5, 1 - 1e-5, n)
    qa, qb = hm("hm_nb_quantile_fast", p, size, mu), oracle.nb_quantile(p, size, mu)
    # a uniform p sits within rounding of a cdf step with probability ~1e-12: the search must agree everywhere
    assert np.array_equal(qa, qb)


def test_device_match_quantiles_fast_equals_plain_and_oracle(hm, oracle):
    rng = np.random.default_rng(22)
    n = 20000
    old_mu = np.exp(rng.normal(4, 2, n)).clip(1e-3, 1e6)
    new_mu = old_mu * np.exp(rng.normal(0, 0.7, n))
    old_phi = np.exp(rng.uniform(np.log(3e-3), np.log(8), n))
    new_phi = old_phi * np.exp(rng.normal(0, 0.3, n))
    x = np.floor(rng.gamma(1 / old_phi, old_mu * old_phi))
    fast = hm("hm_match_quantiles", x, old_mu, old_phi, new_mu, new_phi, extra=(C.c_int(1),))
    plain = hm("hm_match_quantiles", x, old_mu, old_phi, new_mu, new_phi, extra=(C.c_int(0),))
    ref = np.array([oracle.match_quantiles(np.array([[xi]]), np.array([[om]]), np.array([op]), np.array([[nm]]), np.array([np_]))[0, 0]
                    for xi, om, op, nm, np_ in zip(x[:3000], old_mu[:3000], old_phi[:3000], new_mu[:3000], new_phi[:3000])])
    small = hm("hm_match_quantiles", x, old_mu, old_phi, new_mu, new_phi, extra=(C.c_int(2),))
    assert np.array_equal(fast, plain)
    assert np.array_equal(small, plain)                      # direct pmf summation for small counts: same integers
    assert np.array_equal(fast[:3000], ref)
    # Same distribution on both sides: the identity, thanks to qnbinom's 64 eps fuzz — except where R's own two cdf
    # parametrisations (pnbinom_mu for p, pnbinom(size, prob) inside qnbinom, whose 1 - prob rounds) differ by more
    # than the fuzz (size >> mu).  The device routine must reproduce exactly the oracle's set of exceptions.
    ident = hm("hm_match_quantiles", x, old_mu, old_phi, old_mu, old_phi, extra=(C.c_int(1),))
    assert np.array_equal(ident, hm("hm_match_quantiles", x, old_mu, old_phi, old_mu, old_phi, extra=(C.c_int(2),)))
    bad = np.flatnonzero(ident != x)
    assert bad.size <= 1e-3 * n
    for i in bad:
        one = oracle.match_quantiles(np.array([[x[i]]]), np.array([[old_mu[i]]]), np.array([old_phi[i]]),
                                     np.array([[old_mu[i]]]), np.array([old_phi[i]]))[0, 0]
        assert one == ident[i]


def test_fast_log_for_positive_normal_arguments(hm):
    # combat-seq_b200/csrc/fastmath.cuh (host compilation of the same series): within 2 ulp of libm over 600 binades
    rng = np.random.default_rng(5)
    x = np.concatenate([np.exp(rng.uniform(-200, 200, 200000)), 1 + rng.uniform(-0.3, 0.42, 100000),
                        1 + rng.normal(0, 1e-6, 50000), 2.0 ** rng.integers(-300, 300, 1000),
                        np.sqrt(2) * (1 + rng.normal(0, 1e-7, 20000))])
    got, ref = hm("hm_log_pos", x), np.log(x)
    ulp = np.spacing(np.abs(ref))
    err = np.abs(got - ref) / np.maximum(ulp, 5e-324)
    m = ref != 0
    assert err[m].max() <= 2.0, err[m].max()
    assert np.all(got[~m] == 0)


def test_vanishing_dispersion_follows_r35_qnbinom(hm):
    """Dispersions of 1e-20 (the lower end of Brent's bracket on Poisson-like data) make prob = size / (size + mu) round to 1:
    R 3.5's qnbinom returns 0 there (nmath/qnbinom.c: "if (prob == 1 || size == 0) return 0"), so match_quantiles maps every count
    above 1 that is not passed through to 1.  The fast routines (small-count sums included) must do what the plain ones do."""
    rng = np.random.default_rng(0)
    mu = np.exp(rng.normal(3, 1.5, 3000))
    x = rng.poisson(mu).astype(float)
    nmu = mu * np.exp(rng.normal(0, 0.3, 3000))
    for phi in (1e-20, 1e-300):
        plain = hm("hm_match_quantiles", x, mu, phi, nmu, phi, extra=(C.c_int(0),))
        fast = hm("hm_match_quantiles", x, mu, phi, nmu, phi, extra=(C.c_int(2),))
        assert np.array_equal(plain, fast)
        assert set(np.unique(plain[x > 1])) <= {1.0} | set(np.unique(x))
    for phi in (1e-8, 1e-12):                          # still a proper negative binomial, indistinguishable from Poisson
        plain = hm("hm_match_quantiles", x, mu, phi, nmu, phi, extra=(C.c_int(0),))
        fast = hm("hm_match_quantiles", x, mu, phi, nmu, phi, extra=(C.c_int(2),))
        assert np.array_equal(plain, fast)
