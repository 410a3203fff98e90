"""not-gpu tier: the oracle's NB distribution functions (oracle/ref_nmath.cpp, restating R nmath as called at
/root/reference/helper_seq.R:28,53,58) against the golden known answers (tests/golden/nb_kat.npz, mpmath at 50
digits) and against scipy, which shares no code with the restatement."""
import os

import numpy as np
import pytest
from scipy import special, stats

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def kat():
    return np.load(os.path.join(GOLD, "nb_kat.npz"))


def test_pmf_against_mpmath(oracle, kat):
    x, size, mu = kat["x_size_mu"].T
    with np.errstate(divide="ignore"):
        got = np.log(oracle.nb_pmf(x, size, mu))
    m = kat["logpmf"] > -700                                    # dnbinom underflows to 0 below exp(-745)
    assert m.sum() > 350
    err = np.abs(got[m] - kat["logpmf"][m])
    assert np.all(err <= 2e-13 * np.maximum(1.0, np.abs(kat["logpmf"][m])))
    assert np.all(oracle.nb_pmf(x[~m], size[~m], mu[~m]) < 1e-300)


def test_cdf_against_mpmath(oracle, kat):
    x, size, mu = kat["x_size_mu"].T
    m = np.isfinite(kat["cdf"]) & (kat["cdf"] > 1e-300)
    got = oracle.nb_cdf(x[m], size[m], mu[m])
    assert m.sum() > 300
    assert np.max(np.abs(got - kat["cdf"][m]) / kat["cdf"][m]) < 5e-13


def test_quantile_against_mpmath(oracle, kat):
    p, size, mu = kat["q_p_size_mu"].T
    assert p.size > 200
    assert np.array_equal(oracle.nb_quantile(p, size, mu), kat["quantile"])


def test_cdf_two_code_paths_agree(oracle):
    # TOMS 708 bratio vs direct long-double summation of the pmf recurrence (SURVEY 8c "oracle self-consistency")
    rng = np.random.default_rng(3)
    size = 1 / np.exp(rng.uniform(np.log(1e-3), np.log(5), 4000))
    mu = np.exp(rng.uniform(np.log(0.05), np.log(3e3), 4000))
    x = rng.poisson(mu * rng.uniform(0.3, 2.5, 4000)).astype(float)
    a, b = oracle.nb_cdf(x, size, mu), oracle.nb_cdf_bysum(x, size, mu)
    m = b > 1e-290
    assert m.sum() > 3000
    assert np.max(np.abs(a[m] - b[m]) / b[m]) < 1e-12


def test_against_scipy_random_grid(oracle):
    rng = np.random.default_rng(11)
    n = 5000
    size = 1 / np.exp(rng.uniform(np.log(1e-3), np.log(10), n))
    mu = np.exp(rng.uniform(np.log(1e-2), np.log(1e5), n))
    x = np.floor(rng.gamma(size, mu / size) * rng.uniform(0.5, 1.5, n))
    prob = size / (size + mu)
    ref_lp = stats.nbinom.logpmf(x, size, prob)
    with np.errstate(divide="ignore"):
        got_lp = np.log(oracle.nb_pmf(x, size, mu))
    m = ref_lp > -600
    assert np.max(np.abs(got_lp[m] - ref_lp[m]) / np.maximum(1, np.abs(ref_lp[m]))) < 1e-9      # scipy's gammaln-based logpmf is the looser side
    ref_c = special.betainc(size, x + 1, prob)
    got_c = oracle.nb_cdf(x, size, mu)
    m = ref_c > 1e-200
    assert np.max(np.abs(got_c[m] - ref_c[m]) / ref_c[m]) < 1e-10


def test_quantile_is_smallest_with_cdf_at_least_p(oracle):
    rng = np.random.default_rng(5)
    n = 3000
    size = 1 / np.exp(rng.uniform(np.log(5e-3), np.log(5), n))
    mu = np.exp(rng.uniform(np.log(0.1), np.log(2e4), n))
    p = rng.uniform(1e-6, 1 - 1e-6, n)
    y = oracle.nb_quantile(p, size, mu)
    pf = p * (1 - 64 * np.finfo(float).eps)
    assert np.all(oracle.nb_cdf(y, size, mu) >= pf)
    pos = y > 0
    assert np.all(oracle.nb_cdf(y[pos] - 1, size[pos], mu[pos]) < pf[pos])


def test_qnorm_stirlerr_bd0(oracle):
    p = np.concatenate([np.logspace(-300, -1, 200), np.linspace(0.1, 0.9, 81), 1 - np.logspace(-15, -1, 60)])
    ref = special.ndtri(p)
    assert np.max(np.abs(oracle.qnorm(p) - ref) / np.maximum(1, np.abs(ref))) < 1e-13
    n = np.concatenate([np.arange(1, 31) / 2.0, np.array([15.5, 20.25, 36, 81, 501, 1e4, 1e7])])
    import mpmath as mp
    mp.mp.dps = 40
    ref = np.array([float(mp.loggamma(mp.mpf(v) + 1) - (mp.mpf(v) + 0.5) * mp.log(v) + v - mp.log(2 * mp.pi) / 2) for v in n])
    assert np.max(np.abs(oracle.stirlerr(n) - ref) / ref) < 5e-13     # R truncates the series: 1.5e-13 at n = 501
    x = np.array([1.0, 5.0, 90.0, 1e3, 1e5]); npv = x * np.array([0.5, 0.99, 1.0, 1.2, 1.0001])
    ref = x * np.log(x / npv) + npv - x
    assert np.max(np.abs(oracle.bd0(x, npv) - ref) / np.maximum(1e-300, np.abs(ref) + 1e-9)) < 1e-6
    assert np.all(oracle.bd0(x, x) == 0)


def test_pbeta_against_scipy(oracle):
    rng = np.random.default_rng(9)
    a = np.exp(rng.uniform(np.log(0.05), np.log(5e3), 3000))
    b = np.floor(np.exp(rng.uniform(0, np.log(5e4), 3000))) + 1
    xx = a / (a + b * rng.uniform(0.3, 3, 3000))
    ref = special.betainc(a, b, xx)
    m = ref > 1e-250
    assert np.max(np.abs(oracle.pbeta(xx, a, b)[m] - ref[m]) / ref[m]) < 1e-10
