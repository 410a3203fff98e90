"""not-gpu tier: the oracle's edgeR restatement (oracle/ref_edger.cpp; call sites /root/reference/ComBat_seq.R:103-134)
against independent checks: the NB score equations, a numpy/scipy restatement (oracle/pyref.py), brute-force
scans and dense numeric maximisation (SURVEY.md 8c "pins the new repo must create")."""
import os
import sys

import numpy as np
import pytest
from scipy import special

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pyref  # noqa: E402


def nb_counts(rng, G, n, phi=0.2, lo=1.0, hi=8.0):
    mu = np.exp(rng.uniform(lo, hi, G))[:, None] * rng.uniform(0.7, 1.3, (1, n))
    return rng.poisson(rng.gamma(1 / phi, mu * phi)).astype(float)


def designs():
    n = 12
    b = np.repeat([0, 1, 2], 4)
    bm = np.zeros((n, 3)); bm[np.arange(n), b] = 1
    grp = np.tile([0.0, 1.0], 6)
    cov = np.random.default_rng(0).normal(size=n)
    return {"oneway": bm, "levenberg": np.column_stack([bm, grp]), "levenberg_cov": np.column_stack([bm, grp, cov])}


@pytest.mark.parametrize("kind", ["oneway", "levenberg", "levenberg_cov"])
def test_glm_fit_solves_score_equations(oracle, kind):
    rng = np.random.default_rng(1)
    X = designs()[kind]
    G, n = 300, X.shape[0]
    y = nb_counts(rng, G, n)
    lib = y.sum(axis=0)
    phi = rng.uniform(0.05, 0.6, (G, 3))
    cols = np.repeat([0, 1, 2], 4).astype(np.int32)
    coef, mu = oracle.glm_fit(y, X, phi, cols, lib, prior_count=0.0)
    disp = phi[:, cols]
    score = ((y - mu) / (1 + disp * mu)) @ X                  # d loglik / d beta  (log link)
    info = (mu / (1 + disp * mu)) @ (X * X)
    # one-way cells are solved to |step| < 1e-10; Levenberg stops on dl'dbeta < 1e-6  (edgeR tolerances)
    tol = 1e-7 if kind == "oneway" else 2e-3
    assert np.max(np.abs(score) / np.sqrt(info)) < tol
    assert np.allclose(mu, np.exp(coef @ X.T + np.log(lib)[None, :]), rtol=1e-9)


@pytest.mark.parametrize("kind", ["oneway", "levenberg_cov"])
def test_glm_fit_matches_numpy_restatement(oracle, kind):
    rng = np.random.default_rng(2)
    X = designs()[kind]
    G, n = 40, X.shape[0]
    y = nb_counts(rng, G, n)
    y[0, :4] = 0                                               # an all-zero one-way cell / a diverging coefficient
    lib = y.sum(axis=0)
    phi = rng.uniform(0.05, 0.6, (G, 3))
    cols = np.repeat([0, 1, 2], 4).astype(np.int32)
    go = rng.normal(0, 0.3, G)
    coef, mu = oracle.glm_fit(y, X, phi, cols, lib, gene_offset=go, prior_count=1e-4)
    off = np.log(lib)[None, :] + go[:, None]
    c2, m2 = pyref.glm_fit(y, X, phi[:, cols], off, 1e-4)
    ok = np.ones(G, bool); ok[0] = kind == "oneway"           # gene 0 under Levenberg: fragile by design
    assert np.allclose(mu[ok], m2[ok], rtol=1e-6, atol=1e-9)
    assert np.allclose(coef[ok], c2[ok], rtol=1e-6, atol=1e-8)


def test_apl_matches_definition(oracle):
    rng = np.random.default_rng(3)
    X = np.column_stack([np.ones(8), np.tile([0.0, 1.0], 4)])
    G = 60
    y = nb_counts(rng, G, 8)
    lib = y.sum(axis=0)
    phis = np.array([0.01, 0.1, 0.5, 2.0])
    got = oracle.apl(y, X, lib, phis)
    off = np.repeat(np.log(lib)[None, :], G, 0)
    for k, phi in enumerate(phis):
        ref = pyref.adjusted_profile_lik(float(phi), y, X, off)
        assert np.allclose(got[:, k], ref, rtol=1e-9, atol=1e-7)
    # ncoef == 1 branch: adj = log|sum W| / 2
    got1 = oracle.apl(y, None, lib, phis[:2])
    ref1 = pyref.adjusted_profile_lik(float(phis[0]), y, np.ones((8, 1)), off)
    assert np.allclose(got1[:, 0], ref1, rtol=1e-9, atol=1e-7)


def test_unit_deviance_branches(oracle):
    for y, mu, phi in [(5.0, 3.0, 0.2), (0.0, 3.0, 0.2), (5.0, 3.0, 5e-5), (7.0, 1e8, 0.5), (0.0, 0.0, 0.1)]:
        ref = pyref.unit_dev(np.array([y]), np.array([mu]), np.array([phi]))[0]
        assert oracle.unit_nb_deviance(y, mu, phi) == pytest.approx(ref, rel=1e-12, abs=1e-300)
    # general branch equals the textbook NB deviance
    y, mu, phi = 9.0, 4.0, 0.3
    ref = 2 * (y * np.log(y / mu) - (y + 1 / phi) * np.log((1 + phi * y) / (1 + phi * mu)))
    assert oracle.unit_nb_deviance(y, mu, phi) == pytest.approx(ref, rel=1e-7)


def test_fmm_spline_reproduces_cubics_and_interpolates(oracle):
    x = np.linspace(-6, 6, 11)
    y = 0.3 * x ** 3 - 1.1 * x ** 2 + 0.7 * x - 2
    b, c, d = oracle.fmm_spline(x, y)
    assert np.allclose(b, 0.9 * x ** 2 - 2.2 * x + 0.7, atol=1e-10)
    assert np.allclose(c[:-1], (1.8 * x - 2.2)[:-1] / 2, atol=1e-10)
    assert np.allclose(d[:-1], 0.3, atol=1e-10)
    rng = np.random.default_rng(4)
    y = rng.normal(size=11)
    b, c, d = oracle.fmm_spline(x, y)
    h = np.diff(x)
    assert np.allclose(y[:-1] + b[:-1] * h + c[:-1] * h ** 2 + d[:-1] * h ** 3, y[1:], atol=1e-10)       # C0
    assert np.allclose(b[:-1] + 2 * c[:-1] * h + 3 * d[:-1] * h ** 2, b[1:], atol=1e-10)                # C1
    pb, pc, pd = pyref.fmm_spline(x, y)
    assert np.allclose(b, pb) and np.allclose(c, pc) and np.allclose(d, pd)


def test_maximize_interpolant_against_dense_scan(oracle):
    rng = np.random.default_rng(5)
    x = np.linspace(-6, 6, 11)
    for trial in range(200):
        centre = rng.uniform(-7, 7)
        y = -(x - centre) ** 2 * rng.uniform(0.2, 3) + rng.normal(0, 0.3, 11) + 0.01 * (x - centre) ** 3
        got = oracle.maximize_interpolant(x, y)
        assert got == pytest.approx(pyref.maximize_interpolant(x, y), abs=1e-10)
        # dense evaluation of the same spline in the two segments next to the best grid point
        b, c, d = oracle.fmm_spline(x, y)
        k = int(np.argmax(y))
        best, arg = y[k], x[k]
        for seg in (k - 1, k):
            if 0 <= seg < 10:
                t = np.linspace(0, x[seg + 1] - x[seg], 20001)[1:-1]
                v = y[seg] + t * (b[seg] + t * (c[seg] + t * d[seg]))
                if v.max() > best:
                    best, arg = v.max(), x[seg] + t[np.argmax(v)]
        # edgeR only looks at the "sol1" root of the derivative; where it finds an interior maximum it is the dense one
        if got != x[k]:
            assert got == pytest.approx(arg, abs=2e-3)


def test_common_disp_is_argmax_of_summed_apl(oracle):
    rng = np.random.default_rng(6)
    G, n = 400, 6
    y = nb_counts(rng, G, n, phi=0.25)
    X = np.column_stack([np.ones(n), np.tile([0.0, 1.0], 3)])
    common, n_eval = oracle.common_disp(y, X)
    assert 0.15 < common < 0.4 and 8 <= n_eval <= 40
    keep = y.sum(axis=1) >= 5
    lib = y.sum(axis=0)
    delta = np.linspace(max(0.0, common ** 0.25 - 0.05), common ** 0.25 + 0.05, 201)
    f = oracle.apl(y[keep], X, lib, delta ** 4).sum(axis=0)
    assert abs(delta[np.argmax(f)] - common ** 0.25) <= 1e-3        # grid spacing 5e-4, optimize tol 1e-5
    assert common == pytest.approx(pyref.common_disp(y, X), rel=1e-6)


def test_tagwise_disp(oracle):
    rng = np.random.default_rng(7)
    G, n = 60, 8
    y = nb_counts(rng, G, n, phi=0.3)
    y[:3] = 0; y[:3, 0] = [1, 2, 4]                               # rowSums < 5 -> keep the common dispersion
    X = np.column_stack([np.ones(n), np.tile([0.0, 1.0], 4)])
    common, _ = oracle.common_disp(y, X)
    tw = oracle.tagwise_disp(y, X, common)
    assert np.all(tw[:3] == common)
    assert np.all((tw >= common * 2.0 ** -6 * (1 - 1e-12)) & (tw <= common * 2.0 ** 6 * (1 + 1e-12)))
    assert np.allclose(tw, pyref.tagwise_disp(y, X, common), rtol=1e-6)


def test_qr_rank(oracle):
    rng = np.random.default_rng(8)
    X = rng.normal(size=(10, 4))
    assert oracle.qr_rank(X) == 4
    assert oracle.qr_rank(np.column_stack([X, X[:, 0] + X[:, 1]])) == 4
    assert oracle.qr_rank(np.column_stack([np.ones(10), np.ones(10)])) == 1
    assert oracle.qr_rank(np.column_stack([X[:, 0], X[:, 0] * (1 + 1e-9)])) == 1      # tol = 1e-7 as R's qr()


def test_tagwise_prior_df_moderation_two_restatements_agree(oracle):
    """estimateGLMTagwiseDisp(prior.df > 0) — aveLogCPM order, moving-average trend of the APL grid,
    blend (apl + prior.n * trend) / (1 + prior.n).  The C++ oracle against the independent numpy restatement (oracle/pyref.py);
    prior.df = 0 (what ComBat_seq.R:116 passes) must leave the unmoderated result untouched."""
    import pyref
    rng = np.random.default_rng(8)
    G, n = 260, 8
    # well separated abundances: two genes whose aveLogCPM agree to rounding may be ordered differently by two
    # implementations, which moves the edges of every window between them (ill-conditioned in edgeR itself)
    mu = np.exp(np.sort(rng.uniform(4, 9, G)))[:, None] * rng.uniform(0.7, 1.3, (1, n))
    Y = rng.negative_binomial(5, 5 / (5 + mu)).astype(float)
    X = np.column_stack([np.ones(n), np.arange(n) % 2])
    base = oracle.tagwise_disp(Y, X, 0.2, nthreads=2)
    assert np.array_equal(oracle.tagwise_disp(Y, X, 0.2, nthreads=2, prior_df=0.0), base)
    for prior_df in (3.0, 10.0, 50.0):
        a = oracle.tagwise_disp(Y, X, 0.2, nthreads=2, prior_df=prior_df)
        b = pyref.tagwise_disp(Y, X, 0.2, prior_df=prior_df)
        assert np.max(np.abs(a - b) / b) < 1e-8
        assert np.abs(np.log(a) - np.log(base)).mean() > 0.01                   # it does something ...
    strong = oracle.tagwise_disp(Y, X, 0.2, nthreads=2, prior_df=1e4)
    sel = Y.sum(axis=1) >= 5
    assert np.std(np.log(strong[sel])) < 0.5 * np.std(np.log(base[sel]))          # ... and shrinks towards the trend


def test_moving_average_matches_a_direct_window_mean():
    import pyref
    rng = np.random.default_rng(1)
    x = rng.normal(size=(37, 3))
    for width in (1, 2, 5, 8, 37, 50):
        got = pyref.moving_average_by_col(x, width)
        w = min(width, 37)
        h1, h2 = -(-w // 2), w // 2
        want = np.array([x[max(0, r - h1 + 1): min(36, r + h2) + 1].mean(axis=0) for r in range(37)]) if w > 1 else x
        assert np.allclose(got, want, rtol=1e-12, atol=1e-12)
