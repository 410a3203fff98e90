"""not-gpu tier: the C++ oracle's ComBat_seq() (oracle/ref_combat.cpp, restating /root/reference/ComBat_seq.R:22-204
and helper_seq.R:2-64) against the committed golden vectors (tests/golden/pipeline_*.npz, produced by the independent
numpy/scipy restatement oracle/pyref.py) and the structural micro-cases that follow from the reference text alone
(SURVEY.md 8c)."""
import glob
import os

import numpy as np
import pytest

from conftest import rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN = sorted(glob.glob(os.path.join(GOLD, "pipeline_*.npz")))


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[9:-4] for p in GOLDEN])
def test_golden_pipeline(oracle, path):
    g = np.load(path)
    r = oracle.combat_seq(g["counts"].astype(float), g["batch_idx"], g["mod"], shrink=bool(g["shrink"]),
                          shrink_disp=bool(g["shrink_disp"]))
    assert np.array_equal(r["keep"], g["keep"])
    assert np.array_equal(r["lib_size"], g["lib_size"])
    for k in ("disp_common", "phi_hat", "gamma_hat", "gamma_star", "phi_star"):
        assert rel_err(r[k], g[k]) <= 1.0, k                     # |a-b| <= 1e-6 |b| + 1e-9
    assert rel_err(r["mu_hat"], g["mu_hat"]) <= 1.0
    assert np.array_equal(r["adjusted"], g["adjusted"])          # integer output: bit-exact


def test_golden_set_is_complete():
    assert len(GOLDEN) >= 9 and os.path.exists(os.path.join(GOLD, "nb_kat.npz"))


def small_case(synth, pkg, name="C2", **kw):
    case = synth.CASES[name](**kw)
    d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
    return case, d


def test_zero_filtered_genes_pass_through(oracle, synth, pkg):
    case, d = small_case(synth, pkg, G=400)                       # ComBat_seq.R:33-40,198-203
    r = oracle.combat_seq(case.counts.astype(float), d.batch_idx, d.mod)
    dropped = r["keep"] == 0
    assert dropped.any() and not dropped.all()
    allzero_in_some_batch = np.zeros(case.counts.shape[0], bool)
    for b in range(d.batch_idx.max() + 1):
        allzero_in_some_batch |= (case.counts[:, d.batch_idx == b] == 0).all(axis=1)
    assert np.array_equal(dropped, allzero_in_some_batch)
    assert np.array_equal(r["adjusted"][dropped], case.counts[dropped])
    assert np.isnan(r["gamma_hat"][dropped]).all() and not np.isnan(r["gamma_hat"][~dropped]).any()
    assert np.array_equal(r["lib_size"], case.counts[~dropped].sum(axis=0))


def test_counts_at_most_one_pass_through(oracle, synth, pkg):
    case, d = small_case(synth, pkg, G=400)                       # helper_seq.R:50-51
    r = oracle.combat_seq(case.counts.astype(float), d.batch_idx, d.mod)
    m = case.counts <= 1
    assert m.sum() > 50
    assert np.array_equal(r["adjusted"][m], case.counts[m])
    assert (r["adjusted"] >= 0).all() and np.array_equal(r["adjusted"], np.floor(r["adjusted"]))


def test_match_quantiles_identity_and_saturation(oracle):
    rng = np.random.default_rng(0)
    G, n = 500, 6
    mu = np.exp(rng.normal(4, 1.5, (G, n)))
    phi = rng.uniform(0.01, 1.0, G)
    x = rng.poisson(mu).astype(float)
    # same distribution on both sides: identity (needs the 64 eps fuzz of qnbinom)
    assert np.array_equal(oracle.match_quantiles(x, mu, phi, mu, phi), x)
    # |p - 1| < 1e-4  =>  count kept  (helper_seq.R:54-56)
    xs = np.full((1, 1), 5000.0)
    assert oracle.match_quantiles(xs, np.full((1, 1), 10.0), np.array([0.1]), np.full((1, 1), 1000.0), np.array([0.1]))[0, 0] == 5000.0
    # larger mean -> larger counts, monotone in the count
    x2 = np.arange(2, 200, dtype=float).reshape(-1, 1)
    out = oracle.match_quantiles(x2, np.full_like(x2, 50.0), np.full(x2.shape[0], 0.2), np.full_like(x2, 80.0), np.full(x2.shape[0], 0.2))
    sat = np.abs(oracle.nb_cdf(x2[:, 0] - 1, 5.0, 50.0) - 1) < 1e-4     # the reference keeps these counts (not monotone!)
    assert sat.any() and np.array_equal(out[sat, 0], x2[sat, 0])
    assert np.all(np.diff(out[~sat, 0]) >= 0) and np.all(out[~sat, 0] >= x2[~sat, 0])


def test_shrink_options(oracle, synth, pkg):
    case, d = small_case(synth, pkg, "C4", G=120, n=5)
    x = case.counts.astype(float)
    off = oracle.combat_seq(x, d.batch_idx, d.mod, shrink=False)
    assert np.array_equal(off["gamma_star"], off["gamma_hat"], equal_nan=True)          # ComBat_seq.R:166-170
    assert np.array_equal(off["phi_star"], off["phi_hat"], equal_nan=True)
    mean_only = oracle.combat_seq(x, d.batch_idx, d.mod, shrink=True, shrink_disp=False)
    assert np.array_equal(mean_only["phi_star"], mean_only["phi_hat"], equal_nan=True)  # :162-165
    assert not np.array_equal(mean_only["gamma_star"], mean_only["gamma_hat"], equal_nan=True)
    both = oracle.combat_seq(x, d.batch_idx, d.mod, shrink=True, shrink_disp=True)
    assert np.array_equal(both["gamma_star"], mean_only["gamma_star"], equal_nan=True)
    assert not np.array_equal(both["phi_star"], both["phi_hat"], equal_nan=True)
    # posterior means are convex combinations of the other genes' estimates
    k = both["keep"] == 1
    for b in range(4):
        assert np.nanmin(both["gamma_star"][k, b]) >= np.nanmin(both["gamma_hat"][k, b]) - 1e-12
        assert np.nanmax(both["gamma_star"][k, b]) <= np.nanmax(both["gamma_hat"][k, b]) + 1e-12


def test_mc_fallback_when_every_likelihood_underflows(oracle):
    rng = np.random.default_rng(1)                                 # helper_seq.R:28-31
    G, n = 40, 400
    mu = np.exp(rng.normal(6, 1, (G, 1))) * np.ones((1, n))
    phi = rng.uniform(0.1, 0.5, G)
    x = rng.poisson(rng.gamma(1 / phi[:, None], mu * phi[:, None])).astype(float)
    gamma = rng.normal(size=G)
    gs, ps, nfb = oracle.mc_int(x, mu, gamma, phi, mode=0)
    assert nfb == G and np.array_equal(gs, gamma) and np.array_equal(ps, phi)
    gs1, ps1, nfb1 = oracle.mc_int(x, mu, gamma, phi, mode=1)      # log-sum-exp mode never falls back
    assert nfb1 == 0 and not np.array_equal(gs1, gamma)


def test_mc_excludes_self_and_weights_by_likelihood(oracle):
    rng = np.random.default_rng(2)
    G, n = 6, 4
    mu = np.exp(rng.normal(3, 1, (G, n)))
    phi = rng.uniform(0.1, 0.5, G)
    x = rng.poisson(mu).astype(float)
    gamma = rng.normal(size=G)
    from scipy import stats
    gs, ps, _ = oracle.mc_int(x, mu, gamma, phi)
    for i in range(G):
        w = np.array([0.0 if j == i else np.prod(stats.nbinom.pmf(x[i], 1 / phi[j], 1 / (1 + phi[j] * mu[j]))) for j in range(G)])
        assert gs[i] == pytest.approx((w * gamma).sum() / w.sum(), rel=1e-9)
        assert ps[i] == pytest.approx((w * phi).sum() / w.sum(), rel=1e-9)


def test_error_codes_and_reference_messages(oracle):
    rng = np.random.default_rng(0)
    x = rng.negative_binomial(10, 0.1, size=(20, 6)).astype(float)
    ones = np.ones((6, 1))
    with pytest.raises(oracle.OracleError, match="doesn't support 1 sample per batch") as e:      # ComBat_seq.R:29-31
        oracle.combat_seq(x, [0, 0, 0, 1, 1, 2], ones)
    assert e.value.code == 2
    conf = np.column_stack([np.ones(6), [0, 0, 0, 1, 1, 1]])
    with pytest.raises(oracle.OracleError, match="The covariate is confounded with batch") as e:   # :86
        oracle.combat_seq(x, [0, 0, 0, 1, 1, 1], conf)
    assert e.value.code == 3
    two = np.column_stack([np.ones(6), [0, 1, 0, 1, 0, 1], [0, 1, 0, 1, 0, 1]])
    with pytest.raises(oracle.OracleError, match="The covariates are confounded") as e:           # :88
        oracle.combat_seq(x, [0, 0, 0, 1, 1, 1], two)
    assert e.value.code == 4
    mix = np.column_stack([np.ones(6), [0, 1, 0, 1, 0, 1], [0, 0, 0, 1, 1, 1]])
    with pytest.raises(oracle.OracleError, match="At least one covariate is confounded with batch") as e:   # :89
        oracle.combat_seq(x, [0, 0, 0, 1, 1, 1], mix)
    assert e.value.code == 5


def test_threads_do_not_change_results(oracle, synth, pkg):
    case, d = small_case(synth, pkg, "C5", G=150, n=4, B=3)
    x = case.counts.astype(float)
    a = oracle.combat_seq(x, d.batch_idx, d.mod, shrink=True, shrink_disp=True, nthreads=1)
    b = oracle.combat_seq(x, d.batch_idx, d.mod, shrink=True, shrink_disp=True, nthreads=4)
    for k in ("adjusted", "gamma_hat", "phi_hat", "gamma_star", "phi_star", "mu_hat"):
        assert np.array_equal(a[k], b[k], equal_nan=True), k
