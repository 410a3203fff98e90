"""-m gpu: the Monte-Carlo integration kernels (helper_seq.R:8-42) through cbs_mc_int.

Two device algorithms exist: the direct pairwise kernel (every lgamma(x + r) evaluated) and the FP64 tensor-core
contraction with the Chebyshev expansion of the lgamma term (mc_gemm.cuh).  Both must agree with the CPU oracle
within the north_star tolerance for gamma/phi (1e-6 relative, stated form |a-b| <= 1e-6*|b| + 1e-9).
"""
import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu


def make_batch(G, n, seed, phi_lo=0.05, phi_hi=1.0, zero_frac=0.0):
    rng = np.random.default_rng(seed)
    base = np.exp(rng.normal(3.5, 1.5, G)).clip(0.3, 5e4)
    mu = base[:, None] * rng.uniform(0.7, 1.3, (1, n)) * np.exp(rng.normal(0, 0.05, (G, n)))
    phi = np.exp(rng.uniform(np.log(phi_lo), np.log(phi_hi), G))
    shape = 1.0 / phi[:, None]
    x = rng.poisson(rng.gamma(shape, mu / shape)).astype(np.int32)
    if zero_frac:
        dead = rng.random(G) < zero_frac
        mu[dead, 0] = 0.0                       # mu = 0 cells (all-zero one-way cells give fitted value 0)
        x[dead, 0] = 0
    gamma = rng.normal(0, 0.5, G)
    return x, mu, gamma, phi


@pytest.fixture
def algo(pkg):
    yield lambda a: pkg._lib.set_tuning("mc_algo", a)
    pkg._lib.set_tuning("mc_algo", 0)


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("G,n", [(700, 6), (1500, 20), (300, 3)])
def test_contraction_matches_direct_and_oracle(pkg, oracle, algo, mode, G, n):
    x, mu, gamma, phi = make_batch(G, n, seed=G + n, zero_frac=0.01)
    algo(1)
    gs_d, ps_d, nfb_d = pkg._lib.mc_int(x, mu, gamma, phi, mode=mode)
    algo(2)
    pkg._lib.get_diag("reset")
    gs_g, ps_g, nfb_g = pkg._lib.mc_int(x, mu, gamma, phi, mode=mode)
    tail, terms = pkg._lib.get_diag("mc_tail"), pkg._lib.get_diag("mc_terms")
    print("chebyshev terms %d, tail estimate %.2e" % (terms, tail))
    assert terms >= 8 and tail < 1e-8
    gs_o, ps_o, nfb_o = oracle.mc_int(x.astype(float), mu, gamma, phi, mode=mode, nthreads=oracle.max_threads())
    assert nfb_d == nfb_o == nfb_g
    for a, b in ((gs_g, gs_d), (ps_g, ps_d), (gs_g, gs_o), (ps_g, ps_o), (gs_d, gs_o)):
        assert rel_err(a, b) <= 1.0


def test_contraction_wide_dispersion_range(pkg, oracle, algo):
    # phi spans 2^-6 .. 2^6 around 0.3 (the whole range edgeR's tagwise grid can produce)
    x, mu, gamma, phi = make_batch(900, 8, seed=7, phi_lo=0.3 / 64, phi_hi=0.3 * 64)
    algo(2)
    pkg._lib.get_diag("reset")
    gs_g, ps_g, _ = pkg._lib.mc_int(x, mu, gamma, phi, mode=1)
    print("terms", pkg._lib.get_diag("mc_terms"), "tail", pkg._lib.get_diag("mc_tail"))
    assert pkg._lib.get_diag("mc_tail") < 1e-8
    gs_o, ps_o, _ = oracle.mc_int(x.astype(float), mu, gamma, phi, mode=1, nthreads=oracle.max_threads())
    assert rel_err(gs_g, gs_o) <= 1.0 and rel_err(ps_g, ps_o) <= 1.0


def test_contraction_constant_dispersion(pkg, oracle, algo):
    # fallback batches carry one common phi for every gene (ComBat_seq.R:113): degenerate range of r
    x, mu, gamma, phi = make_batch(500, 10, seed=11)
    phi[:] = 0.2
    algo(2)
    gs_g, ps_g, _ = pkg._lib.mc_int(x, mu, gamma, phi, mode=1)
    gs_o, ps_o, _ = oracle.mc_int(x.astype(float), mu, gamma, phi, mode=1, nthreads=oracle.max_threads())
    assert rel_err(gs_g, gs_o) <= 1.0 and rel_err(ps_g, ps_o) <= 1.0


def test_large_counts(pkg, oracle, algo):
    # counts up to ~1e6: the lgamma difference must not lose digits to cancellation
    rng = np.random.default_rng(3)
    G, n = 400, 5
    base = np.exp(rng.uniform(np.log(1e3), np.log(1e6), G))
    mu = base[:, None] * rng.uniform(0.9, 1.1, (G, n))
    phi = rng.uniform(0.01, 0.2, G)
    x = rng.poisson(rng.gamma(1 / phi[:, None], mu * phi[:, None])).astype(np.int32)
    gamma = rng.normal(0, 0.5, G)
    algo(2)
    gs_g, ps_g, _ = pkg._lib.mc_int(x, mu, gamma, phi, mode=1)
    gs_o, ps_o, _ = oracle.mc_int(x.astype(float), mu, gamma, phi, mode=1, nthreads=oracle.max_threads())
    assert rel_err(gs_g, gs_o) <= 1.0 and rel_err(ps_g, ps_o) <= 1.0


def test_gene_subset_uses_direct_kernel(pkg, algo):
    # gene.subset.n (helper_seq.R:17-21): same seed => same answer; a subset of everything == no subset
    x, mu, gamma, phi = make_batch(300, 6, seed=5)
    a = pkg._lib.mc_int(x, mu, gamma, phi, mode=1, gene_subset_n=50, seed=9)
    b = pkg._lib.mc_int(x, mu, gamma, phi, mode=1, gene_subset_n=50, seed=9)
    c = pkg._lib.mc_int(x, mu, gamma, phi, mode=1, gene_subset_n=50, seed=10)
    assert np.array_equal(a[0], b[0]) and not np.array_equal(a[0], c[0])
    full = pkg._lib.mc_int(x, mu, gamma, phi, mode=1)
    allg = pkg._lib.mc_int(x, mu, gamma, phi, mode=1, gene_subset_n=299, seed=1)
    assert rel_err(allg[0], full[0]) <= 1.0
