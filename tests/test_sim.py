"""Synthetic count generator (generator half; cbs_simulate_counts, combat-seq_b200/csrc/rng.cuh).

The reference draws its simulated matrices with polyester::simulate_experiment = rnbinom(mu = reads * fold change, size)
(/root/reference/simulations/sim_DEpipe.R:98-113).  There is no bit-level oracle for a random generator: what is checked is
the DISTRIBUTION (against scipy's negative binomial: chi-square on binned counts, mean, variance, zero fraction), the
Philox4x32-10 block function against its published known-answer vectors, determinism, and that the device produces the
host-compiled sampler's draws.  not-gpu tier: the header compiled for the host; gpu tier: the kernel through the C ABI."""
import ctypes as C

import numpy as np
import pytest
from scipy import stats

f64p = C.POINTER(C.c_double)
i32p = C.POINTER(C.c_int32)


def host_sample(lib, seed, index0, mu, phi):
    mu = np.ascontiguousarray(mu, dtype=np.float64); phi = np.ascontiguousarray(np.broadcast_to(phi, mu.shape), dtype=np.float64)
    out = np.empty(mu.size, dtype=np.int32)
    lib.hm_sample_nb(C.c_int64(mu.size), C.c_uint64(seed), C.c_uint64(index0), mu.ctypes.data_as(f64p), phi.ctypes.data_as(f64p),
                     out.ctypes.data_as(i32p))
    return out


def chi2_against_nb(x, mu, phi):
    """p-value of a chi-square goodness-of-fit test of the draws x against NB(mu, phi) (Poisson when phi == 0)."""
    dist = stats.nbinom(1.0 / phi, 1.0 / (1.0 + mu * phi)) if phi > 0 else stats.poisson(mu)
    edges = np.unique(dist.ppf(np.linspace(0, 1, 41)[1:-1]).astype(int))           # about 40 bins of equal probability
    cdf = np.concatenate([[0.0], dist.cdf(edges), [1.0]])
    expected = np.diff(cdf) * x.size
    keep = expected > 5
    observed = np.histogram(x, bins=np.concatenate([[-0.5], edges + 0.5, [np.inf]]))[0]
    # bins with tiny expectation are merged into their neighbour
    obs, exp = [], []
    ao = ae = 0.0
    for o, e in zip(observed, expected):
        ao += o; ae += e
        if ae > 5:
            obs.append(ao); exp.append(ae); ao = ae = 0.0
    if ae > 0:
        obs[-1] += ao; exp[-1] += ae
    assert keep.any()
    return float(stats.chisquare(obs, np.asarray(exp) * (np.sum(obs) / np.sum(exp))).pvalue)


def test_philox_known_answers(hostmath_lib):
    # Random123 kat_vectors: philox4x32-10
    for ctr, key, want in (((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
                           ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
                           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
                            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))):
        c = (C.c_uint32 * 4)(*ctr)
        hostmath_lib.hm_philox(c, C.c_uint32(key[0]), C.c_uint32(key[1]))
        assert tuple(c) == want


@pytest.mark.parametrize("mu,phi", [(0.3, 0.5), (4.0, 0.2), (9.5, 0.0), (30.0, 0.05), (250.0, 0.4), (5000.0, 0.15),
                                    (2e5, 0.02), (60.0, 3.0)])
def test_host_sampler_distribution(hostmath_lib, mu, phi):
    n = 200000
    x = host_sample(hostmath_lib, 1234, 0, np.full(n, mu), phi).astype(np.float64)
    var = mu + phi * mu * mu
    assert abs(x.mean() - mu) < 5 * np.sqrt(var / n)
    assert abs(x.var() - var) < 0.05 * var + 5 * np.sqrt(var / n)
    assert chi2_against_nb(x, mu, phi) > 1e-4
    p0 = (1 + mu * phi) ** (-1 / phi) if phi > 0 else np.exp(-mu)
    assert abs((x == 0).mean() - p0) < 5 * np.sqrt(max(p0 * (1 - p0), 1e-12) / n) + 1e-9


def test_host_sampler_is_counter_based(hostmath_lib):
    mu = np.full(1000, 50.0)
    a = host_sample(hostmath_lib, 7, 0, mu, 0.3)
    assert np.array_equal(a, host_sample(hostmath_lib, 7, 0, mu, 0.3))               # same (seed, index): same draw
    assert np.array_equal(a[100:200], host_sample(hostmath_lib, 7, 100, mu[:100], 0.3))  # any slice on its own
    assert (a != host_sample(hostmath_lib, 8, 0, mu, 0.3)).mean() > 0.9              # another seed: another matrix
    assert host_sample(hostmath_lib, 1, 0, np.array([0.0, -1.0, np.nan, np.inf]), 0.2).tolist() == [0, 0, 0, 0]
    # neighbouring elements are uncorrelated
    x = host_sample(hostmath_lib, 99, 0, np.full(100000, 20.0), 0.2).astype(float)
    assert abs(np.corrcoef(x[:-1], x[1:])[0, 1]) < 0.02


@pytest.mark.gpu
def test_device_generator_matches_host_sampler_and_recipe(pkg, hostmath_lib):
    L = pkg._lib
    rng = np.random.default_rng(3)
    G, N = 3000, 24
    base = np.exp(rng.normal(4.0, 1.5, G))
    colf = rng.uniform(0.7, 1.3, N)
    level = np.arange(N) % 3
    fold = np.column_stack([np.ones(G), np.where(rng.random(G) < 0.5, 1.5, 1 / 1.5), np.where(rng.random(G) < 0.5, 2.0, 0.5)])
    phi = np.array([0.1, 0.2, 0.3])[level]
    x = L.simulate_counts(base, colf, phi, seed=2024, level=level, fold=fold)
    assert x.shape == (G, N) and x.dtype == np.int32 and x.min() >= 0
    assert np.array_equal(x, L.simulate_counts(base, colf, phi, seed=2024, level=level, fold=fold))
    mu = base[:, None] * colf[None, :] * fold[:, level]
    ref = host_sample(hostmath_lib, 2024, 0, mu.ravel(order="F"), np.broadcast_to(phi[None, :], (G, N)).ravel(order="F")).reshape((G, N), order="F")
    # device and host libm differ in the last place of exp / log: an accept / reject decision can flip, rarely
    assert (x == ref).mean() > 0.999
    # the recipe: E x = mu, Var x = mu + phi mu^2 (standardised residuals over the whole matrix)
    z = (x - mu) / np.sqrt(mu + phi[None, :] * mu * mu)
    assert abs(z.mean()) < 5 / np.sqrt(z.size) and abs(z.var() - 1) < 0.05
    # no fold changes / Poisson limit / device-resident output
    y = L.simulate_counts(np.full(5000, 12.0), np.ones(4), np.zeros(4), seed=5)
    assert abs(y.mean() - 12.0) < 0.1 and abs(y.var() - 12.0) < 0.5
    with pytest.raises(L.CombatSeqError):
        L.simulate_counts(base, colf, phi, seed=1, level=np.full(N, 7), fold=fold)


@pytest.mark.gpu
def test_generated_matrix_runs_through_the_pipeline(pkg):
    """The generator feeds the path: a generated 2-batch matrix with gene-wise mean batch effects (half of the genes doubled, half
    halved in batch 2: sim_DEpipe_helpers.R:27-46) and a dispersion batch effect is adjusted; per gene, the batches' library-size
    normalised means differ by a factor 2 before and agree after."""
    L = pkg._lib
    rng = np.random.default_rng(11)
    G, n = 1500, 12
    base = np.exp(rng.normal(5.0, 1.0, G))
    level = np.repeat([0, 1], n)
    fold = np.column_stack([np.ones(G), np.where(rng.random(G) < 0.5, 2.0, 0.5)])
    x = L.simulate_counts(base, np.ones(2 * n), np.array([0.05, 0.2])[level], seed=77, level=level, fold=fold)
    adj = np.asarray(pkg.ComBat_seq(x, level + 1, verbose=False))

    def batch_log_ratio(m):
        cpm = m / m.sum(0, keepdims=True)
        return np.log2((cpm[:, n:].mean(1) + 1e-9) / (cpm[:, :n].mean(1) + 1e-9))
    assert np.median(np.abs(batch_log_ratio(x))) > 0.6
    assert np.median(np.abs(batch_log_ratio(adj))) < 0.1
