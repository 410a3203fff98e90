"""-m gpu: parity at the shapes that are benchmarked and at the kernel variants the small cases never reach.

C5's real shape is p = 11 design columns (k_glm_fit_lanes<12>) and n_b = 100 samples per batch (contraction length KT = 133,
28-32 Chebyshev terms, screened source-tile lists); the other cases cover k_glm_fit_lanes<16> (12 < p <= 16), the
one-thread-per-gene k_glm_fit (p > 16), C4 at n_b = 50, C2 at its full 20 000 x 24, gene.subset.n against the oracle's
restatement of the engine's sampler, the screened band against the unscreened contraction, very wide batches in the
exact pairwise kernel, input validation at the C ABI and single-process multi-GPU.
Bars as in test_gpu_pipeline.py (north_star): adjusted counts >= 99.99 % exact, parameters within 1e-6 relative.
"""
import os
import zlib

import numpy as np
import pytest

from conftest import rel_err
from test_gpu_pipeline import check_parity, run_both

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def run_vs_golden(pkg, case, tag):
    """The CUDA pipeline against the oracle's stored outputs (tests/golden/make_golden_oracle.py) on regenerated inputs."""
    g = np.load(os.path.join(GOLDEN, "oracle_%s.npz" % tag))
    assert zlib.crc32(np.ascontiguousarray(case.counts, dtype=np.int32).tobytes()) == int(g["input_crc32"]), \
        "synthetic inputs changed: regenerate the golden file"
    res = pkg.ComBat_seq(case.counts, case.batch, return_details=True, verbose=False, **case.kwargs())
    ref = {k: g[k] for k in g.files}
    ref["adjusted"] = ref["adjusted"].astype(np.float64)
    res = dict(res); res["mu_hat"] = res["mu_hat"][::8]; ref["mu_hat"] = ref["mu_hat_sub"]      # MU_STRIDE of the generator
    return res, ref


def test_c5_shape_p11_nb100(pkg, synth):
    # the "1 500 x 1000" case: B = 10, p = 11 -> k_glm_fit_lanes<12>; n_b = 100 -> KT = 133 (oracle cost is quadratic in G only)
    case = synth.c5(G=1500, n=100, B=10)
    pkg._lib.get_diag("reset")
    res, ref = run_vs_golden(pkg, case, "C5shape_1500x1000")
    exact = check_parity(res, ref)
    print("C5-shape exact-match %.6f, chebyshev terms %d, tail %.2e, tiles %d of %d" % (
        exact, pkg._lib.get_diag("mc_terms"), pkg._lib.get_diag("mc_tail"),
        pkg._lib.get_diag("mc_tiles_done"), pkg._lib.get_diag("mc_tiles_all")))
    assert res["design"].mod.shape[1] + 10 - 1 == 11
    assert np.array_equal(res["n_mc_fallback"], ref["n_mc_fallback"])


def test_glm_lanes16_13_batches(pkg, oracle, synth):
    # 12 < p <= 16: 13 batches + group = 14 design columns -> k_glm_fit_lanes<16>
    case = synth.c5(G=400, n=6, B=13)
    res, ref = run_both(pkg, oracle, case)
    check_parity(res, ref)


def test_glm_one_thread_per_gene_19_batches(pkg, oracle, synth):
    # p > 16: 19 batches + group = 20 design columns -> k_glm_fit (one thread per gene)
    case = synth.c5(G=300, n=4, B=19)
    res, ref = run_both(pkg, oracle, case)
    check_parity(res, ref)


def test_c4_nb50(pkg, synth):
    case = synth.c4(G=1200, n=50)
    res, ref = run_vs_golden(pkg, case, "C4_1200_nb50")
    check_parity(res, ref)
    assert np.array_equal(res["n_mc_fallback"], ref["n_mc_fallback"])


def test_c2_full_size(pkg, oracle, synth):
    case = synth.c2()                                            # 20 000 x 24, 3 batches, shrink = FALSE
    res, ref = run_both(pkg, oracle, case)
    exact = check_parity(res, ref)
    print("C2 full size: exact-match %.6f over %d elements" % (exact, case.counts.size))


def test_gene_subset_n_matches_the_oracle_sampler(pkg, oracle, synth):
    # helper_seq.R:17-21 with the engine's sampler (keyed Feistel permutation) restated in the oracle: same subsets,
    # so gamma_star / phi_star must agree at the 1e-6 bar (R's own sample() stream is not reproduced by either)
    from test_gpu_mc import make_batch
    x, mu, gamma, phi = make_batch(700, 8, seed=21)
    for mode in (0, 1):
        gs, ps, nfb = pkg._lib.mc_int(x, mu, gamma, phi, mode=mode, gene_subset_n=120, seed=5)
        gs_o, ps_o, nfb_o = oracle.mc_int(x.astype(float), mu, gamma, phi, mode=mode, gene_subset_n=120, seed=5,
                                          nthreads=oracle.max_threads())
        assert nfb == nfb_o
        assert rel_err(gs, gs_o) <= 1.0 and rel_err(ps, ps_o) <= 1.0
    case = synth.c4(G=500, n=6)
    d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
    res = pkg.ComBat_seq(case.counts, case.batch, group=case.group, shrink=True, gene_subset_n=80, seed=3,
                         return_details=True, verbose=False)
    ref = oracle.combat_seq(case.counts.astype(np.float64), d.batch_idx, d.mod, shrink=True, gene_subset_n=80, seed=3,
                            nthreads=oracle.max_threads())
    check_parity(res, ref)
    with pytest.raises(ValueError, match="larger than the population"):
        pkg.ComBat_seq(case.counts[:60], case.batch, group=case.group, shrink=True, gene_subset_n=500, verbose=False)


def test_screened_band_equals_all_tiles(pkg, synth):
    # the tcgen05 screening pass only removes source tiles whose pairs are all below e^-40 of the row's best weight:
    # results with and without it agree far inside the parity bar, the fallback sets are identical
    from test_gpu_mc import make_batch
    L = pkg._lib
    for (G, n, mode) in ((3000, 40, 1), (2500, 100, 0), (1000, 5, 1)):
        x, mu, gamma, phi = make_batch(G, n, seed=G + n, phi_lo=0.02, phi_hi=2.0)
        try:
            L.set_tuning("mc_algo", 2)
            L.set_tuning("mc_band", 0)
            a = L.mc_int(x, mu, gamma, phi, mode=mode)
            L.set_tuning("mc_band", 1)
            L.get_diag("reset")
            b = L.mc_int(x, mu, gamma, phi, mode=mode)
            done, total = L.get_diag("mc_tiles_done"), L.get_diag("mc_tiles_all")
        finally:
            L.set_tuning("mc_algo", 0); L.set_tuning("mc_band", 1)
        print("G=%d n=%d: source tiles visited %d of %d" % (G, n, done, total))
        assert a[2] == b[2]
        assert rel_err(b[0], a[0], rtol=1e-9, atol=1e-12) <= 1.0 and rel_err(b[1], a[1], rtol=1e-9, atol=1e-12) <= 1.0
        assert 0 < done <= total


def test_direct_kernel_very_wide_batch(pkg, oracle, synth):
    # one batch of 900 samples: the exact pairwise kernel (gene.subset.n, and mc_algo = 1) beyond 64 n bytes of shared memory
    counts, batch, group = synth.simulate(48, [900, 6], (1, 1.5), (0.2, 0.3), seed=17)
    d = pkg.design.build_design(counts.shape[1], batch, group, None, True)
    L = pkg._lib
    cols = np.flatnonzero(d.batch_idx == 0)
    rng = np.random.default_rng(1)
    x = counts[:, cols]
    mu = np.maximum(x.mean(axis=1, keepdims=True), 0.5) * rng.uniform(0.8, 1.2, x.shape)
    phi = rng.uniform(0.05, 0.5, len(x)); gamma = rng.normal(0, 0.3, len(x))
    try:
        L.set_tuning("mc_algo", 1)
        gs, ps, nfb = L.mc_int(x, mu, gamma, phi, mode=1)
        gs2, ps2, nfb2 = L.mc_int(x, mu, gamma, phi, mode=1, gene_subset_n=20, seed=2)
    finally:
        L.set_tuning("mc_algo", 0)
    gs_o, ps_o, _ = oracle.mc_int(x.astype(float), mu, gamma, phi, mode=1, nthreads=oracle.max_threads())
    gs2_o, ps2_o, _ = oracle.mc_int(x.astype(float), mu, gamma, phi, mode=1, gene_subset_n=20, seed=2, nthreads=oracle.max_threads())
    assert rel_err(gs, gs_o) <= 1.0 and rel_err(ps, ps_o) <= 1.0
    assert rel_err(gs2, gs2_o) <= 1.0 and rel_err(ps2, ps2_o) <= 1.0


def test_c_abi_rejects_unrepresentable_counts(pkg):
    rng = np.random.default_rng(0)
    base = rng.negative_binomial(10, 0.1, size=(40, 8)).astype(np.float64)
    d = pkg.design.build_design(8, [1, 1, 1, 1, 2, 2, 2, 2])
    for bad_value, text in ((np.nan, "NA counts not allowed"), (-3.0, "Negative counts not allowed"),
                            (2.5, "non-integer"), (3e9, "2\\^31")):
        x = base.copy(); x[7, 3] = bad_value
        with pytest.raises(pkg._lib.CombatSeqError, match=text) as e:
            pkg._lib.combat_seq_raw(x, d.batch_idx, d.mod)
        assert e.value.code == pkg._lib.CBS_EINVAL
    xi = base.astype(np.int32); xi[0, 0] = -1
    with pytest.raises(pkg._lib.CombatSeqError, match="Negative counts"):
        pkg._lib.combat_seq_raw(xi, d.batch_idx, d.mod)


def test_single_process_multi_gpu_is_bit_identical(pkg, synth):
    # cbs_input.n_gpus: one host thread per device inside one call (what the R .Call reaches); every output must equal
    # the one-device run bit for bit.  Gene counts give every rank work (and a ragged last shard).
    ndev = pkg.device_count()
    if ndev < 2:
        pytest.skip("needs at least 2 visible GPUs")
    L = pkg._lib
    want = ("adjusted", "gamma_hat", "phi_hat", "gamma_star", "phi_star", "mu_hat")
    for case in (synth.c5(G=4000, n=20, B=4), synth.c2(G=5000), synth.c3(G=3000, crossed=True)):
        d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
        one = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=case.shrink, shrink_disp=case.shrink_disp, want=want, n_gpus=1)
        for n in sorted({2, min(ndev, 4), ndev}):
            many = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=case.shrink, shrink_disp=case.shrink_disp, want=want, n_gpus=n)
            for k in want + ("keep", "lib_size", "disp_common", "n_mc_fallback"):
                assert np.array_equal(one[k], many[k], equal_nan=True), "%s differs at n_gpus=%d (%s)" % (k, n, case.name)
            assert (many["row_begin"], many["row_end"]) == (0, case.counts.shape[0])
    # big pageable float64 buffers (what R hands over): every device thread stages its rows through pinned memory
    # (host_bridge.hpp), in and out; 4 600 x 1000 doubles = 36.8 MB, 18 MB per device at n_gpus = 2
    case = synth.c5(G=4600, n=100, B=10)
    d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
    page_in = np.asfortranarray(case.counts.astype(np.float64))
    outs = []
    for n in (1, 2):
        page_out = np.full(case.counts.shape, -1.0, order="F")
        L.combat_seq_into(page_in, d.batch_idx, d.mod, page_out, shrink=False, shrink_disp=False, n_gpus=n)
        outs.append(page_out)
    assert np.array_equal(outs[0], outs[1]) and outs[0].min() >= 0


def test_tagwise_prior_df_moderation(pkg, oracle, synth):
    # estimateGLMTagwiseDisp(prior.df > 0) on the device (aveLogCPM kernel, radix sort, prefix-sum moving
    # average, blend) against the oracle; prior.df = 0 is the unmoderated path ComBat_seq() uses.
    # (abundances well separated: genes whose aveLogCPM agree to rounding can be ordered differently by two implementations,
    # which moves window edges: an ill-conditioning of the moderation itself, see tests/test_oracle_edger.py)
    counts, batch, group = synth.simulate(1500, [8, 8], (1, 1.5), (0.15, 0.3), seed=31, low_frac=0.0, base_logmean=6.5, base_logsd=1.2)
    d = pkg.design.build_design(counts.shape[1], batch, group, None, True)
    cols = np.flatnonzero(d.batch_idx == 1)
    Y, X = counts[:, cols], d.mod[cols]
    base = pkg._lib.tagwise_disp(Y, X, 0.3)
    for prior_df in (5.0, 20.0):
        got = pkg._lib.tagwise_disp(Y, X, 0.3, prior_df=prior_df)
        ref = oracle.tagwise_disp(Y.astype(float), X, 0.3, nthreads=oracle.max_threads(), prior_df=prior_df)
        assert rel_err(got, ref) <= 1.0
        assert not np.allclose(got, base)


def test_glm_wide_second_pass_agrees_with_the_single_pass(pkg, synth):
    # fits that need more than 9 sweeps are redone by a warp per gene (k_glm_fit_wide): the sums associate differently, so
    # the coefficients of those genes may differ in the last places; everything else is bit-identical
    L = pkg._lib
    case = synth.c5(G=6000, n=40, B=4)
    d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
    want = ("adjusted", "gamma_hat", "mu_hat")
    try:
        L.set_tuning("glm_two_pass", 0)
        one = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, want=want)
        L.set_tuning("glm_two_pass", 1)
        L.set_tuning("glm_hist", 1)
        two = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, want=want)
        long_fits = sum(L.get_diag("glm_iter_%d" % k) for k in range(10, 256))
    finally:
        L.set_tuning("glm_two_pass", 1); L.set_tuning("glm_hist", 0)
    print("fits with 10 or more sweeps (second pass):", int(long_fits))
    assert rel_err(two["gamma_hat"], one["gamma_hat"], rtol=1e-9, atol=1e-12) <= 1.0
    assert rel_err(two["mu_hat"], one["mu_hat"], rtol=1e-9, atol=1e-12) <= 1.0
    assert np.array_equal(two["adjusted"], one["adjusted"])
    same = np.all((two["gamma_hat"] == one["gamma_hat"]) | np.isnan(one["gamma_hat"]), axis=1)
    assert same.mean() > 0.99


def test_pageable_buffers_take_the_staged_copies_and_give_the_same_matrix(pkg, synth):
    """Caller buffers of 16 MB and more in pageable memory (an R matrix) are copied by worker threads through pinned staging
    (host_bridge.hpp), pinned ones directly: same bits either way."""
    torch = pytest.importorskip("torch")
    L = pkg._lib
    case = synth.CASES["C5"](G=2200, n=100, B=10)                       # 2200 x 1000 doubles = 17.6 MB in, 17.6 MB out
    d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
    G0, N = case.counts.shape
    page_in = np.asfortranarray(case.counts.astype(np.float64))
    page_out = np.full((G0, N), -1.0, order="F")
    L.combat_seq_into(page_in, d.batch_idx, d.mod, page_out, shrink=False, shrink_disp=False)
    pin_in = torch.empty((N, G0), dtype=torch.float64, pin_memory=True)
    pin_in.numpy()[:] = case.counts.T
    pin_out = torch.full((N, G0), -1.0, dtype=torch.float64, pin_memory=True)
    L.combat_seq_into(pin_in.numpy().T, d.batch_idx, d.mod, pin_out.numpy().T, shrink=False, shrink_disp=False)
    assert np.array_equal(page_out, pin_out.numpy().T)
    assert page_out.min() >= 0 and not np.array_equal(page_out, page_in)
