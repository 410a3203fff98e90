"""-m gpu: BASELINE.json's full-size configuration C5 (60 000 genes x 1000 samples, 10 batches, shrink + shrink.disp).

The oracle needs days for this size (the Monte-Carlo step is quadratic in the genes), so what is checked here are properties
that do not depend on the size and that the reference's code guarantees (file:line into /root/reference):
* genes dropped by the zero filter come back unchanged (ComBat_seq.R:33-40, 198-203);
* counts of 0 and 1 pass through (helper_seq.R:50-51), adjusted counts are non-negative integers;
* with shrink = FALSE the batches' library-size-normalised gene means agree after the adjustment where they differ by the
  simulated batch effect before (the purpose of the method; ComBat_seq.R:173-196);
* the result does not depend on how the work is laid out: match_quantiles over binned elements vs one warp per gene give
  the same adjusted counts bit for bit;
* every Monte-Carlo posterior mean lies inside the range of the gamma_hat it averages (helper_seq.R:31-33: a weighted mean
  with non-negative weights), no posterior dispersion is negative.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_c5_full_size_properties(pkg, synth):
    L = pkg._lib
    case = synth.c5()                                              # the benchmarked workload itself
    G0, N = case.counts.shape
    assert (G0, N) == (60000, 1000)
    d = pkg.design.build_design(N, case.batch, case.group, case.covar_mod, case.full_mod)
    want = ("adjusted", "gamma_hat", "gamma_star", "phi_hat", "phi_star")
    res = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=True, shrink_disp=True, want=want)
    adj, keep = res["adjusted"], res["keep"].astype(bool)
    x = case.counts
    assert 0 < (~keep).sum() < G0 // 10                             # the simulation has genes that a batch never sees
    assert np.array_equal(adj[~keep], x[~keep].astype(np.float64))
    assert np.array_equal(adj, np.rint(adj)) and adj.min() >= 0 and np.isfinite(adj).all()
    low = x[keep] <= 1
    assert np.array_equal(adj[keep][low], x[keep][low].astype(np.float64))
    print("Monte-Carlo fallbacks per batch (reference-mode underflow, helper_seq.R:34-37):", res["n_mc_fallback"].tolist())
    assert res["n_mc_fallback"].sum() < 0.05 * keep.sum() * 10

    # batch effect before and after, on well-expressed genes: |log2 ratio| of two batches' normalised means.  With
    # shrink = FALSE the adjustment removes it.  With shrink = TRUE it does NOT on this simulation, and that is the reference's
    # algorithm, not this engine: at 100 samples per batch the likelihood products of monte_carlo_int_NB are so peaked that a
    # gene's posterior mean is the gamma_hat of its most similar OTHER gene, whose batch effect has a random direction here (the
    # oracle's stored output for the 1 500-gene case of the same shape shows the same: 0.82 before, 1.31 after).
    expressed = keep & (x.mean(1) > 20)
    b0, b9 = d.batch_idx == 0, d.batch_idx == 9

    def batch_log_ratio(m):
        cpm = m / m.sum(0, keepdims=True)
        return np.log2((cpm[:, b9].mean(1) + 1e-12) / (cpm[:, b0].mean(1) + 1e-12))
    before = np.median(np.abs(batch_log_ratio(x[expressed].astype(np.float64))))
    after_shrink = np.median(np.abs(batch_log_ratio(adj[expressed])))
    plain = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=False, shrink_disp=False)["adjusted"]
    after_plain = np.median(np.abs(batch_log_ratio(plain[expressed])))
    print("median |log2 batch ratio| of expressed genes: %.3f before, %.3f after (shrink = FALSE), %.3f after (shrink = TRUE)"
          % (before, after_plain, after_shrink))
    assert before > 0.8 and after_plain < 0.1 * before
    assert np.array_equal(plain[~keep], x[~keep].astype(np.float64)) and np.array_equal(plain, np.rint(plain)) and plain.min() >= 0

    # parameter identities
    gs, gh = res["gamma_star"][keep], res["gamma_hat"][keep]
    for b in range(gs.shape[1]):
        ok = np.isfinite(gs[:, b])
        assert gs[ok, b].min() >= np.nanmin(gh[:, b]) - 1e-9 and gs[ok, b].max() <= np.nanmax(gh[:, b]) + 1e-9
    ps = res["phi_star"][keep]
    print("posterior dispersions: %d of %d not positive (weights that underflow to denormals: the reference mode keeps R's arithmetic)"
          % (int((ps[np.isfinite(ps)] <= 0).sum()), ps.size))
    assert np.all(ps[np.isfinite(ps)] >= 0) and (ps[np.isfinite(ps)] <= 0).mean() < 0.01

    # layout independence
    try:
        L.set_tuning("mq_bins", 0)
        alt = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=True, shrink_disp=True)["adjusted"]
        assert np.array_equal(alt, adj)
    finally:
        L.set_tuning("mq_bins", 1)
