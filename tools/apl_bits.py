"""Is the fast APL path (lgamma table + hoisted invariants) bit-identical to the per-sample calls?  And how does the
common dispersion of the C3b batches compare with the oracle either way?"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg = g.load_package(); L = pkg._lib; orc = g.load_oracle()
synth = importlib.import_module("combat-seq_b200.synth")
case = synth.c3(G=2000, crossed=True)
d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
x = case.counts
keep = np.ones(len(x), bool)
for b in range(3):
    keep &= (x[:, d.batch_idx == b] != 0).any(1)
Y = x[keep]; lib = Y.sum(0).astype(float)
for b in range(3):
    cols = np.flatnonzero(d.batch_idx == b)
    phis = np.array([0.05, 0.2, 0.44713841, 0.44715519, 1.0])
    out = {}
    for fast in (0, 1):
        L.set_tuning("apl_fast", fast)
        out[fast] = (L.apl(Y[:, cols], d.mod[cols], lib[cols], phis), L.common_disp(Y[:, cols], d.mod[cols]))
    ref_apl = orc.apl(Y[:, cols].astype(float), d.mod[cols], lib[cols], phis)
    ref_c, _ = orc.common_disp(Y[:, cols].astype(float), d.mod[cols])
    a0, a1 = out[0][0], out[1][0]
    print("batch %d: apl fast==slow bitwise %s (max |diff| %.3g); max rel err vs oracle slow %.3g fast %.3g; column sums slow %s fast %s oracle %s"
          % (b, np.array_equal(a0, a1), np.max(np.abs(a0 - a1)), np.max(np.abs(a0 - ref_apl) / np.abs(ref_apl)),
             np.max(np.abs(a1 - ref_apl) / np.abs(ref_apl)), a0.sum(0)[2:4], a1.sum(0)[2:4], ref_apl.sum(0)[2:4]))
    print("   common disp: slow %.12g fast %.12g oracle %.12g" % (out[0][1], out[1][1], ref_c))
