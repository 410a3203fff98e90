"""Histogram of Levenberg iterations per gene over the GLM launches of one ComBat_seq call (development helper).
usage: glm_iters.py CASE [G=..] [n=..] [B=..]"""
import sys, os, importlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as g
pkg = g.load_package(); L = pkg._lib
synth = importlib.import_module("combat-seq_b200.synth")
kw = {}
for a in sys.argv[2:]:
    k, v = a.split("="); kw[k] = int(v)
case = synth.CASES[sys.argv[1]](**kw)
d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=False)
L.set_tuning("glm_hist", 1)
res = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=False)
h = np.array([L.get_diag("glm_iter_%d" % k) for k in range(256)])
L.set_tuning("glm_hist", 0)
tot = h.sum()
print("stage ms", {k: round(v, 2) for k, v in res["stage_ms"].items()})
print("fits", int(tot), "mean iterations %.2f" % ((h * np.arange(256)).sum() / max(tot, 1)))
cum = np.cumsum(h[::-1])[::-1]
for k in (0, 1, 2, 4, 6, 8, 10, 12, 16, 20, 24, 32, 48, 64, 100, 150, 200, 250, 255):
    print("  iterations >= %3d: %d fits (%.4f %%)" % (k, cum[k], 100.0 * cum[k] / max(tot, 1)))
print("nonzero bins", {int(k): int(v) for k, v in enumerate(h) if v})
