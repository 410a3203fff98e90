"""List the genes whose GPU APL differs from the oracle's (C3b batch 3) and dump their inputs for offline study."""
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
b = 2
cols = np.flatnonzero(d.batch_idx == b)
phis = np.array([0.2, 0.44713841, 0.44715519, 0.7])
a = L.apl(Y[:, cols], d.mod[cols], lib[cols], phis)
r = orc.apl(Y[:, cols].astype(float), d.mod[cols], lib[cols], phis)
rel = np.abs(a - r) / np.abs(r)
bad = np.argwhere(rel > 1e-10)
print("genes x evals with rel diff > 1e-10:", len(bad))
for gi, e in bad[:20]:
    print("gene", gi, "phi", phis[e], "gpu %.15g oracle %.15g rel %.3g" % (a[gi, e], r[gi, e], rel[gi, e]), "counts", Y[gi, cols].tolist())
coef_g, mu_g = L.glm_fit(Y[:, cols], d.mod[cols], np.full((len(Y), 1), phis[1]), None, lib[cols])
coef_o, mu_o = orc.glm_fit(Y[:, cols].astype(float), d.mod[cols], np.full((len(Y), 1), phis[1]), None, lib[cols])
rc = np.abs(coef_g - coef_o).max(1)
for gi in np.argsort(-rc)[:5]:
    print("glm gene", gi, "max |coef diff| %.3g" % rc[gi], "gpu", coef_g[gi], "oracle", coef_o[gi])
np.savez("gpurun_out/apl_diff_inputs.npz", Y=Y[:, cols], mod=d.mod[cols], lib=lib[cols], phis=phis, apl_gpu=a)
