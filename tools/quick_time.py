"""Ad-hoc stage timing of one config on the GPU (development helper, not the bench)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as g
pkg = g.load_package()
import importlib
synth = importlib.import_module("combat-seq_b200.synth")
name = sys.argv[1]
kw = {}
for a in sys.argv[2:]:
    k, v = a.split("="); kw[k] = int(v)
case = synth.CASES[name](**kw)
d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
for it in range(2):
    t = time.time()
    res = pkg._lib.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=case.shrink, shrink_disp=case.shrink_disp)
    dt = time.time() - t
    print(name, kw, case.counts.shape, "wall %.3fs" % dt, {k: round(v, 1) for k, v in res["stage_ms"].items()},
          "fallback", res["n_mc_fallback"].tolist(), "disp", np.round(res["disp_common"], 4).tolist(), flush=True)
