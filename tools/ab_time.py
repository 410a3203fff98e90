"""A/B of the tuning knobs on one config (development helper): stage times per variant and equality of the results."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as g
pkg = g.load_package()
import importlib
synth = importlib.import_module("combat-seq_b200.synth")
L = pkg._lib
name = sys.argv[1] if len(sys.argv) > 1 else "C5"
case = synth.CASES[name]()
d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
variants = [("base: glm local, mq tiles", {"glm_scr": 0, "mq_rows": 1}),
            ("mq tiles/out-of-line", {"glm_scr": 0, "mq_rows": 4}),
            ("glm smem scratch", {"glm_scr": 1, "mq_rows": 1}),
            ("mq warp/inlined", {"glm_scr": 1, "mq_rows": 2}),
            ("mq warp/out-of-line", {"glm_scr": 1, "mq_rows": 3})]
ref = None
for label, knobs in variants:
    for k, v in knobs.items():
        L.set_tuning(k, v)
    best = None
    for it in range(3):
        res = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=case.shrink, shrink_disp=case.shrink_disp,
                               want=("adjusted", "gamma_hat", "phi_hat", "gamma_star"))
        st = res["stage_ms"]
        if best is None or st["total"] < best["total"]:
            best = dict(st)
    same = ""
    if ref is None:
        ref = res
    else:
        same = ", ".join("%s identical %s" % (k, np.array_equal(ref[k], res[k], equal_nan=True))
                         for k in ("adjusted", "gamma_hat", "phi_hat", "gamma_star", "disp_common") if k in ref and k in res)
    print("%-28s" % label, {k: round(v, 1) for k, v in best.items() if k in ("common_disp", "tagwise_disp", "glm", "mc", "match_quantiles", "total")}, same, flush=True)
