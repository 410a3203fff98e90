"""Time the Monte-Carlo stage alone (cbs profile hooks) on one synthetic batch: exact pairwise kernel (algo 1), the
contraction over all source tiles (band 0) and over the screened lists (band 1).  usage: mc_time.py G n [algo:band ...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import __graft_entry__ as g
pkg = g.load_package(); L = pkg._lib
from test_gpu_mc import make_batch
G, n = int(sys.argv[1]), int(sys.argv[2])
variants = [tuple(int(v) for v in a.split(":")) for a in sys.argv[3:]] or [(2, 0), (2, 1)]
x, mu, gamma, phi = make_batch(G, n, seed=1, phi_lo=0.01, phi_hi=3.0)
print("fp64 fma peak %.2f TF, dmma peak %.2f TF" % (L.fp64_peak_tflops(), L.fp64_mma_peak_tflops()))
ref = None
for algo, band in variants:
    L.set_tuning("mc_algo", algo); L.set_tuning("mc_band", band)
    for mode in (0, 1):
        L.mc_int(x, mu, gamma, phi, mode=mode)
        L.get_diag("reset")
        L.profile_enable(True)
        r = L.mc_int(x, mu, gamma, phi, mode=mode)
        prof = L.profile_dump(); L.profile_enable(False)
        tot = sum(v[0] for v in prof.values())
        if ref is None or ref[0] != mode: ref = (mode, r) if band == 0 or algo == 1 else ref
        dev = ""
        if ref is not None and ref[0] == mode and ref[1] is not r:
            dev = " max rel dev vs first: %.2e" % float(np.nanmax(np.abs(r[0] - ref[1][0]) / (np.abs(ref[1][0]) + 1e-12)))
        print("algo %d band %d mode %d G=%d n=%d: kernels %.2f ms, fallback %d, Kc %d, tiles %d/%d%s" % (
            algo, band, mode, G, n, tot, r[2], L.get_diag("mc_terms"), L.get_diag("mc_tiles_done"), L.get_diag("mc_tiles_all"), dev),
            {k: round(v[0], 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])[:8]}, flush=True)
