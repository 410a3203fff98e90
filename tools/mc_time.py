"""Time the Monte-Carlo stage alone (cbs profile hooks) for both algorithms on one synthetic batch."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import __graft_entry__ as g
pkg = g.load_package(); L = pkg._lib
from test_gpu_mc import make_batch
G, n = int(sys.argv[1]), int(sys.argv[2])
algos = [int(a) for a in sys.argv[3:]] or [2]
x, mu, gamma, phi = make_batch(G, n, seed=1)
print("fp64 fma peak %.2f TF, dmma peak %.2f TF" % (L.fp64_peak_tflops(), L.fp64_mma_peak_tflops()))
for algo in algos:
    L.set_tuning("mc_algo", algo)
    for mode in (0, 1):
        L.mc_int(x, mu, gamma, phi, mode=mode)
        L.profile_enable(True)
        r = L.mc_int(x, mu, gamma, phi, mode=mode)
        prof = L.profile_dump(); L.profile_enable(False)
        tot = sum(v[0] for v in prof.values())
        print("algo %d mode %d G=%d n=%d: kernels %.2f ms, terms/s %.3g, fallback %d, Kc %d" % (algo, mode, G, n, tot, G * (G - 1.0) * n / tot * 1e3, r[2], L.get_diag("mc_terms")),
              {k: round(v[0], 3) for k, v in prof.items()}, "checksum %.12e %.12e" % (float(np.nansum(r[0])), float(np.nansum(r[1]))))
