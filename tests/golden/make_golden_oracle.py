"""Regenerates tests/golden/oracle_*.npz: outputs of the C++ CPU oracle (oracle/liboracle.so) on the seeded synthetic
cases whose Monte-Carlo stage is too slow to recompute inside every GPU test run (the oracle is quadratic in the gene
count: 170 s on 8 threads for the C5-shaped case).  Run here, in the build container; committed with its outputs.
The inputs are not stored: the tests regenerate them from the same seeds (combat-seq_b200/synth.py) and check a checksum.
PARITY UNPINNED against real R/edgeR (see oracle/oracle.py).

adjusted is stored as int32 (the oracle's doubles hold integers), mu_hat for every 8th gene only.
"""
import importlib
import os
import sys
import time
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

CASES = {                     # tag -> (synth factory name, kwargs)
    "C5shape_1500x1000": ("c5", dict(G=1500, n=100, B=10)),
    "C4_1200_nb50": ("c4", dict(G=1200, n=50)),
}
MU_STRIDE = 8


def checksum(counts):
    return zlib.crc32(np.ascontiguousarray(counts, dtype=np.int32).tobytes())


def main():
    import oracle as orc
    synth = importlib.import_module("combat-seq_b200.synth")
    design = importlib.import_module("combat-seq_b200.design")
    for tag, (fac, kw) in CASES.items():
        case = getattr(synth, fac)(**kw)
        d = design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
        t = time.time()
        r = orc.combat_seq(case.counts.astype(np.float64), d.batch_idx, d.mod, shrink=case.shrink,
                           shrink_disp=case.shrink_disp, nthreads=orc.max_threads())
        adj = r["adjusted"]
        assert np.array_equal(adj, np.rint(adj)) and adj.max() < 2 ** 31
        np.savez_compressed(os.path.join(HERE, "oracle_%s.npz" % tag), input_crc32=np.uint32(checksum(case.counts)),
                            adjusted=adj.astype(np.int32), keep=r["keep"], lib_size=r["lib_size"],
                            disp_common=r["disp_common"], phi_hat=r["phi_hat"], gamma_hat=r["gamma_hat"],
                            gamma_star=r["gamma_star"], phi_star=r["phi_star"], mu_hat_sub=r["mu_hat"][::MU_STRIDE],
                            n_mc_fallback=r["n_mc_fallback"])
        print("%s: oracle %.1f s" % (tag, time.time() - t))


if __name__ == "__main__":
    main()
