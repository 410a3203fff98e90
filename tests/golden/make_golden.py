"""Regenerates the golden vectors under tests/golden/ (run here, in the build container; committed with its outputs).

The reference (/root/reference) ships no tests, fixtures or stored outputs and R/edgeR are not
installed, so the pins are:
  nb_kat.npz       NB log-pmf / cdf / quantile known answers from mpmath at 50 digits (independent of every
                   implementation in this repo) on a grid spanning phi in [1e-4, 10], mu in [1e-3, 1e6];
  pipeline_*.npz   every intermediate of ComBat_seq() on the README configurations and on small C2/C3/C3b/C4-style
                   matrices, computed by oracle/pyref.py (numpy/scipy restatement of ComBat_seq.R + helper_seq.R that
                   shares no code with the C++ oracle).
PARITY UNPINNED against real R/edgeR.
"""
import importlib
import os
import sys

import mpmath as mp
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
mp.mp.dps = 50


def nb_logpmf_mp(x, size, mu):
    x, size, mu = mp.mpf(x), mp.mpf(size), mp.mpf(mu)
    return (mp.loggamma(x + size) - mp.loggamma(size) - mp.loggamma(x + 1) + size * mp.log(size / (size + mu))
            + x * mp.log(mu / (size + mu)))


def nb_cdf_table(size, mu, xmax):
    """cdf(0..xmax) by exact summation of the pmf recurrence in 50-digit arithmetic (b = x+1 is an integer, so the
    regularised incomplete beta function is a finite sum)"""
    size, mu = mp.mpf(size), mp.mpf(mu)
    q = mu / (size + mu)
    p = (size / (size + mu)) ** size
    out, acc = [], mp.mpf(0)
    for k in range(xmax + 1):
        acc += p
        out.append(acc)
        p = p * (k + size) / (k + 1) * q
    return out


CDF_MAX_X = 400000


def make_kat():
    rng = np.random.default_rng(20240607)
    rows, cdf, qrows, qans = [], [], [], []
    fuzz = 1 - 64 * mp.mpf(2) ** -52
    for phi in (1e-4, 1e-3, 0.01, 0.05, 0.15, 0.5, 1.0, 3.0, 10.0):
        for mu in (1e-3, 0.1, 1.0, 7.5, 90.0, 1e3, 2.5e4, 1e6):
            size = 1.0 / phi
            sd = np.sqrt(mu + phi * mu * mu)
            xs = sorted({0, 1, 2, int(mu), int(mu + sd), int(max(0, mu - sd)), int(mu + 3 * sd),
                         int(rng.integers(0, int(mu + 4 * sd) + 2))})
            tab = nb_cdf_table(size, mu, xs[-1]) if xs[-1] <= CDF_MAX_X else None
            for x in xs:
                rows.append((float(x), size, mu))
                cdf.append(float(tab[x]) if tab is not None else np.nan)
                # quantile case: a probability mid-way between two steps of the cdf, so the answer is x by construction
                if tab is not None and x >= 1 and tab[x] - tab[x - 1] > mp.mpf(10) ** -12 and tab[x] < 1 - mp.mpf(10) ** -9:
                    p = float((tab[x - 1] + tab[x]) / 2)
                    pf = mp.mpf(p) * fuzz
                    y = next(k for k in range(max(0, x - 2), x + 3) if tab[k] >= pf)
                    qrows.append((p, size, mu)); qans.append(float(y))
    arr = np.array(rows)
    logpmf = np.array([float(nb_logpmf_mp(*r)) for r in rows])
    np.savez_compressed(os.path.join(HERE, "nb_kat.npz"), x_size_mu=arr, logpmf=logpmf, cdf=np.array(cdf),
                        q_p_size_mu=np.array(qrows), quantile=np.array(qans))
    print("nb_kat: %d pmf rows, %d cdf rows, %d quantile rows" % (len(rows), int(np.isfinite(cdf).sum()), len(qrows)))


PIPELINE_CASES = [  # (file tag, synth case, synth kwargs, option overrides)
    ("C1", "C1", {}, {}),
    ("C1g", "C1g", {}, {}),
    ("C1c", "C1c", {}, {}),
    ("C1g_shrink", "C1g", {}, dict(shrink=True, shrink_disp=True)),
    ("C1g_shrink_meanonly", "C1g", {}, dict(shrink=True, shrink_disp=False)),
    ("C2_60", "C2", dict(G=60), {}),
    ("C3_40", "C3", dict(G=40), {}),
    ("C3b_40", "C3b", dict(G=40), {}),
    ("C4_40x5", "C4", dict(G=40, n=5), {}),
]


def make_pipeline():
    import pyref
    synth = importlib.import_module("combat-seq_b200.synth")
    design = importlib.import_module("combat-seq_b200.design")
    for tag, name, kw, opt in PIPELINE_CASES:
        case = synth.CASES[name](**kw)
        d = design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
        shrink = opt.get("shrink", case.shrink); sd = opt.get("shrink_disp", case.shrink_disp)
        r = pyref.combat_seq(case.counts, d.batch_idx, d.mod, shrink, sd)
        np.savez_compressed(os.path.join(HERE, "pipeline_%s.npz" % tag), counts=case.counts.astype(np.int32),
                            batch_idx=d.batch_idx, mod=d.mod, shrink=shrink, shrink_disp=sd, **r)
        print("pipeline_%s: %s, changed %d of %d counts" % (tag, case.counts.shape, int((r["adjusted"] != case.counts).sum()), case.counts.size))


if __name__ == "__main__":
    make_kat()
    make_pipeline()
