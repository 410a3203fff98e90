"""Seeded synthetic count matrices of the BASELINE.json configs (C1..C5).

The recipe follows the reference's simulation set-up (/root/reference/simulations/sim_DEpipe.R:27-33,90-95 and
sim_DEpipe_helpers.R:11-14,27-46): a random half of the genes is shifted up in one batch and the other half in
another, each batch has its own dispersion, and a subset of genes carries a biological fold change.  Draws are
gamma-Poisson (numpy PCG64), so the marginals are negative binomial.  C1 is the README example
(/root/reference/README.md:35-38).
"""
import numpy as np


class Case:
    def __init__(self, name, counts, batch, group=None, covar_mod=None, full_mod=True, shrink=False,
                 shrink_disp=False, gene_subset_n=None):
        self.name, self.counts, self.batch, self.group, self.covar_mod = name, counts, batch, group, covar_mod
        self.full_mod, self.shrink, self.shrink_disp, self.gene_subset_n = full_mod, shrink, shrink_disp, gene_subset_n

    def kwargs(self):
        return dict(group=self.group, covar_mod=self.covar_mod, full_mod=self.full_mod, shrink=self.shrink,
                    shrink_disp=self.shrink_disp, gene_subset_n=self.gene_subset_n)


def nb_matrix(rng, mu, phi):
    """NB(mean mu, dispersion phi) draws; mu G x N, phi broadcastable."""
    shape = 1.0 / np.broadcast_to(phi, mu.shape)
    lam = rng.gamma(shape, mu / shape)
    return rng.poisson(lam).astype(np.int32)


def simulate(G, n_per_batch, batch_fc, batch_phi, seed, group_in_batch=True, bio_fc=1.5, n_de=None,
             low_frac=0.02, base_logmean=4.0, base_logsd=1.5):
    rng = np.random.default_rng(seed)
    B = len(n_per_batch)
    N = int(sum(n_per_batch))
    batch = np.repeat(np.arange(1, B + 1), n_per_batch)
    group = np.concatenate([np.arange(n) % 2 for n in n_per_batch]) if group_in_batch else None
    base = np.exp(rng.normal(base_logmean, base_logsd, G)).clip(0.5, 1e5)
    low = rng.random(G) < low_frac
    base[low] = rng.uniform(0.02, 0.5, low.sum())              # exercises the zero filter and x <= 1 pass-through
    libf = rng.uniform(0.7, 1.3, N)
    mu = base[:, None] * libf[None, :]
    direction = rng.random(G) < 0.5                             # sim_DEpipe_helpers.R:27-46: half up, half down
    for b in range(B):
        cols = batch == b + 1
        fc = np.where(direction, batch_fc[b], 1.0 / batch_fc[b])
        mu[:, cols] *= fc[:, None]
    if group is not None and bio_fc != 1:
        n_de = n_de if n_de is not None else max(2, G // 20)
        de = rng.permutation(G)[: 2 * n_de]
        mu[np.ix_(de[:n_de], group == 1)] *= bio_fc
        mu[np.ix_(de[n_de:], group == 1)] /= bio_fc
    phi = np.empty((1, N))
    for b in range(B):
        phi[0, batch == b + 1] = batch_phi[b]
    counts = nb_matrix(rng, mu, phi)
    return counts, batch, group


def c1(seed=1, G=50):
    rng = np.random.default_rng(seed)
    counts = rng.negative_binomial(10, 0.1, size=(G, 8)).astype(np.int32)        # README.md:35
    return Case("C1", counts, np.array([1, 1, 1, 1, 2, 2, 2, 2]))


def c1_group(seed=1, G=50):
    c = c1(seed, G); c.name = "C1-group"; c.group = np.array([0, 1, 0, 1, 0, 1, 0, 1]); return c   # README.md:44-45


def c1_covar(seed=1, G=50):
    c = c1(seed, G); c.name = "C1-covar"
    rng = np.random.default_rng(seed + 100)
    c.group = np.array([0, 1, 0, 1, 0, 1, 0, 1])
    c.covar_mod = np.column_stack([np.array([0, 0, 1, 1, 0, 0, 1, 1.0]), rng.normal(size=8)])      # README.md:51-54
    return c


def c2(seed=2, G=20000):
    counts, batch, group = simulate(G, [8, 8, 8], (1, 1.5, 2), (0.15, 0.30, 0.45), seed, n_de=max(2, G // 20))
    return Case("C2", counts, batch, group)


def c3(seed=3, G=23000, crossed=False):
    n_b = [17, 12, 18]
    counts, batch, _ = simulate(G, n_b, (1, 1.5, 2), (0.15, 0.30, 0.45), seed, group_in_batch=False)
    rng = np.random.default_rng(seed + 100)
    if crossed:                                                  # C3b: crossed 2-level group -> within-batch Levenberg
        group = np.concatenate([np.arange(n) % 2 for n in n_b])
    else:                                                        # nested as gfrn_application.R:23-38
        group = np.concatenate([[0] * 12 + [1] * 5, [0] * 6 + [2] * 6, [0] * 9 + [3] * 9])
    N = sum(n_b)
    cov = np.column_stack([(rng.random(N) < 0.5).astype(float), rng.normal(size=N)])
    cov[:, 1] -= cov[:, 1].mean()
    return Case("C3b" if crossed else "C3", counts, batch, group, covar_mod=cov)


def c4(seed=4, G=20000, n=50):
    counts, batch, group = simulate(G, [n] * 4, (1, 1.5, 2, 3), (0.15, 0.30, 0.45, 0.60), seed)
    return Case("C4", counts, batch, group, shrink=True, shrink_disp=False)


def c5(seed=5, G=60000, n=100, B=10):
    fc = np.linspace(1, 3, B)
    phi = np.linspace(0.1, 0.5, B)
    counts, batch, group = simulate(G, [n] * B, fc, phi, seed)
    return Case("C5", counts, batch, group, shrink=True, shrink_disp=True)


CASES = dict(C1=c1, C1g=c1_group, C1c=c1_covar, C2=c2, C3=c3, C3b=lambda **k: c3(crossed=True, **k), C4=c4, C5=c5)
