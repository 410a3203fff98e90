"""Time the device count generator (cbs_simulate_counts) on the C5 shape against numpy's generator on the host."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import __graft_entry__ as g
pkg = g.load_package(); L = pkg._lib
G, N, B = 60000, 1000, 10
rng = np.random.default_rng(5)
base = np.exp(rng.normal(4.0, 1.5, G)).clip(0.5, 1e5)
colf = rng.uniform(0.7, 1.3, N)
level = np.repeat(np.arange(B), N // B)
fold = np.where(rng.random(G)[:, None] < 0.5, np.linspace(1, 3, B)[None, :], 1 / np.linspace(1, 3, B)[None, :])
phi = np.linspace(0.1, 0.5, B)[level]
d = torch.empty((N, G), dtype=torch.int32, device="cuda")
for it in range(3):
    torch.cuda.synchronize(); t = time.perf_counter()
    L.simulate_counts(base, colf, phi, seed=it, level=level, fold=fold, counts_dev_ptr=d.data_ptr())
    torch.cuda.synchronize(); dt = time.perf_counter() - t
    print("device generator (incl. parameter upload), %d x %d: %.2f ms, mean count %.1f" % (G, N, dt * 1e3, d.float().mean().item()), flush=True)
t = time.perf_counter()
mu = base[:, None] * colf[None, :] * fold[:, level]
r = 1.0 / phi[None, :]
x = rng.negative_binomial(np.broadcast_to(r, mu.shape), r / (r + mu))
print("numpy negative_binomial on the host, same shape: %.2f s, mean count %.1f" % (time.perf_counter() - t, x.mean()))
