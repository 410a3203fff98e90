"""-m gpu, collected last: cbs_mc_int_static, the Monte-Carlo integration of one batch without any allocation.

The entry point has no counterpart in the reference (added for callers that cannot
call cudaMalloc); its scratch is one static device array.  Checked through the C ABI against the existing entry point
cbs_mc_int on the same inputs, within the tolerance the Monte-Carlo tests use (|a-b| <= 1e-6*|b| + 1e-9, fallback counts
exact), on two calls in a row with different shapes (the second call finds the first call's contents in the array), in
both likelihood modes; plus: device memory in use does not change across the calls, oversize problems are rejected.
"""
import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu


def make_batch(G, n, seed, zero_frac=0.01):
    rng = np.random.default_rng(seed)
    base = np.exp(rng.normal(3.5, 1.5, G)).clip(0.3, 5e4)
    mu = base[:, None] * rng.uniform(0.7, 1.3, (1, n)) * np.exp(rng.normal(0, 0.05, (G, n)))
    phi = np.exp(rng.uniform(np.log(0.05), np.log(1.0), G))
    shape = 1.0 / phi[:, None]
    x = rng.poisson(rng.gamma(shape, mu / shape)).astype(np.int32)
    dead = rng.random(G) < zero_frac
    mu[dead, 0] = 0.0
    x[dead, 0] = 0
    return x, mu, rng.normal(0, 0.5, G), phi


def test_static_scratch_entry_matches_cbs_mc_int_on_two_calls_in_a_row(pkg):
    torch = pytest.importorskip("torch")
    L = pkg._lib
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cases = [(1500, 20, 0), (700, 6, 1), (1500, 20, 1)]              # (G, n, mode): big, smaller (stale scratch behind it), big again
    prepared = []
    for k, (G, n, mode) in enumerate(cases):
        x, mu, gamma, phi = make_batch(G, n, seed=100 + k)
        want = L.mc_int(x, mu, gamma, phi, mode=mode)              # existing entry point (allocates; host buffers)
        bufs = dict(x=torch.from_numpy(np.ascontiguousarray(x.T)).to(dev),            # (n, G) row-major = G x n column-major
                    mu=torch.from_numpy(np.ascontiguousarray(mu.T)).to(dev),
                    gamma=torch.from_numpy(gamma).to(dev), phi=torch.from_numpy(phi).to(dev),
                    gs=torch.full((G,), float("nan"), dtype=torch.float64, device=dev),
                    ps=torch.full((G,), float("nan"), dtype=torch.float64, device=dev),
                    nfb=torch.full((1,), -1, dtype=torch.int64, device=dev))
        prepared.append((G, n, mode, want, bufs))
    # set-up: the first call on a device looks the static array's address up (and the driver loads the kernels' code lazily)
    G, n, mode, _, b = prepared[0]
    L.mc_int_static(G, n, b["x"].data_ptr(), b["mu"].data_ptr(), b["gamma"].data_ptr(), b["phi"].data_ptr(),
                    b["gs"].data_ptr(), b["ps"].data_ptr(), b["nfb"].data_ptr(), mc_mode=mode)
    torch.cuda.synchronize()
    b["gs"].fill_(float("nan")); b["ps"].fill_(float("nan")); b["nfb"].fill_(-1)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    n0 = L.kernel_launches()
    for G, n, mode, want, b in prepared:                            # calls in a row on one stream: no overlap on the GPU
        L.mc_int_static(G, n, b["x"].data_ptr(), b["mu"].data_ptr(), b["gamma"].data_ptr(), b["phi"].data_ptr(),
                        b["gs"].data_ptr(), b["ps"].data_ptr(), b["nfb"].data_ptr(), mc_mode=mode)
    torch.cuda.synchronize()
    assert L.kernel_launches() - n0 == 5 * len(prepared)
    assert torch.cuda.mem_get_info()[0] == free_before, "the entry point must not allocate device memory"
    for G, n, mode, (gs, ps, nfb), b in prepared:
        assert int(b["nfb"].item()) == nfb
        assert rel_err(b["gs"].cpu().numpy(), gs) <= 1.0
        assert rel_err(b["ps"].cpu().numpy(), ps) <= 1.0
    # n_fallback is optional
    G, n, mode, (gs, ps, nfb), b = prepared[1]
    b["gs"].fill_(float("nan"))
    L.mc_int_static(G, n, b["x"].data_ptr(), b["mu"].data_ptr(), b["gamma"].data_ptr(), b["phi"].data_ptr(),
                    b["gs"].data_ptr(), b["ps"].data_ptr(), None, mc_mode=mode)
    torch.cuda.synchronize()
    assert rel_err(b["gs"].cpu().numpy(), gs) <= 1.0


def test_static_scratch_entry_rejects_what_does_not_fit(pkg):
    torch = pytest.importorskip("torch")
    L = pkg._lib
    d = torch.zeros(16, dtype=torch.float64, device="cuda").data_ptr()
    G, n = 60000, 100                                               # 8 * 104 * 60000 = 50 MB of scratch
    assert 8 * (n + 4) * G > L.STATIC_SCRATCH_BYTES
    with pytest.raises(L.CombatSeqError) as e:
        L.mc_int_static(G, n, d, d, d, d, d, d)
    assert e.value.code == L.CBS_EINVAL and "static array holds" in str(e.value) and "cbs_mc_int" in str(e.value)
    with pytest.raises(L.CombatSeqError):
        L.mc_int_static(100, 4, d, d, d, d, None, d)                # null output
