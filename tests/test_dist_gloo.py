"""not-gpu tier: the gene-sharded (N > 1) data flow on CPU, world_size 2 over gloo.

The multi-GPU path shards the kept genes by rows; the only cross-rank steps are (i) the all-reduce
of the per-chunk partial sums of the common-dispersion objective, (ii) one all-gather of mu_hat / gamma_hat / phi_hat
before the Monte-Carlo integration, (iii) row-range assembly of the adjusted matrix.  Here two gloo ranks run exactly
that flow with the library's own partition (cbs_shard_range, host-only) and the CPU oracle as the per-gene compute, and
the assembled adjusted counts must equal the unsharded oracle run exactly (parameters within the 1e-6 bar: the
unsharded oracle sums the common-dispersion objective gene by gene, the sharded flow chunk by chunk).  That is the
property the NCCL path relies on: per-gene stages do not depend on the shard, and the one cross-gene reduction is summed
in a fixed chunk order whatever the number of ranks.
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHUNK = 128


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _chunk_sums(v, g0, G):
    """per-chunk partial sums of a shard's per-gene values, placed at the chunk's global index (zeros elsewhere)"""
    nchunk = (G + CHUNK - 1) // CHUNK
    out = np.zeros(nchunk)
    for c in range(g0 // CHUNK, nchunk):
        lo, hi = max(c * CHUNK, g0), min((c + 1) * CHUNK, g0 + len(v))
        if hi > lo:
            out[c] = v[lo - g0:hi - g0].sum()
    return out


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import __graft_entry__ as graft
    import pyref
    pkg = graft.load_package()
    orc = graft.load_oracle()
    import importlib
    synth = importlib.import_module("combat-seq_b200.synth")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        case = synth.c5(G=2600, n=4, B=3)                        # 21 reduction chunks of 128 genes: shards of 1408 and ~1190 genes
        d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
        counts = case.counts.astype(np.float64)
        bidx, mod = d.batch_idx, d.mod
        B = int(bidx.max()) + 1
        N = counts.shape[1]
        # replicated prep: zero filter + library sizes need all kept genes (every rank holds the count matrix)
        keep = np.ones(counts.shape[0], bool)
        for b in range(B):
            keep &= (counts[:, bidx == b] != 0).any(axis=1)
        Y = counts[keep]
        G = Y.shape[0]
        lib = Y.sum(axis=0)
        g0, g1 = pkg._lib.shard_range(G, rank, world)
        Ys = Y[g0:g1]
        # (i) common dispersion: Brent in lock-step on every rank; each evaluation all-reduces the chunk partials
        common = np.empty(B)
        for b in range(B):
            cols = np.flatnonzero(bidx == b)
            Xb = mod[cols]
            sel = Ys[:, cols].sum(axis=1) >= 5

            def f(delta):
                apl = orc.apl(Ys[:, cols], Xb, lib[cols], np.array([delta ** 4]))[:, 0]
                part = torch.from_numpy(_chunk_sums(np.where(sel, apl, 0.0), g0, G))
                dist.all_reduce(part)                            # disjoint chunks: exact zeros from the other rank
                total = 0.0
                for v in part.numpy().tolist():                  # summed in chunk order on every rank
                    total += v
                return -total
            common[b] = pyref.brent_fmin(0.0, 4 ** 0.25, f, 1e-5) ** 4
        # per-gene stages on the shard
        # tagwise dispersion: the offsets are log(colSums) of the WHOLE kept matrix, so go through the APL entry point
        phi_s = np.empty((g1 - g0, B))
        pts = np.linspace(-6, 6, 11)
        for b in range(B):
            cols = np.flatnonzero(bidx == b)
            apl = orc.apl(Ys[:, cols], mod[cols], lib[cols], common[b] * 2.0 ** pts)
            sel = Ys[:, cols].sum(axis=1) >= 5
            phi_s[:, b] = [common[b] * 2.0 ** orc.maximize_interpolant(pts, a) if s else common[b] for a, s in zip(apl, sel)]
        design = d.design
        coef1, _ = orc.glm_fit(Ys, design, phi_s, bidx, lib, prior_count=1e-4)
        alpha = coef1[:, :B] @ (np.bincount(bidx) / N)
        coef2, mu_s = orc.glm_fit(Ys, design, phi_s, bidx, lib, gene_offset=alpha, prior_count=1e-4)
        gam_s = np.ascontiguousarray(coef2[:, :B])

        # (ii) all-gather of the per-gene fit results
        def allgather(a):
            a = np.ascontiguousarray(a)
            sizes = [pkg._lib.shard_range(G, r, world) for r in range(world)]
            bufs = [torch.empty((e - s,) + a.shape[1:], dtype=torch.float64) for s, e in sizes]
            bufs[rank].copy_(torch.from_numpy(a))
            for r in range(world):                               # shards are uneven (chunk-rounded): one broadcast per rank
                dist.broadcast(bufs[r], r)
            return np.concatenate([b_.numpy() for b_ in bufs], axis=0)
        mu_all, gam_all, phi_all = allgather(mu_s), allgather(gam_s), allgather(phi_s)
        # Monte-Carlo: this rank's targets against all sources
        gstar = np.empty((g1 - g0, B)); pstar = np.empty((g1 - g0, B))
        for b in range(B):
            cols = np.flatnonzero(bidx == b)
            gs, ps, _ = orc.mc_int(Y[:, cols], mu_all[:, cols], gam_all[:, b], phi_all[:, b], targets=np.arange(g0, g1))
            gstar[:, b], pstar[:, b] = gs[g0:g1], ps[g0:g1]
        # match_quantiles on the shard
        mu_star = np.exp(np.log(mu_s) - gstar[:, bidx])
        phi_star = pstar.mean(axis=1)
        adj_s = np.empty_like(Ys)
        for b in range(B):
            cols = np.flatnonzero(bidx == b)
            adj_s[:, cols] = orc.match_quantiles(Ys[:, cols], mu_s[:, cols], phi_s[:, b], mu_star[:, cols], phi_star)
        # (iii) assembly (stands for the per-rank row-range D2H copies)
        adj_all, gstar_all = allgather(adj_s), allgather(gstar)
        if rank == 0:
            np.savez(os.path.join(out_dir, "sharded.npz"), adjusted=adj_all, common=common, gamma_star=gstar_all,
                     phi_hat=phi_all, keep=keep)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_rank_gene_sharding_reproduces_unsharded_counts(tmp_path, oracle, pkg, synth):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = np.load(os.path.join(str(tmp_path), "sharded.npz"))
    case = synth.c5(G=2600, n=4, B=3)
    d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
    ref = oracle.combat_seq(case.counts.astype(np.float64), d.batch_idx, d.mod, shrink=True, shrink_disp=True)
    keep = ref["keep"] == 1
    assert np.array_equal(got["keep"], keep)
    assert keep.sum() > 2 * CHUNK                                   # both ranks own genes, rank 1's start mid-matrix
    assert np.allclose(got["common"], ref["disp_common"], rtol=1e-6)
    assert np.allclose(got["phi_hat"], ref["phi_hat"][keep], rtol=1e-6)
    assert np.allclose(got["gamma_star"], ref["gamma_star"][keep], rtol=1e-6, atol=1e-9)
    assert np.array_equal(got["adjusted"], ref["adjusted"][keep])


def test_reference_arm_runs_on_rank0_only():
    """bench.py --impl reference under torchrun: ranks != 0 exit 0 without work or output."""
    import subprocess
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
