#!/usr/bin/env python
"""bench.py — genes x samples adjusted per second per ComBat_seq call (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C5] [--impl reference]

One step = one whole ComBat_seq numeric body (zero filter, library sizes, per-batch dispersions, the two NB GLM
fits, Monte-Carlo EB when shrink=TRUE, quantile matching, re-insertion) over one synthetic count matrix of the
named workload.  `value` is measured with the counts already resident in HBM and the adjusted matrix left in
HBM; `e2e` goes through the host-buffer C-ABI call (H2D of the counts and D2H of the adjusted matrix inside the
timed region, pinned host memory).  Under torchrun (N > 1) the genes are sharded over the ranks (weak = False:
total work is fixed, so "scaling": "strong").

--impl reference times the CPU oracle (oracle/liboracle.so: a C++ restatement of the reference's R/edgeR path;
R itself is not installed in this image) on a bounded row-sample of the same workload and extrapolates by the
stages' known complexity; see cpu_baseline.sample in the JSON line.
"""
import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

METRIC = "genes_x_samples_adjusted_per_sec_per_ComBat_seq_call"
UNIT = "gene-samples/s"

# Algorithmic work of the dominant kernels (DESIGN.md "Kernels and rooflines"):
#   k_mc_*: one NB term (i, j, s) = 1 DFMA for the contraction x_is*A_js + one lgamma(x_is + r_j) by the Stirling
#           series (1 log, 1 reciprocal, 5-term Horner) ~= 38 FP64 flop in total (counted from the SASS of the
#           inner loop: DFMA = 2, DADD/DMUL = 1).
#   k_match_quantiles: 4 B int32 count + 8 B mu_hat read + 8 B adjusted written per element = 20 B.
MC_FLOP_PER_TERM = 38.0
MQ_BYTES_PER_ELEM = 20.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default=os.environ.get("CBS_BENCH_WORKLOAD", "C5"))
    ap.add_argument("--genes", type=int, default=0, help="override the workload's gene count (debugging only)")
    ap.add_argument("--cpu-genes", type=int, default=0, help="rows of the bounded CPU sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--tune", action="append", default=[], metavar="KEY=VALUE", help="cbs_set_tuning knob for A/B runs (development)")
    ap.add_argument("--inproc", type=int, default=0, help="single-process multi-GPU (cbs_input.n_gpus): time the host-buffer "
                                                           "call on this many devices of one process (what the R .Call reaches)")
    return ap.parse_args()


def make_case(pkg, name, genes):
    synth = importlib.import_module("combat-seq_b200.synth")
    kw = {"G": genes} if genes else {}
    case = synth.CASES[name](**kw)
    d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
    return case, d


def describe(case, d):
    G0, N = case.counts.shape
    nb = np.bincount(d.batch_idx)
    return {"workload": "%s: %d genes x %d samples, %d batches (%s per batch), p=%d design columns, shrink=%s, "
                        "shrink.disp=%s, gene.subset.n=NULL" % (case.name, G0, N, len(nb), "/".join(map(str, sorted(set(nb.tolist())))),
                                                                   len(nb) + d.mod.shape[1] - 1, case.shrink, case.shrink_disp),
            "genes": int(G0), "samples": int(N), "batches": int(len(nb))}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is not None:
            self.p.terminate()
            try:
                self.p.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def cpu_sample(case, d, n_rows, threads):
    """Time the CPU oracle on the first n_rows kept genes (all samples, same design/options) and extrapolate to the
    whole workload: every stage is linear in the gene count except the Monte-Carlo integration, which is
    quadratic (each target gene is integrated against every other gene)."""
    orc = graft.load_oracle()
    G0, N = case.counts.shape
    nz = np.ones(G0, bool)
    for b in range(d.batch_idx.max() + 1):
        nz &= (case.counts[:, d.batch_idx == b] != 0).any(axis=1)
    rows = np.flatnonzero(nz)[:n_rows]
    sub = np.asfortranarray(case.counts[rows].astype(np.float64))
    t = time.perf_counter()
    res = orc.combat_seq(sub, d.batch_idx, d.mod, shrink=case.shrink, shrink_disp=case.shrink_disp, nthreads=threads)
    wall = time.perf_counter() - t
    t_mc = res["stage_ms"]["mc"] * 1e-3
    t_lin = wall - t_mc
    scale = float(nz.sum()) / float(len(rows))
    t_full = t_lin * scale + t_mc * scale * scale
    return {"rows": int(len(rows)), "wall_s": wall, "mc_s": t_mc, "t_full_s": t_full, "value": G0 * N / t_full}


def measured_pair(pkg, L, name):
    """An UNEXTRAPOLATED GPU / CPU pair: a workload the CPU oracle finishes in seconds (C2: 20 000 x 24, shrink = FALSE),
    whole call on both sides, same inputs, host buffers in and out."""
    case, d = make_case(pkg, name, 0)
    G0, N = case.counts.shape
    orc = graft.load_oracle()
    threads = cpu_threads()
    x = np.asfortranarray(case.counts.astype(np.float64))
    t = time.perf_counter()
    ref = orc.combat_seq(x, d.batch_idx, d.mod, shrink=case.shrink, shrink_disp=case.shrink_disp, nthreads=threads)
    cpu_s = time.perf_counter() - t
    out = np.empty((G0, N), order="F")
    L.combat_seq_into(x, d.batch_idx, d.mod, out, shrink=case.shrink, shrink_disp=case.shrink_disp)
    ts = []
    for _ in range(5):
        t = time.perf_counter()
        L.combat_seq_into(x, d.batch_idx, d.mod, out, shrink=case.shrink, shrink_disp=case.shrink_disp)
        ts.append(time.perf_counter() - t)
    gpu_s = statistics.median(ts)
    return {"workload": describe(case, d)["workload"], "cpu_s": cpu_s, "cpu_threads": threads, "cpu_kind": "port (C++ oracle, not R)",
            "gpu_s": gpu_s, "gpu_value": G0 * N / gpu_s, "cpu_value": G0 * N / cpu_s, "unit": UNIT,
            "exact_match": float((out == ref["adjusted"]).mean()), "extrapolated": False}


def auto_cpu_rows(case):
    G0, N = case.counts.shape
    if not case.shrink:
        return int(min(G0, max(200, 4.0e6 / N)))          # linear stages only: ~10 s on 8 threads
    return int(min(G0, max(100, (1.5e8 / N) ** 0.5 * 1.0)))  # G_s^2 * N ~ 1.5e8 NB terms: ~10 s on 8 threads


def cpu_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pkg = graft.load_package()
    case, d = make_case(pkg, args.workload, args.genes)
    G0, N = case.counts.shape
    threads = cpu_threads()                       # explicit num_threads(): torchrun sets OMP_NUM_THREADS=1
    rows = args.cpu_genes or auto_cpu_rows(case)
    for _ in range(args.warmup):
        cpu_sample(case, d, max(50, rows // 4), threads)
    t_full, t_wall = [], 0.0
    for _ in range(args.steps):
        r = cpu_sample(case, d, rows, threads)
        t_full.append(r["t_full_s"]); t_wall += r["wall_s"]
    tf = sum(t_full) / len(t_full)
    value = G0 * N / tf
    sample = ("CPU oracle (C++ restatement of ComBat_seq.R + edgeR 3.22 + nmath; R is not installed) on the first %d kept "
              "genes x all %d samples per step, %d OpenMP threads; extrapolated to %d genes: linear stages x%.1f, "
              "Monte-Carlo stage x%.1f^2 (measured %.2f s per sampled step)" % (r["rows"], N, threads, G0, G0 / r["rows"],
                                                                                G0 / r["rows"], t_wall / args.steps))
    cfg = describe(case, d)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": tf * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def run_native(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    pkg = graft.load_package()
    L = pkg._lib
    for kv in args.tune:
        L.set_tuning(kv.split("=")[0], float(kv.split("=")[1]))
    if pkg.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: libcombatseq has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    comm = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        uid = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(L.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        comm = L.comm_create(bytes(uid.cpu().numpy().tobytes()), rank, world, local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    case, d = make_case(pkg, args.workload, args.genes)
    G0, N = case.counts.shape
    opts = dict(shrink=case.shrink, shrink_disp=case.shrink_disp, device=local, comm=comm)
    # pinned host buffers (e2e arm) and device-resident buffers (value arm)
    h_counts = torch.empty((N, G0), dtype=torch.int32, pin_memory=True)             # column-major G0 x N
    h_counts.numpy()[:] = np.ascontiguousarray(case.counts.T)
    h_adj = torch.empty((N, G0), dtype=torch.float64, pin_memory=True)
    d_counts = h_counts.to(dev)
    d_adj = torch.empty((N, G0), dtype=torch.float64, device=dev)
    counts_view = h_counts.numpy().T                                                # F-ordered (G0, N) view, no copy
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)                   # > 126 MB L2
    small = G0 * N * 4 < (200 << 20)

    def step_resident():
        return L.combat_seq_raw(None, d.batch_idx, d.mod, want=(), counts_dev_ptr=d_counts.data_ptr(),
                                adjusted_dev_ptr=d_adj.data_ptr(), shape=(G0, N), **opts)

    def step_e2e():
        return L.combat_seq_into(counts_view, d.batch_idx, d.mod, h_adj.numpy().T, **opts)

    def timed(fn, steps, warmup, sampler=None):
        for _ in range(warmup):
            fn()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        if sampler:
            sampler.start()
        res = None
        for a, b in ev:
            if small:
                flush.zero_()                            # L2 flush between timed iterations (outside the event pair)
            a.record()
            res = fn()
            b.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = sum(a.elapsed_time(b) for a, b in ev)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), res, clocks

    # ---- value arm: counts resident in HBM, kernels profiled with CUDA events on the launching stream ----
    for _ in range(args.warmup):
        step_resident()
    L.profile_enable(True)
    L.get_diag("reset")
    n0 = L.kernel_launches()
    ms_val, res, clocks = timed(step_resident, args.steps, 0, ClockSampler(local))
    launches = L.kernel_launches() - n0
    prof = L.profile_dump()
    L.profile_enable(False)
    diag = {k: L.get_diag(k) for k in ("mc_terms", "mc_tiles_done", "mc_tiles_all", "mc_tail")}
    # ---- e2e arm: host buffers through the C ABI ----
    ms_e2e, res_e, _ = timed(step_e2e, args.steps, min(args.warmup, 1) if not small else args.warmup)
    checksum = float(np.asarray(h_adj.numpy()[:, :: max(1, G0 // 997)]).sum())
    # ---- the same call from pageable host memory (what an R matrix is): one warm-up, `steps` timed calls, wall clock ----
    pg_counts = np.asfortranarray(case.counts.astype(np.float64))      # R's numeric matrix: column-major doubles
    pg_adj = np.empty((G0, N), order="F")
    L.combat_seq_into(pg_counts, d.batch_idx, d.mod, pg_adj, **opts)
    barrier()
    t_pg = time.perf_counter()
    for _ in range(args.steps):
        L.combat_seq_into(pg_counts, d.batch_idx, d.mod, pg_adj, **opts)
    barrier()
    t_pg = torch.tensor([(time.perf_counter() - t_pg) * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_pg, op=dist.ReduceOp.MAX)
    pageable = {"value": float(G0) * N * args.steps / (float(t_pg.item()) * 1e-3), "unit": UNIT, "ms_per_step": float(t_pg.item()) / args.steps,
                "h2d_bytes_per_step": int(G0) * N * 8, "d2h_bytes_per_step": int(G0) * N * 8,
                "note": "pageable (not pinned) float64 count matrix in, float64 matrix out, host wall clock: the buffers an R caller has"}
    del pg_counts, pg_adj
    c2_pair = None
    if rank == 0 and world == 1 and not args.no_cpu:
        c2_pair = measured_pair(pkg, L, "C2")

    if rank == 0:
        units = float(G0) * N * args.steps
        G = int(res["keep"].sum())
        nb = np.bincount(d.batch_idx)
        line = {"metric": METRIC, "value": units / (ms_val * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_val / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": describe(case, d), "clocks": clocks,
                "e2e": {"value": units / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                        "h2d_bytes_per_step": int(G0) * N * 4, "d2h_bytes_per_step": int(G0) * N * 8,
                        "checksum": checksum},
                "gpu_launches": int(launches)}
        line["config"]["l2"] = "L2 flushed between timed steps" if small else "inputs larger than L2 (%.0f MB counts)" % (G0 * N * 4 / 1e6)
        line["config"]["kept_genes"] = G
        line["stage_ms"] = {k: round(v, 3) for k, v in res["stage_ms"].items()}
        total_k = sum(v[0] for v in prof.values()) or 1.0
        line["kernels"] = {k: {"ms_per_step": round(v[0] / args.steps, 4), "launches_per_step": v[1] / args.steps,
                               "share": round(v[0] / total_k, 4)} for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}
        ra = rooflines(L, prof, G, int(G0), N, nb, world, args.steps, diag)
        line["roofline"] = ra[0] if ra else None
        line["roofline_all"] = ra
        line["mc_tiles"] = {"visited_per_step": diag["mc_tiles_done"] / args.steps, "all_per_step": diag["mc_tiles_all"] / args.steps,
                            "chebyshev_terms": diag["mc_terms"], "tail_estimate": diag["mc_tail"]}
        line["e2e_pageable"] = pageable
        if c2_pair is not None:
            line["c2_pair"] = c2_pair
        if world == 1 and not args.no_cpu:
            threads = cpu_threads()                       # explicit num_threads(): torchrun sets OMP_NUM_THREADS=1
            r = cpu_sample(case, d, args.cpu_genes or auto_cpu_rows(case), threads)
            line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": "CPU oracle (C++ restatement; R/edgeR not installed) on the first %d kept genes x %d "
                                              "samples, %.1f s wall (Monte-Carlo %.1f s); extrapolated to %d genes: linear stages "
                                              "x%.1f, Monte-Carlo x%.1f^2" % (r["rows"], N, r["wall_s"], r["mc_s"], G0,
                                                                              G0 / r["rows"], G0 / r["rows"])}
        print(json.dumps(line))
    if comm is not None:
        L.comm_destroy(comm)
    if world > 1:
        dist.destroy_process_group()


def ncu_table():
    """Per-launch figures taken from the committed `ncu --set full` captures of the 1-GPU C5 workload
    (profiles/ncu_kernels.json, written by tools/ncu_kernels.py from the captures' raw pages): DRAM bytes and executed
    FP64 operations (DFMA = 2 flop, DADD / DMUL = 1) of one launch."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_kernels.json")))
    except (OSError, ValueError):
        return {}


def ncu_entry(tab, name):
    for k, v in tab.items():
        if name.startswith(k):
            return v
    return None


def rooflines(L, prof, G, G0, N, nb, world, steps, diag):
    """One roofline entry per kernel with at least 2 % of the step's device time (north_star: every kernel against its
    roofline).  The dominant kernel's entry is also the line's `roofline` object.

    bound / algorithmic work per launch:
      k_mc_gemm        FP64 tensor pipe; flops = 2 * KTP * 64 * 128 per (target tile, source tile) pair actually visited
                       (the screened lists; `pairs_equiv` also gives the figure for all G^2 pairs, which the round-1 kernel
                       executed and the reference evaluates)
      k_scr_mma        bf16 tensor cores (tcgen05); flops = 2 * 128 * 256 * 16 * 3 per chunk of 16 columns and tile pair
      k_apl_lanes, k_glm_fit_lanes, k_match_quantiles_warp, k_mcg_prep_*
                       FP64 pipe; executed FP64 flops per launch counted by ncu on the committed capture of the same
                       workload (null when the shape differs from the capture)
      streaming kernels HBM; algorithmic bytes"""
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    bf16 = float(peaks.get("bf16_tflops", 1590.0))
    src = "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)"
    fp64 = L.fp64_peak_tflops()
    dmma = L.fp64_mma_peak_tflops()
    tab = ncu_table()
    captured_shape = world == 1 and (G0, N) == (60000, 1000)
    total_ms = sum(v[0] for v in prof.values()) or 1.0
    n_b = float(N) / len(nb)
    kc = diag["mc_terms"]
    kt = n_b + kc + 4 + 1
    ktp = (int(kt) + 7) // 8 * 8
    tiles_done, tiles_all = diag["mc_tiles_done"], diag["mc_tiles_all"]
    out = []
    for name, (ms, n) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
        if ms / total_ms < 0.02:
            continue
        avg_s = ms / n * 1e-3
        e = {"kernel": name, "share": round(ms / total_ms, 4), "avg_launch_ms": ms / n, "launches": n, "traffic": None}
        cap = ncu_entry(tab, name) if captured_shape else None
        if cap:
            e["traffic"] = cap.get("dram_bytes_per_launch")
        if name.startswith("k_mc_gemm"):
            flops = 2.0 * ktp * 64 * 128 * tiles_done / max(n, 1)          # executed: rank 0's visited tile pairs per launch
            ach = flops / avg_s / 1e12
            e.update(bound="tensor", achieved=ach, peak=dmma, unit="TFLOP/s", frac=ach / dmma,
                     note="FP64 tensor pipe (mma.sync m8n8k4 f64 = DMMA; tcgen05 has no FP64 kind). peak = DMMA microbenchmark with "
                          "shared-memory operands measured in this run (cbs_fp64_mma_peak; MEASURED_PEAKS.json has no FP64 figure). "
                          "inner length KTP = %d (n_b %d + %d Chebyshev terms + 5, padded to 8); source tiles visited %d of %d "
                          "(screened band); pairs_equiv = all G^2 pairs / time" % (ktp, n_b, kc, tiles_done, tiles_all),
                     pairs_equiv_per_s=float(G) / world * G / avg_s)
        elif name.startswith("k_scr_mma"):
            chunks = (int(n_b) + int(kc) - 1 + 4 + 3 + 15) // 16       # sample columns, Chebyshev orders >= 1, one-hot, 3 pieces of c_j
            rows = (G + 127) // 128 * 128
            flops = 2.0 * 128 * 256 * 16 * 3 * chunks * (rows / 128.0 / world) * ((rows / 128 + 1) // 2)
            ach = flops / avg_s / 1e12
            e.update(bound="tensor", achieved=ach, peak=bf16, unit="TFLOP/s", frac=ach / bf16,
                     note="bf16 tcgen05.mma 128x256x16, 3 MMAs (hi*hi, hi*lo, lo*hi) per 16-column chunk, %d chunks; peak = %s "
                          "bf16_tflops (cuBLAS burst)" % (chunks, src))
        elif cap and cap.get("fp64_flop_per_launch"):
            ach = cap["fp64_flop_per_launch"] / avg_s / 1e12
            e.update(bound="fp64", achieved=ach, peak=fp64, unit="TFLOP/s", frac=ach / fp64,
                     note="FP64 pipe: executed DFMA (2 flop) + DADD + DMUL per launch counted by ncu on the committed capture of this "
                          "workload (%s); peak = DFMA microbenchmark measured in this run (cbs_fp64_peak)" % cap.get("source", "profiles/"))
        elif name.startswith(("k_keep_mask", "k_compact_colsum", "k_counts_to_f64", "k_rowsum_sel", "k_mcg_sort_key", "k_scr_convert")):
            per = {"k_keep_mask": 4.0 * G0 * N, "k_compact_colsum": 8.0 * G * N, "k_counts_to_f64": 12.0 * G0 * N,
                   "k_rowsum_sel": 4.0 * G * N, "k_mcg_sort_key": 8.0 * G * n_b,
                   "k_scr_convert": 8.0 * ktp * G + 2.0 * 160 * G}[name.split("<")[0]]
            ach = per / avg_s / 1e9
            e.update(bound="hbm", achieved=ach, peak=hbm, unit="GB/s", frac=ach / hbm, note="algorithmic bytes per launch / time; peak = %s hbm_gbs" % src)
        else:
            e.update(bound="fp64", achieved=None, peak=fp64, unit="TFLOP/s", frac=None,
                     note="no committed ncu capture for this shape: executed-flop count unavailable")
        out.append(e)
    return out


def run_inproc(args):
    """One process, one cbs_combat_seq call sharded over args.inproc devices by the library's own host threads."""
    import torch
    pkg = graft.load_package()
    L = pkg._lib
    for kv in args.tune:
        L.set_tuning(kv.split("=")[0], float(kv.split("=")[1]))
    case, d = make_case(pkg, args.workload, args.genes)
    G0, N = case.counts.shape
    h_counts = torch.empty((N, G0), dtype=torch.int32, pin_memory=True)
    h_counts.numpy()[:] = np.ascontiguousarray(case.counts.T)
    h_adj = torch.empty((N, G0), dtype=torch.float64, pin_memory=True)
    opts = dict(shrink=case.shrink, shrink_disp=case.shrink_disp, n_gpus=args.inproc)
    for _ in range(max(args.warmup, 1)):
        res = L.combat_seq_into(h_counts.numpy().T, d.batch_idx, d.mod, h_adj.numpy().T, **opts)
    t = time.perf_counter()
    for _ in range(args.steps):
        res = L.combat_seq_into(h_counts.numpy().T, d.batch_idx, d.mod, h_adj.numpy().T, **opts)
    ms = (time.perf_counter() - t) * 1e3 / args.steps
    print(json.dumps({"metric": METRIC, "mode": "single process, cbs_input.n_gpus", "n_gpus": args.inproc, "value": G0 * N / (ms * 1e-3),
                      "unit": UNIT, "ms_per_step": ms, "steps": args.steps, "warmup": args.warmup, "config": describe(case, d),
                      "stage_ms_rank0": {k: round(v, 3) for k, v in res["stage_ms"].items()},
                      "checksum": float(np.asarray(h_adj.numpy()[:, :: max(1, G0 // 997)]).sum())}))


if __name__ == "__main__":
    a = parse_args()
    if a.inproc:
        run_inproc(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_native(a)
