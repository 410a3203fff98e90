"""Under torchrun (one rank per GPU): the gene-sharded pipeline against the single-GPU pipeline on the same inputs.
Every rank runs both on its own device and compares the G0-row range it owns."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import __graft_entry__ as g
pkg = g.load_package(); L = pkg._lib
synth = importlib.import_module("combat-seq_b200.synth")
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
uid = torch.zeros(128, dtype=torch.uint8, device=dev)
if rank == 0:
    uid.copy_(torch.frombuffer(bytearray(L.comm_unique_id()), dtype=torch.uint8))
dist.broadcast(uid, 0)
comm = L.comm_create(bytes(uid.cpu().numpy().tobytes()), rank, world, local)
ok = True
for name, kw in (("C2", dict(G=9000)), ("C3b", dict(G=5000)), ("C5", dict(G=7000, n=20, B=4))):
    case = synth.CASES[name](**kw)
    d = pkg.design.build_design(case.counts.shape[1], case.batch, case.group, case.covar_mod, case.full_mod)
    want = ("adjusted", "phi_hat", "gamma_hat", "gamma_star", "phi_star")
    one = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=case.shrink, shrink_disp=case.shrink_disp, device=local, want=want)
    sh = L.combat_seq_raw(case.counts, d.batch_idx, d.mod, shrink=case.shrink, shrink_disp=case.shrink_disp, device=local, comm=comm, want=want)
    r0, r1 = sh["row_begin"], sh["row_end"]
    same_disp = np.array_equal(one["disp_common"], sh["disp_common"])
    exact = float((one["adjusted"][r0:r1] == sh["adjusted"][r0:r1]).mean()) if r1 > r0 else 1.0
    errs = {}
    for k in want[1:]:
        a, b = sh[k][r0:r1], one[k][r0:r1]
        m = ~np.isnan(b)
        errs[k] = float(np.max(np.abs(a[m] - b[m]) / (1e-6 * np.abs(b[m]) + 1e-9))) if m.any() else 0.0
    good = same_disp and exact >= 0.9999 and max(errs.values()) <= 1.0
    ok &= good
    print("rank %d %s rows [%d,%d) of %d: disp_common identical %s, adjusted exact %.6f, err/tol %s -> %s"
          % (rank, name, r0, r1, case.counts.shape[0], same_disp, exact, {k: round(v, 4) for k, v in errs.items()}, "OK" if good else "FAIL"), flush=True)
t = torch.tensor([1 if ok else 0], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MIN)
L.comm_destroy(comm); dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 1 else 1)
