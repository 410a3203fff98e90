"""ComBat_seq() — host-side mirror of the reference's one public function (/root/reference/ComBat_seq.R:22-23).

Same arguments, defaults, progress messages and error strings as the R closure; the numeric body
(ComBat_seq.R:33-203) runs in libcombatseq.so (hand-written sm_100a kernels) through the C ABI of
include/combatseq.h.  The factor/model.matrix bookkeeping (ComBat_seq.R:25-31,53-94) stays on the host, exactly
as it stays in R in the `.Call` wrapper (combat-seq_b200/R/ComBat_seq.R).
"""
import sys

import numpy as np

from . import _lib
from .design import build_design

_ERRORS = {
    _lib.CBS_ESINGLE_SAMPLE_BATCH: ValueError, _lib.CBS_ECONF_COVAR_BATCH: ValueError,
    _lib.CBS_ECONF_COVARS: ValueError, _lib.CBS_ECONF_COVAR_WITH_BATCH: ValueError, _lib.CBS_EINVAL: ValueError,
}


def ComBat_seq(counts, batch, group=None, covar_mod=None, full_mod=True, shrink=False, shrink_disp=False,
               gene_subset_n=None, *, seed=0, mc_mode="reference", device=-1, n_gpus=1, comm=None, verbose=True,
               return_details=False):
    """Adjust a gene x sample count matrix for batch effects (negative-binomial ComBat).

    counts, batch, group, covar_mod, full_mod, shrink, shrink_disp (R: shrink.disp), gene_subset_n
    (R: gene.subset.n) have the reference's meaning.  Extras (keyword-only, defaults = reference behaviour):
    seed (gene_subset_n sampling), mc_mode "reference" | "stable", device, n_gpus (single-process sharding over
    devices 0..n_gpus-1; 0 = all visible), comm (one process per GPU, see below), verbose (the reference's cat()
    messages), return_details (also return every intermediate).
    Returns a float64 matrix with the shape of `counts` (ComBat_seq.R:199-203); pandas row/column labels are kept.
    With comm= this process is one rank of a gene-sharded job and owns rows [row_begin, row_end) only: the function
    then always returns the details dict (rows of other ranks are NaN; "row_begin"/"row_end" give the range).
    """
    def say(msg):
        if verbose:
            sys.stdout.write(msg + "\n")

    labels = None
    if hasattr(counts, "index") and hasattr(counts, "columns"):        # pandas.DataFrame: keep dimnames
        labels = (counts.index, counts.columns)
        counts = counts.to_numpy()
    counts = np.asarray(counts)                                          # as.matrix  (ComBat_seq.R:25)
    if counts.ndim != 2:
        raise ValueError("counts must be a gene x sample matrix")
    d = build_design(counts.shape[1], batch, group, covar_mod, full_mod)
    for m in d.messages:
        say(m)
    n_na = int(np.isnan(counts).sum()) if counts.dtype.kind == "f" else 0
    if n_na:                                                             # ComBat_seq.R:93-94
        say("Found %d Missing Data Values" % n_na)
        raise ValueError("missing values in counts are not supported (edgeR's DGEList rejects NA counts)")
    say("Estimating dispersions")                                       # :98
    say("Fitting the GLM model")                                        # :129
    if shrink:
        say("Apply shrinkage - computing posterior estimates for parameters")          # :143
        if gene_subset_n is not None:
            say("Using %s random genes for Monte Carlo integration" % gene_subset_n)   # helper_seq.R:18
        else:
            say("Using all genes for Monte Carlo integration; the function runs very slow for large number of genes")
        if not shrink_disp:
            say("Apply shrinkage to mean only")                         # :163
    else:
        say("Shrinkage off - using GLM estimates for parameters")       # :167
    say("Adjusting the data")                                           # :182
    want = ("adjusted", "phi_hat", "gamma_hat", "gamma_star", "phi_star", "mu_hat") if return_details else ("adjusted",)
    try:
        res = _lib.combat_seq_raw(counts, d.batch_idx, d.mod, shrink=shrink, shrink_disp=shrink_disp,
                                  gene_subset_n=gene_subset_n or 0, seed=seed,
                                  mc_mode=_lib.CBS_MC_STABLE if mc_mode == "stable" else _lib.CBS_MC_REFERENCE,
                                  device=device, comm=comm, n_gpus=n_gpus, want=want)
    except _lib.CombatSeqError as e:
        raise _ERRORS.get(e.code, RuntimeError)(str(e)) from None
    adj = res["adjusted"]
    if labels is not None:
        import pandas as pd
        adj = pd.DataFrame(adj, index=labels[0], columns=labels[1])
    if return_details or comm is not None:
        res["adjusted"] = adj
        res["design"] = d
        return res
    return adj
