"""Host-side design bookkeeping of ComBat_seq() — the part that stays in the wrapper.

Restates /root/reference/ComBat_seq.R:25-31 (batch factor, one-sample check), :53-81 (model matrices, dropping
all-ones columns) and :84-90 (confounding checks with the reference's stop() strings), plus the base-R pieces
they rely on: as.factor (sorted levels), model.matrix(~f) (treatment contrasts), qr()$rank (LINPACK dqrdc2,
tol = 1e-7).  Pure numpy; no device work happens here.
"""
import numpy as np

MSG_ONE_SAMPLE = "ComBat-seq doesn't support 1 sample per batch yet"                                   # ComBat_seq.R:30
MSG_CONF_COVAR_BATCH = "The covariate is confounded with batch! Remove the covariate and rerun ComBat-Seq"  # :86
MSG_CONF_COVARS = ("The covariates are confounded! Please remove one or more of the covariates so the design "
                   "is not confounded")                                                                # :88
MSG_CONF_COVAR_WITH_BATCH = ("At least one covariate is confounded with batch! Please remove confounded "
                             "covariates and rerun ComBat-Seq")                                        # :89


def as_factor(x):
    """R as.factor(): levels = sort(unique(x)); returns (levels, 0-based codes).  None -> zero levels."""
    if x is None:
        return np.array([]), np.zeros(0, dtype=np.int32)
    x = np.asarray(x)
    if x.dtype.kind in "OUS":
        x = x.astype(str)
    levels, codes = np.unique(x, return_inverse=True)
    return levels, codes.astype(np.int32)


def model_matrix_factor(codes, nlevels):
    """model.matrix(~f): intercept + treatment-contrast dummies for levels 2..L."""
    n = codes.shape[0]
    m = np.zeros((n, nlevels), dtype=np.float64)
    m[:, 0] = 1.0
    for lv in range(1, nlevels):
        m[:, lv] = (codes == lv)
    return m


def model_matrix_column(col):
    """model.matrix(~col) for one data.frame column: numeric -> [1, col]; anything else -> factor dummies."""
    col = np.asarray(col)
    if col.dtype.kind in "fiub":
        return np.column_stack([np.ones(col.shape[0]), col.astype(np.float64)])
    levels, codes = as_factor(col)
    return model_matrix_factor(codes, len(levels))


def qr_rank(X, tol=1e-7):
    """Rank as R's qr(X)$rank: LINPACK dqrdc2 (Householder QR, limited column pivoting, tol = 1e-7)."""
    x = np.array(X, dtype=np.float64, order="F", copy=True)
    if x.ndim == 1:
        x = x.reshape(-1, 1)
    n, p = x.shape
    if p == 0:
        return 0
    qraux = np.sqrt((x * x).sum(axis=0))
    w1 = qraux.copy()
    w2 = qraux.copy()
    w2[w2 == 0] = 1.0
    lup = min(n, p)
    k = p + 1
    for l in range(1, lup + 1):
        l0 = l - 1
        while l < k and qraux[l0] < w2[l0] * tol:
            x[:, l0:] = np.roll(x[:, l0:], -1, axis=1)
            for v in (qraux, w1, w2):
                v[l0:] = np.roll(v[l0:], -1)
            k -= 1
        if l == n:
            break
        nrmxl = np.sqrt((x[l0:, l0] ** 2).sum())
        if nrmxl != 0:
            if x[l0, l0] != 0:
                nrmxl = np.copysign(nrmxl, x[l0, l0])
            x[l0:, l0] /= nrmxl
            x[l0, l0] += 1.0
            for j in range(l0 + 1, p):
                t = -(x[l0:, l0] @ x[l0:, j]) / x[l0, l0]
                x[l0:, j] += t * x[l0:, l0]
                if qraux[j] != 0:
                    tt = max(1.0 - (abs(x[l0, j]) / qraux[j]) ** 2, 0.0)
                    if abs(tt) < 1e-6:
                        qraux[j] = np.sqrt((x[l0 + 1:, j] ** 2).sum())
                        w1[j] = qraux[j]
                    else:
                        qraux[j] *= np.sqrt(tt)
            qraux[l0] = x[l0, l0]
            x[l0, l0] = -nrmxl
    return min(k - 1, n)


class Design:
    """Everything the C ABI needs: batch_idx (N), mod (N x p_mod, col 0 = intercept), plus messages."""

    def __init__(self, batch_levels, batch_idx, mod, messages):
        self.batch_levels = batch_levels
        self.batch_idx = batch_idx
        self.mod = mod
        self.messages = messages

    @property
    def n_batch(self):
        return len(self.batch_levels)

    @property
    def design(self):
        """cbind(batchmod, mod) without all-ones columns (ComBat_seq.R:75-80)."""
        n = self.batch_idx.shape[0]
        bm = np.zeros((n, self.n_batch))
        bm[np.arange(n), self.batch_idx] = 1.0
        full = np.column_stack([bm, self.mod])
        check = np.all(full == 1, axis=0)
        return full[:, ~check]


def build_design(n_samples, batch, group=None, covar_mod=None, full_mod=True):
    """ComBat_seq.R:27-31 and :53-90.  Raises ValueError with the reference's stop() strings."""
    messages = []
    batch_levels, batch_idx = as_factor(batch)
    if batch_idx.shape[0] != n_samples:
        raise ValueError("length(batch) must equal ncol(counts)")
    n_batch = len(batch_levels)
    n_batches = np.bincount(batch_idx, minlength=n_batch)
    if np.any(n_batches <= 1):
        raise ValueError(MSG_ONE_SAMPLE)
    messages.append("Found %d batches" % n_batch)                                     # ComBat_seq.R:51
    g_levels, g_codes = as_factor(group)
    if full_mod and len(g_levels) > 1:
        messages.append("Using full model in ComBat-seq.")                            # :59
        if g_codes.shape[0] != n_samples:
            raise ValueError("length(group) must equal ncol(counts)")
        mod = model_matrix_factor(g_codes, len(g_levels))
    else:
        messages.append("Using null model in ComBat-seq.")                            # :62
        mod = np.ones((n_samples, 1))
    if covar_mod is not None:
        if hasattr(covar_mod, "columns") and hasattr(covar_mod, "iloc"):               # data.frame  (:67-69)
            cm = np.column_stack([model_matrix_column(covar_mod.iloc[:, i].to_numpy())
                                  for i in range(covar_mod.shape[1])])
        else:
            cm = np.asarray(covar_mod, dtype=np.float64)
            if cm.ndim == 1:
                cm = cm.reshape(-1, 1)
        if cm.shape[0] != n_samples:
            raise ValueError("nrow(covar_mod) must equal ncol(counts)")
        cm = cm[:, ~np.all(cm == 1, axis=0)]                                          # :70
        mod = np.column_stack([mod, cm])                                              # :73
    d = Design(batch_levels, batch_idx, np.asfortranarray(mod), messages)
    # The C ABI rebuilds design as [batch indicators | mod[:, 1:]]; drop any further all-ones columns of mod so
    # both sides agree with "design[, !check]" (ComBat_seq.R:76-78).
    extra_ones = np.all(mod[:, 1:] == 1, axis=0) if mod.shape[1] > 1 else np.zeros(0, dtype=bool)
    if extra_ones.any():
        keep = np.concatenate([[True], ~extra_ones])
        d.mod = np.asfortranarray(mod[:, keep])
    design = d.design
    messages.append("Adjusting for %d covariate(s) or covariate level(s)" % (design.shape[1] - n_batch))  # :81
    if qr_rank(design) < design.shape[1]:                                             # :84-90
        if design.shape[1] == n_batch + 1:
            raise ValueError(MSG_CONF_COVAR_BATCH)
        if design.shape[1] > n_batch + 1:
            if qr_rank(design[:, n_batch:]) < design.shape[1] - n_batch:
                raise ValueError(MSG_CONF_COVARS)
            raise ValueError(MSG_CONF_COVAR_WITH_BATCH)
    return d
