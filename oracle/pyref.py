"""ORACLE (second, independent restatement) — TEST INFRASTRUCTURE ONLY.

A plain numpy / scipy restatement of /root/reference/ComBat_seq.R:22-204 and /root/reference/helper_seq.R:2-64
for SMALL cases (pure-Python loops over genes).  It shares no code with oracle/ref_*.cpp: the NB distribution
functions come from scipy.stats / scipy.special (not from the nmath restatement), the GLM fits are solved by
the same published edgeR algorithms but written independently in numpy.  tests/ use it (i) to cross-check the
C++ oracle on the README configurations and (ii) through tests/golden/make_golden.py, which froze its outputs as
golden vectors.  PARITY UNPINNED against real R/edgeR (not installed in this image).

Cited edgeR 3.22 / R 3.5 algorithms: SURVEY.md Appendix A (edgeR and nmath are third-party and not under
/root/reference).
"""
import math

import numpy as np
from scipy import special, stats

EPS = np.finfo(float).eps


# ---- edgeR: one-group Newton (src/glm_one_group.cpp) ---------------------------------------------------------
def glm_one_group(y, offset, disp, maxit=50, tol=1e-10):
    nz = y > 1e-10
    if not nz.any():
        return -np.inf
    beta = math.log((y[nz] / np.exp(offset[nz])).sum() / y.size)
    for _ in range(maxit):
        mu = np.exp(beta + offset)
        den = 1 + mu * disp
        step = ((y - mu) / den).sum() / (mu / den).sum()
        beta += step
        if abs(step) < tol:
            break
    return beta


def unit_dev(y, mu, phi):                                   # src/nbdev.cpp
    y = y + 1e-8
    mu = mu + 1e-8
    out = np.empty_like(mu)
    for k in range(mu.size):
        if phi[k] < 1e-4:
            r = y[k] - mu[k]
            out[k] = 2 * (y[k] * math.log(y[k] / mu[k]) - r - 0.5 * r * r * phi[k] * (1 + phi[k] * (0 * r - y[k])))   # edgeR 3.22: 2/3 is C++ integer division
        elif mu[k] * phi[k] > 1e6:
            out[k] = 2 * ((y[k] - mu[k]) / mu[k] - math.log(y[k] / mu[k])) * mu[k] / (1 + mu[k] * phi[k])
        else:
            iv = 1 / phi[k]
            out[k] = 2 * (y[k] * math.log(y[k] / mu[k]) + (y[k] + iv) * math.log((mu[k] + iv) / (y[k] + iv)))
    return out


def design_groups(X):
    """designAsFactor: group index of each row, levels ordered by first appearance is NOT what edgeR does (it sorts
    the hash), but the coefficients recovered below do not depend on the level order."""
    rows = [tuple(r) for r in X]
    lv = sorted(set(rows))
    return np.array([lv.index(r) for r in rows]), len(lv)


def glm_levenberg(y, X, disp, offset, maxit=250, tol=1e-6):       # src/glm_levenberg.cpp, null start
    n, p = X.shape
    if y.max() < 1e-10:
        return np.full(p, np.nan), np.zeros(n)
    N_s = np.exp(offset)
    w = N_s / (1 + disp * N_s)
    c = math.log((y * w / N_s).sum() / w.sum())
    beta = np.linalg.lstsq(X, np.full(n, c), rcond=None)[0]
    mu = np.exp(offset + X @ beta)
    dev = unit_dev(y, mu, disp).sum()
    ymax = y.max()
    max_info, lam = -1.0, 0.0
    for it in range(1, maxit + 1):
        den = 1 + mu * disp
        W = mu / den
        z = (y - mu) / den
        XtWX = X.T @ (X * W[:, None])
        dl = X.T @ z
        max_info = max(max_info, XtWX.diagonal().max())
        if it == 1:
            lam = max(max_info * 1e-6, 1e-13)
        lev, low_dev, failed = 0, False, False
        while True:
            lev += 1
            while True:
                try:
                    Lc = np.linalg.cholesky(XtWX + lam * np.eye(p))
                    break
                except np.linalg.LinAlgError:
                    lam = max(lam * 10, 1e-100)
            dbeta = np.linalg.solve(Lc.T, np.linalg.solve(Lc, dl))
            beta_new = beta + dbeta
            mu_new = np.exp(offset + X @ beta_new)
            dev_new = unit_dev(y, mu_new, disp).sum()
            if dev_new / ymax < 1e-13:
                low_dev = True
            if dev_new <= dev or low_dev:
                beta, mu, dev = beta_new, mu_new, dev_new
                break
            lam = max(lam * 2, 1e-100)
            if lam / max_info > 1e13:
                failed = True
                break
        if failed or low_dev:
            break
        if dl @ dbeta < tol:
            break
        if lev == 1:
            lam /= 10
    return beta, mu


def glm_fit_plain(Y, X, disp, offset):
    """glmFit.default(prior.count = 0): (coef G x p, fitted G x n).  disp, offset are G x n."""
    G, n = Y.shape
    p = X.shape[1]
    grp, nlev = design_groups(X)
    coef = np.empty((G, p))
    fitted = np.empty((G, n))
    if nlev == p:                                            # mglmOneWay shortcut
        first = np.array([np.flatnonzero(grp == k)[0] for k in range(nlev)])
        Xg = X[first]
        for g in range(G):
            bg = np.empty(nlev)
            for k in range(nlev):
                m = grp == k
                bg[k] = max(glm_one_group(Y[g, m], offset[g, m], disp[g, m]), -1e8)
            fitted[g] = np.exp(bg[grp] + offset[g])
            coef[g] = np.linalg.solve(Xg, bg)
    else:
        for g in range(G):
            coef[g], fitted[g] = glm_levenberg(Y[g], X, disp[g], offset[g])
    return coef, fitted


def glm_fit(Y, X, disp, offset, prior_count):
    """glmFit.default: coefficients from the predFC refit when prior.count > 0 (R/glmfit.R, R/predFC.R,
    src/add_prior.cpp), fitted values from the plain fit."""
    coef, fitted = glm_fit_plain(Y, X, disp, offset)
    if prior_count > 0:
        lib = np.exp(offset)
        prior = prior_count * lib / lib.mean(axis=1, keepdims=True)
        coef, _ = glm_fit_plain(Y + prior, X, disp, np.log(lib + 2 * prior))
    return coef, fitted


def adjusted_profile_lik(phi, Y, X, offset):                   # R/adjustedProfileLik.R, src/adj_coxreid.cpp
    G, n = Y.shape
    disp = np.full((G, n), phi) if np.isscalar(phi) else np.repeat(np.asarray(phi)[:, None], n, 1)
    _, mu = glm_fit_plain(Y, X, disp, offset)
    out = np.empty(G)
    for g in range(G):
        m = mu[g] != 0
        r = 1 / disp[g, m]
        y, mg = Y[g, m], mu[g, m]
        ll = (y * np.log(mg) - y * np.log(mg + r) + r * np.log(r) - r * np.log(mg + r)
              + special.gammaln(y + r) - special.gammaln(y + 1) - special.gammaln(r)).sum()
        W = np.zeros(n)
        W[m] = mg / (1 + disp[g, m] * mg)
        A = X.T @ (X * W[:, None])
        if X.shape[1] == 1:
            adj = 0.5 * math.log(abs(A[0, 0]))
        else:
            # LDL' of an SPD matrix: D = squares of the Cholesky diagonal; tiny pivots clamp to log(1e-10)
            try:
                dd = np.linalg.cholesky(A).diagonal() ** 2
            except np.linalg.LinAlgError:
                dd = np.linalg.eigvalsh(A)
            adj = 0.5 * sum(math.log(d) if (np.isfinite(d) and d >= 1e-10) else math.log(1e-10) for d in dd)
        out[g] = ll - adj
    return out


def brent_fmin(ax, bx, f, tol):                                # R optimize(): Brent_fmin
    c = (3 - math.sqrt(5)) * 0.5
    eps = math.sqrt(EPS)
    a, b = ax, bx
    v = w = x = a + c * (b - a)
    d = e = 0.0
    fx = fv = fw = f(x)
    tol3 = tol / 3
    while True:
        xm = (a + b) * 0.5
        tol1 = eps * abs(x) + tol3
        t2 = tol1 * 2
        if abs(x - xm) <= t2 - (b - a) * 0.5:
            break
        p = q = r = 0.0
        if abs(e) > tol1:
            r = (x - w) * (fx - fv)
            q = (x - v) * (fx - fw)
            p = (x - v) * q - (x - w) * r
            q = (q - r) * 2
            if q > 0:
                p = -p
            else:
                q = -q
            r = e
            e = d
        if abs(p) >= abs(q * 0.5 * r) or p <= q * (a - x) or p >= q * (b - x):
            e = (b - x) if x < xm else (a - x)
            d = c * e
        else:
            d = p / q
            u = x + d
            if u - a < t2 or b - u < t2:
                d = tol1 if x < xm else -tol1
        u = x + (d if abs(d) >= tol1 else (tol1 if d > 0 else -tol1))
        fu = f(u)
        if fu <= fx:
            if u < x:
                b = x
            else:
                a = x
            v, w, x = w, x, u
            fv, fw, fx = fw, fx, fu
        else:
            if u < x:
                a = u
            else:
                b = u
            if fu <= fw or w == x:
                v, fv, w, fw = w, fw, u, fu
            elif fu <= fv or v == x or v == w:
                v, fv = u, fu
    return x


def common_disp(Y, X):                                         # R/dispCoxReid.R via estimateGLMCommonDisp
    off = np.log(Y.sum(axis=0))
    keep = Y.sum(axis=1) >= 5
    Ys = Y[keep]
    offm = np.repeat(off[None, :], Ys.shape[0], 0)
    x = brent_fmin(0.0, 4 ** 0.25, lambda d: -adjusted_profile_lik(d ** 4, Ys, X, offm).sum(), 1e-5)
    return x ** 4


def fmm_spline(x, y):                                          # R's fmm_spline (src/fmm_spline.c), n >= 4
    n = len(x)
    b = np.zeros(n); c = np.zeros(n); d = np.zeros(n)
    nm1 = n - 1
    d[0] = x[1] - x[0]
    c[1] = (y[1] - y[0]) / d[0]
    for i in range(1, nm1):
        d[i] = x[i + 1] - x[i]
        b[i] = 2 * (d[i - 1] + d[i])
        c[i + 1] = (y[i + 1] - y[i]) / d[i]
        c[i] = c[i + 1] - c[i]
    b[0] = -d[0]
    b[nm1] = -d[nm1 - 1]
    c[0] = c[nm1] = 0.0
    if n > 3:
        c[0] = c[2] / (x[3] - x[1]) - c[1] / (x[2] - x[0])
        c[nm1] = c[nm1 - 1] / (x[nm1] - x[nm1 - 2]) - c[nm1 - 2] / (x[nm1 - 1] - x[nm1 - 3])
        c[0] = c[0] * d[0] * d[0] / (x[3] - x[0])
        c[nm1] = -c[nm1] * d[nm1 - 1] * d[nm1 - 1] / (x[nm1] - x[nm1 - 3])
    for i in range(1, n):
        t = d[i - 1] / b[i - 1]
        b[i] -= t * d[i - 1]
        c[i] -= t * c[i - 1]
    c[nm1] /= b[nm1]
    for i in range(nm1 - 1, -1, -1):
        c[i] = (c[i] - d[i] * c[i + 1]) / b[i]
    b[nm1] = (y[nm1] - y[nm1 - 1]) / d[nm1 - 1] + d[nm1 - 1] * (c[nm1 - 1] + 2 * c[nm1])
    for i in range(nm1):
        b[i] = (y[i + 1] - y[i]) / d[i] - d[i] * (c[i + 1] + 2 * c[i])
        d[i] = (c[i + 1] - c[i]) / d[i]
        c[i] = 3 * c[i]
    c[nm1] = 3 * c[nm1]
    d[nm1] = d[nm1 - 1]
    return b, c, d


def maximize_interpolant(x, y):                                # src/interpolator.cpp
    k = int(np.argmax(y))
    xmax, best = x[k], y[k]
    b, c, d = fmm_spline(x, y)
    for seg in (k - 1, k):
        if seg < 0 or seg >= len(x) - 1:
            continue
        A, Bq, Cq = 3 * d[seg], 2 * c[seg], b[seg]
        disc = Bq * Bq - 4 * A * Cq
        if disc >= 0 and A != 0:
            t = (-Bq - math.sqrt(disc)) / (2 * A)
            if 0 < t < x[seg + 1] - x[seg]:
                val = ((d[seg] * t + c[seg]) * t + b[seg]) * t + y[seg]
                if val > best:
                    best, xmax = val, x[seg] + t
    return xmax


def ave_log_cpm(Y, lib):                                       # R/aveLogCPM.R: prior.count 2, dispersion 0.05
    prior = 2.0 * lib / lib.mean()
    off = np.log(lib + 2 * prior)
    return np.array([(glm_one_group(y + prior, off, np.full(len(lib), 0.05)) + np.log(1e6)) / np.log(2) for y in Y])


def moving_average_by_col(x, width):                           # R/movingAverageByCol.R, full.length = TRUE
    n = x.shape[0]
    if width <= 1:
        return x
    width = min(width, n)
    half1, half2 = -(-width // 2), width // 2
    z = np.vstack([np.zeros((half1, x.shape[1])), x, np.zeros((half2, x.shape[1]))])
    cs = np.cumsum(z, axis=0)
    out = cs[width:] - cs[:-width]
    w = np.full(n, float(width))
    if half1 > 1:
        w[:half1 - 1] = width - np.arange(half1 - 1, 0, -1)
    if half2 > 0:
        w[n - half2:] = width - np.arange(1, half2 + 1)
    return out / w[:, None]


def tagwise_disp(Y, X, common, prior_df=0.0):                  # dispCoxReidInterpolateTagwise; ComBat-seq passes prior.df = 0
    G, n = Y.shape
    lib = Y.sum(axis=0).astype(float)
    off = np.repeat(np.log(lib)[None, :], G, 0)
    out = np.full(G, common)
    sel = Y.sum(axis=1) >= 5
    pts = np.linspace(-6, 6, 11)
    apl = np.stack([adjusted_profile_lik(common * 2 ** t, Y[sel], X, off[sel]) for t in pts], axis=1)
    if prior_df > 0:                                           # weighted-likelihood EB moderation towards the abundance trend
        prior_n = prior_df / (n - X.shape[1])
        span = (10.0 / G) ** 0.23 if G > 10 else 1.0
        width = int(np.floor(span * sel.sum()))
        o = np.argsort(ave_log_cpm(Y[sel].astype(float), lib), kind="stable")
        smooth = np.empty_like(apl)
        smooth[o] = moving_average_by_col(apl[o], width)
        apl = (apl + prior_n * smooth) / (1 + prior_n)
    out[sel] = [common * 2 ** maximize_interpolant(pts, a) for a in apl]
    return out


# ---- R nmath through scipy -------------------------------------------------------------------------------------
def pnbinom(q, mu, size):
    return stats.nbinom.cdf(q, size, size / (size + mu))


def qnbinom(p, mu, size):
    """R 3.5 qnbinom: smallest y >= 0 with pnbinom(y) >= p*(1 - 64 eps)."""
    prob = size / (size + mu)
    pf = p * (1 - 64 * EPS)
    y = max(0.0, float(stats.nbinom.ppf(pf, size, prob)))
    while y > 0 and stats.nbinom.cdf(y - 1, size, prob) >= pf:
        y -= 1
    while stats.nbinom.cdf(y, size, prob) < pf:
        y += 1
    return y


def match_quantiles(counts_sub, old_mu, old_phi, new_mu, new_phi):         # helper_seq.R:46-64
    new = np.empty_like(counts_sub, dtype=float)
    for a in range(counts_sub.shape[0]):
        for b in range(counts_sub.shape[1]):
            x = counts_sub[a, b]
            if x <= 1:
                new[a, b] = x
                continue
            tmp_p = pnbinom(x - 1, old_mu[a, b], 1 / old_phi[a])
            if abs(tmp_p - 1) < 1e-4:
                new[a, b] = x
            else:
                new[a, b] = 1 + qnbinom(tmp_p, new_mu[a, b], 1 / new_phi[a])
    return new


def monte_carlo_int_NB(dat, mu, gamma, phi):                   # helper_seq.R:8-42, gene.subset.n = NULL
    G = dat.shape[0]
    gs = np.empty(G); ps = np.empty(G)
    for i in range(G):
        m = np.arange(G) != i
        LH = np.array([np.prod(stats.nbinom.pmf(dat[i], 1 / phi[j], (1 / phi[j]) / (1 / phi[j] + mu[j])))
                       for j in np.flatnonzero(m)])
        LH[np.isnan(LH)] = 0
        if LH.sum() == 0 or np.isnan(LH.sum()):
            gs[i], ps[i] = gamma[i], phi[i]
        else:
            gs[i] = (gamma[m] * LH).sum() / LH.sum()
            ps[i] = (phi[m] * LH).sum() / LH.sum()
    return gs, ps


# ---- ComBat_seq.R:22-204 ---------------------------------------------------------------------------------------
def qr_rank(X):
    return np.linalg.matrix_rank(X, tol=1e-7 * max(1.0, np.abs(X).max()))


def combat_seq(counts, batch_idx, mod, shrink=False, shrink_disp=False):
    counts = np.asarray(counts, dtype=float)
    batch_idx = np.asarray(batch_idx)
    G0, N = counts.shape
    B = int(batch_idx.max()) + 1
    batches = [np.flatnonzero(batch_idx == b) for b in range(B)]
    keep = np.ones(G0, bool)
    for cols in batches:                                                    # ComBat_seq.R:33-37
        keep &= (counts[:, cols] != 0).any(axis=1)
    Y = counts[keep]
    G = Y.shape[0]
    lib = Y.sum(axis=0)                                                     # DGEList  :43
    batchmod = np.zeros((N, B))
    batchmod[np.arange(N), batch_idx] = 1
    design = np.column_stack([batchmod, mod[:, 1:]])                        # :75-78
    p_mod = mod.shape[1]
    disp_common = np.empty(B)
    phi_hat = np.empty((G, B))
    for b, cols in enumerate(batches):                                      # :99-119
        modb = mod[cols]
        fallback = len(cols) <= p_mod or qr_rank(modb) < p_mod
        Xb = np.ones((len(cols), 1)) if fallback else modb
        disp_common[b] = common_disp(Y[:, cols], Xb)
        phi_hat[:, b] = disp_common[b] if fallback else tagwise_disp(Y[:, cols], Xb, disp_common[b])
    phi_matrix = phi_hat[:, batch_idx]                                      # :122-125
    off = np.repeat(np.log(lib)[None, :], G, 0)
    coef1, _ = glm_fit(Y, design, phi_matrix, off, 1e-4)                    # :130
    alpha = coef1[:, :B] @ (np.array([len(c) for c in batches]) / N)        # :131
    off2 = off + alpha[:, None]                                             # :132-133
    coef2, mu_hat = glm_fit(Y, design, phi_matrix, off2, 1e-4)              # :134
    gamma_hat = coef2[:, :B]
    if shrink:                                                              # :141-165
        gamma_star = np.empty((G, B)); phi_star_mat = np.empty((G, B))
        for b, cols in enumerate(batches):
            gamma_star[:, b], phi_star_mat[:, b] = monte_carlo_int_NB(Y[:, cols], mu_hat[:, cols], gamma_hat[:, b], phi_hat[:, b])
        if not shrink_disp:
            phi_star_mat = phi_hat
    else:
        gamma_star, phi_star_mat = gamma_hat, phi_hat
    mu_star = np.exp(np.log(mu_hat) - gamma_star[:, batch_idx])             # :174-177
    phi_star = phi_star_mat.mean(axis=1)                                    # :178
    adjust = np.empty((G, N))
    for b, cols in enumerate(batches):                                      # :183-193
        adjust[:, cols] = match_quantiles(Y[:, cols], mu_hat[:, cols], phi_hat[:, b], mu_star[:, cols], phi_star)
    adjusted = counts.copy()                                                # :198-203
    adjusted[keep] = adjust

    def scat(m):
        o = np.full((G0, m.shape[1]), np.nan)
        o[keep] = m
        return o
    return dict(adjusted=adjusted, keep=keep.astype(np.uint8), lib_size=lib, disp_common=disp_common,
                phi_hat=scat(phi_hat), gamma_hat=scat(gamma_hat), gamma_star=scat(gamma_star),
                phi_star=scat(phi_star_mat), mu_hat=scat(mu_hat))
