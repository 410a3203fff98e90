"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/liboracle.so, the CPU restatement of /root/reference/ComBat_seq.R:22-204,
/root/reference/helper_seq.R:2-64 and the edgeR 3.22 / R 3.5 nmath routines they call.  May be imported only
by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; the product package
(combat-seq_b200/) never imports it.  PARITY UNPINNED against real R/edgeR (not installed in this image).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

f64p = C.POINTER(C.c_double)
i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
u8p = C.POINTER(C.c_uint8)


class CbsInput(C.Structure):
    _fields_ = [("G0", C.c_int32), ("N", C.c_int32), ("B", C.c_int32), ("p_mod", C.c_int32),
                ("counts_f64", f64p), ("counts_i32", i32p), ("batch_idx", i32p), ("mod", f64p),
                ("use_fallback", u8p), ("shrink", C.c_int32), ("shrink_disp", C.c_int32),
                ("gene_subset_n", C.c_int32), ("seed", C.c_uint64), ("mc_mode", C.c_int32),
                ("device", C.c_int32), ("comm", C.c_void_p), ("counts_on_device", C.c_int32),
                ("n_gpus", C.c_int32)]


class CbsOutput(C.Structure):
    _fields_ = [("adjusted", f64p), ("keep", u8p), ("lib_size", f64p), ("disp_common", f64p),
                ("phi_hat", f64p), ("gamma_hat", f64p), ("gamma_star", f64p), ("phi_star", f64p),
                ("mu_hat", f64p), ("stage_ms", f64p), ("n_mc_fallback", i64p),
                ("row_begin", C.c_int64), ("row_end", C.c_int64), ("adjusted_on_device", C.c_int32)]


def build():
    subprocess.run(["make", "-s", "-C", _HERE], check=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.ref_last_error.restype = C.c_char_p
        _LIB.ref_common_disp.restype = C.c_double
        _LIB.ref_unit_nb_deviance.restype = C.c_double
        _LIB.ref_unit_nb_deviance.argtypes = [C.c_double] * 3
        _LIB.ref_maximize_interpolant.restype = C.c_double
    return _LIB


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _p(a):
    return a.ctypes.data_as(f64p) if a is not None else None


def _elementwise(fn, *arrs):
    arrs = np.broadcast_arrays(*[np.asarray(a, dtype=np.float64) for a in arrs])
    shape = arrs[0].shape
    arrs = [_d(a.ravel()) for a in arrs]
    out = np.empty(arrs[0].size)
    getattr(lib(), fn)(C.c_int64(out.size), *[_p(a) for a in arrs], _p(out))
    return out.reshape(shape)


def nb_pmf(x, size, mu): return _elementwise("ref_nb_pmf", x, size, mu)
def nb_cdf(x, size, mu): return _elementwise("ref_nb_cdf", x, size, mu)
def nb_cdf_bysum(x, size, mu): return _elementwise("ref_nb_cdf_bysum", x, size, mu)
def nb_quantile(p, size, mu): return _elementwise("ref_nb_quantile", p, size, mu)
def qnorm(p): return _elementwise("ref_qnorm", p)
def stirlerr(x): return _elementwise("ref_stirlerr", x)
def bd0(x, np_): return _elementwise("ref_bd0", x, np_)
def pbeta(x, a, b): return _elementwise("ref_pbeta", x, a, b)


def qr_rank(X):
    X = np.asfortranarray(X, dtype=np.float64)
    return lib().ref_qr_rank(_p(X), C.c_int(X.shape[0]), C.c_int(X.shape[1]))


def unit_nb_deviance(y, mu, phi):
    return lib().ref_unit_nb_deviance(y, mu, phi)


def maximize_interpolant(x, y):
    x = _d(x); y = _d(y)
    return lib().ref_maximize_interpolant(C.c_int(x.size), _p(x), _p(y))


def fmm_spline(x, y):
    x = _d(x); y = _d(y); n = x.size
    b = np.empty(n); c = np.empty(n); d = np.empty(n)
    lib().ref_fmm_spline(C.c_int(n), _p(x), _p(y), _p(b), _p(c), _p(d))
    return b, c, d


def _F(a, dtype=np.float64):
    return np.asfortranarray(a, dtype=dtype)


def glm_fit(y, design, disp, disp_cols, lib_size, gene_offset=None, prior_count=0.0, nthreads=1):
    y = _F(y); G, n = y.shape
    design = _F(design); p = design.shape[1]
    disp = _F(np.asarray(disp, dtype=np.float64).reshape(G, -1))
    cols = np.ascontiguousarray(disp_cols, dtype=np.int32) if disp_cols is not None else np.zeros(n, np.int32)
    lib_size = _d(lib_size)
    go = _d(gene_offset) if gene_offset is not None else None
    coef = np.empty((G, p), order="F"); fitted = np.empty((G, n), order="F")
    lib().ref_glm_fit(C.c_int64(G), C.c_int32(n), _p(y), _p(design), C.c_int32(p), _p(disp), C.c_int32(disp.shape[1]),
                      cols.ctypes.data_as(i32p), _p(lib_size), _p(go), C.c_double(prior_count), _p(coef), _p(fitted),
                      C.c_int(nthreads))
    return coef, fitted


def apl(y, design, lib_size, disp, disp_per_gene=False, nthreads=1):
    y = _F(y); G, n = y.shape
    des = _F(design) if design is not None else None
    p = des.shape[1] if des is not None else 1
    disp = _F(disp) if disp_per_gene else _d(disp)
    D = disp.shape[1] if disp_per_gene else disp.size
    out = np.empty((G, D), order="F")
    lib().ref_apl(C.c_int64(G), C.c_int32(n), _p(y), _p(des), C.c_int32(p), _p(_d(lib_size)), _p(disp), C.c_int32(D),
                  C.c_int32(1 if disp_per_gene else 0), _p(out), C.c_int(nthreads))
    return out


def common_disp(y, design=None, nthreads=1):
    y = _F(y); G, n = y.shape
    des = _F(design) if design is not None else None
    p = des.shape[1] if des is not None else 1
    ne = C.c_int(0)
    v = lib().ref_common_disp(C.c_int64(G), C.c_int32(n), _p(y), _p(des), C.c_int32(p), C.c_int(nthreads), C.byref(ne))
    return v, ne.value


def tagwise_disp(y, design, common, nthreads=1, prior_df=0.0):
    y = _F(y); G, n = y.shape
    des = _F(design) if design is not None else None
    p = des.shape[1] if des is not None else 1
    out = np.empty(G)
    if prior_df:
        lib().ref_tagwise_disp_prior(C.c_int64(G), C.c_int32(n), _p(y), _p(des), C.c_int32(p), C.c_double(common),
                                     C.c_double(prior_df), _p(out), C.c_int(nthreads))
        return out
    lib().ref_tagwise_disp(C.c_int64(G), C.c_int32(n), _p(y), _p(des), C.c_int32(p), C.c_double(common), _p(out),
                           C.c_int(nthreads))
    return out


def mc_int(dat, mu, gamma, phi, mode=0, gene_subset_n=0, seed=0, targets=None, nthreads=1):
    dat = _F(dat); mu = _F(mu); G, n = dat.shape
    gamma = _d(gamma); phi = _d(phi)
    gs = np.empty(G); ps = np.empty(G); nfb = C.c_int64(0)
    t = np.ascontiguousarray(targets, dtype=np.int64) if targets is not None else None
    lib().ref_mc_int(C.c_int64(G), C.c_int32(n), _p(dat), _p(mu), _p(gamma), _p(phi), C.c_int32(mode),
                     C.c_int32(gene_subset_n), C.c_uint64(seed), _p(gs), _p(ps), C.byref(nfb),
                     t.ctypes.data_as(i64p) if t is not None else None, C.c_int64(t.size if t is not None else 0),
                     C.c_int(nthreads))
    return gs, ps, nfb.value


def match_quantiles(counts, old_mu, old_phi, new_mu, new_phi, nthreads=1):
    counts = _F(counts); G, n = counts.shape
    out = np.empty((G, n), order="F")
    lib().ref_match_quantiles(C.c_int64(G), C.c_int32(n), _p(counts), _p(_F(old_mu)), _p(_d(old_phi)), _p(_F(new_mu)),
                              _p(_d(new_phi)), _p(out), C.c_int(nthreads))
    return out


def max_threads():
    return lib().ref_max_threads()


class OracleError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def combat_seq(counts, batch_idx, mod, shrink=False, shrink_disp=False, gene_subset_n=0, seed=0, mc_mode=0,
               use_fallback=None, nthreads=1, mc_target_stride=1):
    """Run the oracle pipeline; returns a dict of every intermediate (G0-row layout, NaN rows for dropped genes)."""
    counts = _F(counts); G0, N = counts.shape
    batch_idx = np.ascontiguousarray(batch_idx, dtype=np.int32)
    B = int(batch_idx.max()) + 1
    mod = _F(mod); p_mod = mod.shape[1]
    inp = CbsInput()
    inp.G0, inp.N, inp.B, inp.p_mod = G0, N, B, p_mod
    inp.counts_f64 = _p(counts); inp.counts_i32 = None
    inp.batch_idx = batch_idx.ctypes.data_as(i32p); inp.mod = _p(mod)
    fb = None
    if use_fallback is not None:
        fb = np.ascontiguousarray(use_fallback, dtype=np.uint8); inp.use_fallback = fb.ctypes.data_as(u8p)
    inp.shrink, inp.shrink_disp = int(shrink), int(shrink_disp)
    inp.gene_subset_n, inp.seed, inp.mc_mode = int(gene_subset_n or 0), int(seed), int(mc_mode)
    inp.device = -1; inp.comm = None; inp.counts_on_device = 0; inp.n_gpus = 1
    res = dict(adjusted=np.empty((G0, N), order="F"), keep=np.zeros(G0, np.uint8), lib_size=np.empty(N),
               disp_common=np.empty(B), phi_hat=np.empty((G0, B), order="F"), gamma_hat=np.empty((G0, B), order="F"),
               gamma_star=np.empty((G0, B), order="F"), phi_star=np.empty((G0, B), order="F"),
               mu_hat=np.empty((G0, N), order="F"), n_mc_fallback=np.zeros(B, np.int64))
    out = CbsOutput()
    for k in ("adjusted", "lib_size", "disp_common", "phi_hat", "gamma_hat", "gamma_star", "phi_star", "mu_hat"):
        setattr(out, k, _p(res[k]))
    out.keep = res["keep"].ctypes.data_as(u8p)
    out.n_mc_fallback = res["n_mc_fallback"].ctypes.data_as(i64p)
    stage = np.zeros(9); out.stage_ms = _p(stage)
    rc = lib().ref_combat_seq(C.byref(inp), C.byref(out), C.c_int(nthreads), C.c_int(mc_target_stride))
    if rc != 0:
        raise OracleError(rc, lib().ref_last_error().decode())
    res["stage_ms"] = dict(zip(("h2d", "prep", "common_disp", "tagwise_disp", "glm", "mc", "match_quantiles", "d2h", "total"),
                               stage.tolist()))
    return res
