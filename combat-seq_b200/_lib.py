"""ctypes binding of libcombatseq.so (include/combatseq.h) — the same C ABI the R `.Call` shim binds.

There is no CPU path: if the shared library has not been built, loading fails loudly; if it is built but no CUDA
device is visible, every compute entry point returns CBS_ENOGPU and `CombatSeqError` is raised.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("COMBATSEQ_LIB") or os.path.join(HERE, "libcombatseq.so")   # (override: development builds)


def _prefer_bundled_nccl():
    """If this Python has the NCCL wheel PyTorch uses (nvidia/nccl/lib/libnccl.so.2), let the library dlopen that one
    (COMBATSEQ_NCCL) instead of an older system NCCL: whichever libnccl.so.2 is mapped first also serves a later `import torch`."""
    if os.environ.get("COMBATSEQ_NCCL"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for root in (spec.submodule_search_locations if spec else []):
            cand = os.path.join(root, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["COMBATSEQ_NCCL"] = cand
                return
    except (ImportError, ValueError, AttributeError):
        pass


_prefer_bundled_nccl()

f64p = C.POINTER(C.c_double)
i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
u8p = C.POINTER(C.c_uint8)

CBS_OK, CBS_EINVAL, CBS_ESINGLE_SAMPLE_BATCH, CBS_ECONF_COVAR_BATCH, CBS_ECONF_COVARS, \
    CBS_ECONF_COVAR_WITH_BATCH, CBS_ECUDA, CBS_ENCCL, CBS_ENOGPU, CBS_ENOMEM = range(10)
CBS_MC_REFERENCE, CBS_MC_STABLE = 0, 1
STAGES = ("h2d", "prep", "common_disp", "tagwise_disp", "glm", "mc", "match_quantiles", "d2h", "total")

# every symbol include/combatseq.h declares (tests check that the library exports all of them)
EXPORTS = ("cbs_version", "cbs_device_count", "cbs_last_error", "cbs_shutdown", "cbs_kernel_launches",
           "cbs_comm_unique_id", "cbs_comm_create", "cbs_comm_destroy", "cbs_shard_range", "cbs_combat_seq", "cbs_nb_logpmf",
           "cbs_nb_cdf", "cbs_nb_quantile", "cbs_glm_fit", "cbs_apl", "cbs_common_disp", "cbs_tagwise_disp",
           "cbs_mc_int", "cbs_match_quantiles", "cbs_profile_enable", "cbs_profile_dump", "cbs_fp64_peak",
           "cbs_fp64_mma_peak", "cbs_set_tuning", "cbs_get_diag", "cbs_tagwise_disp_prior", "cbs_simulate_counts", "cbs_mc_int_static")


class CbsInput(C.Structure):
    _fields_ = [("G0", C.c_int32), ("N", C.c_int32), ("B", C.c_int32), ("p_mod", C.c_int32),
                ("counts_f64", C.c_void_p), ("counts_i32", C.c_void_p), ("batch_idx", i32p), ("mod", f64p),
                ("use_fallback", u8p), ("shrink", C.c_int32), ("shrink_disp", C.c_int32),
                ("gene_subset_n", C.c_int32), ("seed", C.c_uint64), ("mc_mode", C.c_int32),
                ("device", C.c_int32), ("comm", C.c_void_p), ("counts_on_device", C.c_int32),
                ("n_gpus", C.c_int32)]


class CbsOutput(C.Structure):
    _fields_ = [("adjusted", C.c_void_p), ("keep", u8p), ("lib_size", f64p), ("disp_common", f64p),
                ("phi_hat", f64p), ("gamma_hat", f64p), ("gamma_star", f64p), ("phi_star", f64p),
                ("mu_hat", f64p), ("stage_ms", f64p), ("n_mc_fallback", i64p),
                ("row_begin", C.c_int64), ("row_end", C.c_int64), ("adjusted_on_device", C.c_int32)]


class CombatSeqError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


_LIB = None


def lib():
    """Load libcombatseq.so.  Raises if it has not been built (run `python -c 'import __graft_entry__ as g; g.build()'`)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: the CUDA extension has not been built and this package has no "
                              "CPU fallback.  Build it with combat-seq_b200/build.py (needs nvcc)." % LIB_PATH)
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        L.cbs_last_error.restype = C.c_char_p
        L.cbs_kernel_launches.restype = C.c_uint64
        L.cbs_combat_seq.argtypes = [C.POINTER(CbsInput), C.POINTER(CbsOutput)]
        L.cbs_comm_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.cbs_comm_destroy.argtypes = [C.c_void_p]
        L.cbs_comm_unique_id.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def check(rc):
    if rc != CBS_OK:
        raise CombatSeqError(rc, lib().cbs_last_error().decode())


def device_count():
    return lib().cbs_device_count()


def kernel_launches():
    return int(lib().cbs_kernel_launches())


def profile_enable(on=True):
    check(lib().cbs_profile_enable(int(on)))


def profile_dump():
    """{kernel name: (total device ms, launches)} since profile_enable(True)."""
    buf = C.create_string_buffer(1 << 16)
    check(lib().cbs_profile_dump(buf, C.c_size_t(len(buf))))
    out = {}
    for line in buf.value.decode().splitlines():
        name, ms, n = line.rsplit(" ", 2)
        out[name] = (float(ms), int(n))
    return out


def fp64_peak_tflops():
    v = C.c_double(0)
    check(lib().cbs_fp64_peak(C.byref(v)))
    return v.value


def fp64_mma_peak_tflops():
    v = C.c_double(0)
    check(lib().cbs_fp64_mma_peak(C.byref(v)))
    return v.value


def set_tuning(key, value):
    check(lib().cbs_set_tuning(key.encode(), C.c_double(value)))


def get_diag(key):
    v = C.c_double(0)
    check(lib().cbs_get_diag(key.encode(), C.byref(v)))
    return v.value


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _F(a, dtype=np.float64):
    return np.asfortranarray(a, dtype=dtype)


def _p(a, t=f64p):
    return a.ctypes.data_as(t) if a is not None else None


# ---- NB primitives (helper_seq.R:28,53,58) ------------------------------------------------------------------
def _elementwise(fn, *arrs):
    arrs = np.broadcast_arrays(*[np.asarray(a, dtype=np.float64) for a in arrs])
    shape = arrs[0].shape
    arrs = [_d(a.ravel()) for a in arrs]
    out = np.empty(arrs[0].size)
    check(getattr(lib(), fn)(C.c_int64(out.size), *[_p(a) for a in arrs], _p(out)))
    return out.reshape(shape)


def nb_logpmf(x, size, mu): return _elementwise("cbs_nb_logpmf", x, size, mu)
def nb_cdf(x, size, mu): return _elementwise("cbs_nb_cdf", x, size, mu)
def nb_quantile(p, size, mu): return _elementwise("cbs_nb_quantile", p, size, mu)


# ---- per-subsystem entry points ------------------------------------------------------------------------------
STATIC_SCRATCH_BYTES = 8 << 20


def mc_int_static(G, n, counts_dev, mu_dev, gamma_dev, phi_dev, gamma_star_dev, phi_star_dev, n_fallback_dev=None,
                  mc_mode=CBS_MC_REFERENCE, stream=None):
    """cbs_mc_int_static: every argument but the sizes is a device address (int); enqueues on `stream`, does not synchronise."""
    vp = lambda a: C.c_void_p(a) if a else None
    check(lib().cbs_mc_int_static(C.c_int64(G), C.c_int32(n), vp(counts_dev), vp(mu_dev), vp(gamma_dev), vp(phi_dev), C.c_int32(mc_mode),
                                  vp(gamma_star_dev), vp(phi_star_dev), vp(n_fallback_dev), vp(stream)))


def simulate_counts(base_mu, col_factor, phi_col, seed, level=None, fold=None, counts_dev_ptr=None):
    """Synthetic NB count matrix generated on the device (cbs_simulate_counts): G x N int32, column-major."""
    base_mu = _d(base_mu); col_factor = _d(col_factor); phi_col = _d(phi_col)
    G, N = base_mu.size, col_factor.size
    n_levels = 0
    if fold is not None:
        fold = _F(np.asarray(fold, dtype=np.float64).reshape(G, -1)); n_levels = fold.shape[1]
        level = np.ascontiguousarray(level, dtype=np.int32)
    out = np.empty((G, N), dtype=np.int32, order="F") if counts_dev_ptr is None else None
    check(lib().cbs_simulate_counts(C.c_int64(G), C.c_int32(N), _p(base_mu), _p(col_factor), _p(level, i32p) if fold is not None else None,
                                    C.c_int32(n_levels), _p(fold) if fold is not None else None, _p(phi_col), C.c_uint64(seed),
                                    _p(out, i32p) if out is not None else None, C.c_void_p(counts_dev_ptr) if counts_dev_ptr else None))
    return out


def glm_fit(y, design, disp, disp_cols, lib_size, gene_offset=None, prior_count=0.0):
    y = _F(y, np.int32); G, n = y.shape
    design = _F(design); p = design.shape[1]
    disp = _F(np.asarray(disp, dtype=np.float64).reshape(G, -1))
    cols = np.ascontiguousarray(disp_cols, dtype=np.int32) if disp_cols is not None else None
    go = _d(gene_offset) if gene_offset is not None else None
    coef = np.empty((G, p), order="F"); fitted = np.empty((G, n), order="F")
    check(lib().cbs_glm_fit(C.c_int64(G), C.c_int32(n), _p(y, i32p), _p(design), C.c_int32(p), _p(disp),
                            C.c_int32(disp.shape[1]), _p(cols, i32p), _p(_d(lib_size)), _p(go),
                            C.c_double(prior_count), _p(coef), _p(fitted)))
    return coef, fitted


def apl(y, design, lib_size, disp, disp_per_gene=False):
    y = _F(y, np.int32); G, n = y.shape
    des = _F(design) if design is not None else None
    p = des.shape[1] if des is not None else 1
    disp = _F(disp) if disp_per_gene else _d(disp)
    D = disp.shape[1] if disp_per_gene else disp.size
    out = np.empty((G, D), order="F")
    check(lib().cbs_apl(C.c_int64(G), C.c_int32(n), _p(y, i32p), _p(des), C.c_int32(p), _p(_d(lib_size)), _p(disp),
                        C.c_int32(D), C.c_int32(1 if disp_per_gene else 0), _p(out)))
    return out


def common_disp(y, design=None):
    y = _F(y, np.int32); G, n = y.shape
    des = _F(design) if design is not None else None
    p = des.shape[1] if des is not None else 1
    v = C.c_double(0)
    check(lib().cbs_common_disp(C.c_int64(G), C.c_int32(n), _p(y, i32p), _p(des), C.c_int32(p), C.byref(v)))
    return v.value


def tagwise_disp(y, design, common, prior_df=0.0):
    y = _F(y, np.int32); G, n = y.shape
    des = _F(design) if design is not None else None
    p = des.shape[1] if des is not None else 1
    out = np.empty(G)
    if prior_df:
        check(lib().cbs_tagwise_disp_prior(C.c_int64(G), C.c_int32(n), _p(y, i32p), _p(des), C.c_int32(p), C.c_double(common),
                                           C.c_double(prior_df), _p(out)))
        return out
    check(lib().cbs_tagwise_disp(C.c_int64(G), C.c_int32(n), _p(y, i32p), _p(des), C.c_int32(p), C.c_double(common),
                                 _p(out)))
    return out


def mc_int(dat, mu, gamma, phi, mode=CBS_MC_REFERENCE, gene_subset_n=0, seed=0):
    dat = _F(dat, np.int32); mu = _F(mu); G, n = dat.shape
    gs = np.empty(G); ps = np.empty(G); nfb = C.c_int64(0)
    check(lib().cbs_mc_int(C.c_int64(G), C.c_int32(n), _p(dat, i32p), _p(mu), _p(_d(gamma)), _p(_d(phi)),
                           C.c_int32(mode), C.c_int32(gene_subset_n or 0), C.c_uint64(seed), _p(gs), _p(ps),
                           C.byref(nfb)))
    return gs, ps, nfb.value


def match_quantiles(counts, old_mu, old_phi, new_mu, new_phi):
    counts = _F(counts, np.int32); G, n = counts.shape
    out = np.empty((G, n), order="F")
    check(lib().cbs_match_quantiles(C.c_int64(G), C.c_int32(n), _p(counts, i32p), _p(_F(old_mu)), _p(_d(old_phi)),
                                    _p(_F(new_mu)), _p(_d(new_phi)), _p(out)))
    return out


# ---- communicator (one process per GPU) ------------------------------------------------------------------------
def comm_unique_id():
    buf = (C.c_char * 128)()
    check(lib().cbs_comm_unique_id(buf))
    return bytes(buf)


def comm_create(uid, rank, world, device):
    buf = (C.c_char * 128).from_buffer_copy(uid)
    h = C.c_void_p()
    check(lib().cbs_comm_create(buf, rank, world, device, C.byref(h)))
    return h


def shard_range(G, rank, world):
    """Kept-gene rows [g0, g1) of `rank` (cbs_shard_range; host-only)."""
    a, b = C.c_int64(0), C.c_int64(0)
    check(lib().cbs_shard_range(C.c_int64(G), C.c_int32(rank), C.c_int32(world), C.byref(a), C.byref(b)))
    return a.value, b.value


def comm_destroy(h):
    if h:
        lib().cbs_comm_destroy(h)


# ---- the whole pipeline ------------------------------------------------------------------------------------------
def combat_seq_raw(counts, batch_idx, mod, shrink=False, shrink_disp=False, gene_subset_n=0, seed=0,
                   mc_mode=CBS_MC_REFERENCE, use_fallback=None, device=-1, comm=None, n_gpus=1, want=("adjusted",),
                   counts_dev_ptr=None, adjusted_dev_ptr=None, shape=None, adjusted_out=None):
    """cbs_combat_seq().  `want` selects the optional outputs; returns a dict.  counts may be float64 or int32
    (column-major is used as is; anything else is copied).  counts_dev_ptr/adjusted_dev_ptr (+ shape) run the
    device-resident variant used by bench.py's kernel-only timing.  n_gpus: 1 = one device, k > 1 = single-process
    sharding over devices 0..k-1 (one host thread each inside the library), 0 = all visible devices.  With comm (one
    process per GPU) only rows [row_begin, row_end) of the matrix outputs are this rank's; the rest is NaN."""
    inp = CbsInput(); out = CbsOutput()
    if counts_dev_ptr is not None:
        G0, N = shape
        inp.counts_i32 = C.c_void_p(counts_dev_ptr); inp.counts_f64 = None; inp.counts_on_device = 1
        hold = None
    else:
        counts = np.asarray(counts)
        if counts.dtype == np.int32:
            hold = _F(counts, np.int32); inp.counts_i32 = hold.ctypes.data_as(C.c_void_p); inp.counts_f64 = None
        else:
            hold = _F(counts, np.float64); inp.counts_f64 = hold.ctypes.data_as(C.c_void_p); inp.counts_i32 = None
        G0, N = hold.shape
        inp.counts_on_device = 0
    batch_idx = np.ascontiguousarray(batch_idx, dtype=np.int32)
    B = int(batch_idx.max()) + 1
    mod = _F(mod)
    inp.G0, inp.N, inp.B, inp.p_mod = G0, N, B, mod.shape[1]
    inp.batch_idx = _p(batch_idx, i32p); inp.mod = _p(mod)
    fb = None
    if use_fallback is not None:
        fb = np.ascontiguousarray(use_fallback, dtype=np.uint8); inp.use_fallback = _p(fb, u8p)
    inp.shrink, inp.shrink_disp = int(bool(shrink)), int(bool(shrink_disp))
    inp.gene_subset_n, inp.seed, inp.mc_mode = int(gene_subset_n or 0), int(seed), int(mc_mode)
    inp.device = int(device); inp.comm = comm; inp.n_gpus = int(n_gpus)
    res = {}
    shapes = dict(adjusted=(G0, N), phi_hat=(G0, B), gamma_hat=(G0, B), gamma_star=(G0, B), phi_star=(G0, B),
                  mu_hat=(G0, N))
    for k in want:
        if k in shapes:
            if k == "adjusted" and adjusted_dev_ptr is not None:
                continue
            if k == "adjusted" and adjusted_out is not None:
                assert adjusted_out.flags.f_contiguous and adjusted_out.dtype == np.float64 and adjusted_out.shape == (G0, N)
                res[k] = adjusted_out
            else:
                res[k] = np.full(shapes[k], np.nan, order="F") if comm is not None else np.empty(shapes[k], order="F")
            setattr(out, k, res[k].ctypes.data_as(C.c_void_p) if k == "adjusted" else _p(res[k]))
    if adjusted_dev_ptr is not None:
        out.adjusted = C.c_void_p(adjusted_dev_ptr); out.adjusted_on_device = 1
    res["keep"] = np.zeros(G0, np.uint8); out.keep = _p(res["keep"], u8p)
    res["lib_size"] = np.empty(N); out.lib_size = _p(res["lib_size"])
    res["disp_common"] = np.empty(B); out.disp_common = _p(res["disp_common"])
    res["n_mc_fallback"] = np.zeros(B, np.int64); out.n_mc_fallback = _p(res["n_mc_fallback"], i64p)
    stage = np.zeros(len(STAGES)); out.stage_ms = _p(stage)
    check(lib().cbs_combat_seq(C.byref(inp), C.byref(out)))
    res["stage_ms"] = dict(zip(STAGES, stage.tolist()))
    res["row_begin"], res["row_end"] = int(out.row_begin), int(out.row_end)
    return res


def combat_seq_into(counts, batch_idx, mod, adjusted_out, **kw):
    """cbs_combat_seq() writing the adjusted matrix into a caller-owned column-major float64 buffer (e.g. pinned)."""
    return combat_seq_raw(counts, batch_idx, mod, want=("adjusted",), adjusted_out=adjusted_out, **kw)
