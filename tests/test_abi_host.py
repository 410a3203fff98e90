"""not-gpu tier: the drop-in boundary without a device.

* libcombatseq.so loads and exports every symbol include/combatseq.h declares;
* with no CUDA device every compute entry point fails loudly with CBS_ENOGPU (there is no CPU fallback);
* argument validation happens before any device work and keeps the reference's stop() strings
  (/root/reference/ComBat_seq.R:29-31,84-90);
* the host mirror of ComBat_seq() (combat-seq_b200/api.py, design.py) reproduces the reference's factor / model.matrix
  bookkeeping, messages and errors (ComBat_seq.R:25-31,51-94).
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "combatseq.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cbs_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    names = declared_symbols()
    assert len(names) >= 25 and "cbs_combat_seq" in names and "cbs_shard_range" in names
    lib = pkg._lib.lib()
    for n in names:
        assert hasattr(lib, n), "libcombatseq.so does not export %s" % n
    assert set(names) == set(pkg._lib.EXPORTS), "the ctypes binding and the header list different entry points"
    assert lib.cbs_version() == 200


def test_struct_layout_matches_header(pkg):
    # the ctypes mirrors must have the C layout of cbs_input / cbs_output (x86-64 SysV): spot-check sizes and offsets
    I, O = pkg._lib.CbsInput, pkg._lib.CbsOutput
    assert I.G0.offset == 0 and I.counts_f64.offset == 16 and I.batch_idx.offset == 32 and I.use_fallback.offset == 48
    assert I.shrink.offset == 56 and I.seed.offset == 72 and I.comm.offset == 88 and C.sizeof(I) == 104
    assert O.adjusted.offset == 0 and O.stage_ms.offset == 72 and O.row_begin.offset == 88 and C.sizeof(O) == 112


needs_no_gpu = pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="this test describes the device-less behaviour")


@needs_no_gpu
def test_no_cpu_fallback(pkg):
    L = pkg._lib
    assert L.device_count() == 0
    rng = np.random.default_rng(0)
    counts = rng.negative_binomial(10, 0.1, size=(30, 8))
    n0 = L.kernel_launches()
    with pytest.raises(RuntimeError, match="no CUDA device"):
        pkg.ComBat_seq(counts, [1, 1, 1, 1, 2, 2, 2, 2], verbose=False)
    for call in (lambda: L.nb_cdf(np.arange(4.0), 2.0, 3.0),
                 lambda: L.match_quantiles(counts[:, :4], np.ones((30, 4)), np.ones(30), np.ones((30, 4)), np.ones(30)),
                 lambda: L.common_disp(counts[:, :4]),
                 lambda: L.mc_int(counts[:, :4], np.ones((30, 4)), np.zeros(30), np.ones(30)),
                 lambda: L.fp64_peak_tflops()):
        with pytest.raises(L.CombatSeqError) as e:
            call()
        assert e.value.code == L.CBS_ENOGPU
    assert L.kernel_launches() == n0


def test_c_abi_validation_uses_reference_messages(pkg):
    L = pkg._lib
    rng = np.random.default_rng(0)
    x = rng.negative_binomial(10, 0.1, size=(20, 6)).astype(np.float64)
    ones = np.ones((6, 1))

    def code(batch, mod):
        with pytest.raises(L.CombatSeqError) as e:
            L.combat_seq_raw(x, batch, mod)
        return e.value.code, str(e.value)
    c, msg = code([0, 0, 0, 1, 1, 2], ones)
    assert c == L.CBS_ESINGLE_SAMPLE_BATCH and msg == "ComBat-seq doesn't support 1 sample per batch yet"
    c, msg = code([0, 0, 0, 1, 1, 1], np.column_stack([np.ones(6), [0, 0, 0, 1, 1, 1]]))
    assert c == L.CBS_ECONF_COVAR_BATCH and msg.startswith("The covariate is confounded with batch!")
    c, msg = code([0, 0, 0, 1, 1, 1], np.column_stack([np.ones(6), [0, 1, 0, 1, 0, 1], [0, 1, 0, 1, 0, 1]]))
    assert c == L.CBS_ECONF_COVARS and msg.startswith("The covariates are confounded!")
    c, msg = code([0, 0, 0, 1, 1, 1], np.column_stack([np.ones(6), [0, 1, 0, 1, 0, 1], [0, 0, 0, 1, 1, 1]]))
    assert c == L.CBS_ECONF_COVAR_WITH_BATCH and msg.startswith("At least one covariate is confounded with batch!")
    c, msg = code([0, 0, 0, 1, 1, 1], np.column_stack([np.zeros(6)]))
    assert c == L.CBS_EINVAL and "intercept" in msg
    c, msg = code([0, 0, 0, 0, 0, 0], ones)
    assert c == L.CBS_EINVAL


def test_shard_range_partitions_in_reduction_chunks(pkg):
    L = pkg._lib
    for G in (1, 1023, 1024, 1025, 59933, 60000, 200000):
        for world in (1, 2, 3, 4, 8):
            edges = [L.shard_range(G, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == G
            for (a0, a1), (b0, b1) in zip(edges, edges[1:]):
                assert a1 == b0 and a0 <= a1
            for g0, g1 in edges:
                assert g0 % 128 == 0 or g0 == G                 # shard starts on a reduction-chunk boundary
    with pytest.raises(L.CombatSeqError):
        L.shard_range(10, 2, 2)


# ---- host mirror of the R wrapper ---------------------------------------------------------------------------------
def test_design_matches_model_matrix_semantics(pkg):
    D = pkg.design
    d = D.build_design(8, [1, 1, 1, 1, 2, 2, 2, 2])                              # README.md:36-38, group = NULL
    assert d.messages == ["Found 2 batches", "Using null model in ComBat-seq.",
                          "Adjusting for 0 covariate(s) or covariate level(s)"]
    assert d.mod.shape == (8, 1) and d.design.shape == (8, 2)
    d = D.build_design(8, ["b", "b", "a", "a", "b", "a", "b", "a"], group=[0, 1, 0, 1, 0, 1, 0, 1])
    assert list(d.batch_levels) == ["a", "b"] and d.batch_idx.tolist() == [1, 1, 0, 0, 1, 0, 1, 0]   # sorted levels
    assert "Using full model in ComBat-seq." in d.messages and d.mod[:, 1].tolist() == [0, 1, 0, 1, 0, 1, 0, 1]
    assert d.messages[-1] == "Adjusting for 1 covariate(s) or covariate level(s)"
    # full_mod = FALSE, or a one-level group, gives the null model (ComBat_seq.R:57-64)
    assert D.build_design(8, [1] * 4 + [2] * 4, group=[0, 1] * 4, full_mod=False).mod.shape == (8, 1)
    assert D.build_design(8, [1] * 4 + [2] * 4, group=[3] * 8).mod.shape == (8, 1)
    # covar_mod: all-ones columns are dropped, a single remaining column survives
    cov = np.column_stack([np.ones(8), [0, 0, 1, 1, 0, 0, 1, 1]])
    d = D.build_design(8, [1] * 4 + [2] * 4, group=[0, 1] * 4, covar_mod=cov)
    assert d.mod.shape == (8, 3) and d.design.shape == (8, 4)
    # 3-level group: treatment contrasts against the first (sorted) level
    d = D.build_design(6, [1, 1, 1, 2, 2, 2], group=["c", "a", "b", "c", "a", "b"])
    assert d.mod.tolist() == [[1, 0, 1], [1, 0, 0], [1, 1, 0], [1, 0, 1], [1, 0, 0], [1, 1, 0]]


def test_design_errors_match_reference_strings(pkg):
    D = pkg.design
    with pytest.raises(ValueError, match="ComBat-seq doesn't support 1 sample per batch yet"):
        D.build_design(6, [1, 1, 1, 2, 2, 3])
    with pytest.raises(ValueError, match="The covariate is confounded with batch! Remove the covariate and rerun ComBat-Seq"):
        D.build_design(6, [1, 1, 1, 2, 2, 2], group=[0, 0, 0, 1, 1, 1])
    with pytest.raises(ValueError, match="The covariates are confounded! Please remove one or more"):
        D.build_design(6, [1, 1, 1, 2, 2, 2], group=[0, 1, 0, 1, 0, 1], covar_mod=np.array([0, 1, 0, 1, 0, 1.0]))
    with pytest.raises(ValueError, match="At least one covariate is confounded with batch!"):
        D.build_design(6, [1, 1, 1, 2, 2, 2], group=[0, 1, 0, 1, 0, 1], covar_mod=np.array([0, 0, 0, 1, 1, 1.0]))


def test_qr_rank_matches_oracle_dqrdc2(pkg, oracle):
    rng = np.random.default_rng(4)
    for trial in range(30):
        n, p = int(rng.integers(3, 12)), int(rng.integers(1, 6))
        X = rng.integers(0, 2, size=(n, p)).astype(float)
        if trial % 3 == 0:
            X = np.column_stack([X, X[:, 0]])
        assert pkg.design.qr_rank(X) == oracle.qr_rank(X) == np.linalg.matrix_rank(X)


def test_messages_follow_the_reference(pkg, capsys):
    with pytest.raises(Exception):                     # no device here (or a device: then it simply runs)
        pkg.ComBat_seq(np.ones((4, 4)), [1, 1, 2, 2], shrink=True, gene_subset_n=3)
        raise RuntimeError("ran")
    out = capsys.readouterr().out.splitlines()
    assert out[:3] == ["Found 2 batches", "Using null model in ComBat-seq.", "Adjusting for 0 covariate(s) or covariate level(s)"]
    assert "Estimating dispersions" in out and "Fitting the GLM model" in out
    assert "Apply shrinkage - computing posterior estimates for parameters" in out
    assert "Using 3 random genes for Monte Carlo integration" in out
    assert "Apply shrinkage to mean only" in out and out[-1] == "Adjusting the data"


def test_r_shim_compiles_against_a_stub_of_the_r_api():
    """The .Call shim cannot be built against R here (no R in the image); it is at least compiled (-fsyntax-only) against a
    minimal stand-in of the R API declarations it uses, with the real include/combatseq.h, so that field names, types and
    the entry point's arity stay in step with the C ABI."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    shim = os.path.join(root, "combat-seq_b200", "R", "pkg", "src", "r_shim.c")
    r = subprocess.run(["gcc", "-fsyntax-only", "-Wall", "-Werror", "-I" + os.path.join(root, "tests", "r_stub"),
                        "-I" + os.path.join(root, "include"), shim], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    wrapper = open(os.path.join(root, "combat-seq_b200", "R", "pkg", "R", "ComBat_seq.R")).read()
    assert wrapper.count("as.integer(n.gpus)") == 1 and "n_gpus" in open(shim).read()
