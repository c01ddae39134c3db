"""-m gpu: K5 on the device — the model tables derived by device kernels against the host builder
that preceded them (option "host_tables" = 1, kept as the checker: bit-equality of every table), the device-resident
M-step (sgmm_mstep) against MixtureModel::finalize on the same statistics, and the transactional behaviour of
sgmm_set_model (a rejected model leaves the context usable and unchanged)."""
import numpy as np
import pytest

from tests import helpers
from speechrecognition_b200 import synth

pytestmark = pytest.mark.gpu

TABLES = ["meanT", "invT", "mean_tail", "inv_tail", "meanR", "invU", "var_of", "norm", "logw", "mu32", "h32", "mu32T", "c32",
          "sguard", "Bimg", "sguard_tc", "Bimg16", "sguard16", "tile_info", "col_off"]


DTYPES = {"var_of": np.uint32, "tile_info": np.uint32, "col_off": np.uint32, "mu32": np.float32, "h32": np.float32,
          "mu32T": np.float32, "c32": np.float32, "sguard": np.float32, "Bimg": np.float32, "sguard_tc": np.float32,
          "Bimg16": np.float16, "sguard16": np.float32}


def same_table(name, a, b):
    """Bit equality, except that NaN entries (densities that received no frame: 0/0 means) only have to be NaN on both
    sides — x86 and the GPU produce different NaN sign / payload bits in conversions."""
    if a.size != b.size:
        return False
    dt = DTYPES.get(name, np.float64)
    av, bv = a.view(dt), b.view(dt)
    if np.issubdtype(dt, np.floating):
        nan = np.isnan(av)
        if not np.array_equal(nan, np.isnan(bv)):
            return False
        ib = {np.float64: np.uint64, np.float32: np.uint32, np.float16: np.uint16}[dt]
        return bool(np.array_equal(av.view(ib)[~nan], bv.view(ib)[~nan]))
    return bool(np.array_equal(av, bv))


def packed(t, dim):
    mi, vi = t["mean_idx"].astype(np.int64), t["var_idx"].astype(np.int64)
    return (dim, t["mix_offsets"], t["means"].reshape(-1, dim)[mi], t["vars_inv"].reshape(-1, dim)[vi], t["norm"][vi],
            t["mean_wlog"][mi])


def model_cases():
    rng = np.random.default_rng(3)
    lex = synth.Lexicon()
    yield "pooled_33x160", 33, synth.random_model(rng, 33, lex.num_states, 160, "mixture")
    yield "pooled_33x256", 33, synth.random_model(rng, 33, lex.num_states, 256, "mixture")
    yield "global_25x20", 25, synth.random_model(rng, 25, lex.num_states, 20, "global")
    yield "none_25x3", 25, synth.random_model(rng, 25, lex.num_states, 3, "none")
    yield "none_12x70", 12, synth.random_model(rng, 12, lex.num_states, 70, "none")
    yield "pooled_48x64", 48, synth.random_model(rng, 48, 300, 64, "mixture")
    yield "pooled_70x40", 70, synth.random_model(rng, 70, 20, 40, "mixture")       # no f32 tables (dim > 64)
    t = synth.random_model(rng, 33, lex.num_states, 130, "mixture")                 # tiny and huge entries: fp16 flush terms
    t["means"] = t["means"] * np.where(rng.random(t["means"].shape) < 0.3, 1e-7, 1.0)
    yield "pooled_33x130_tiny", 33, t
    O, corp, ali = helpers.trained_oracle(dim=25, n_utts=24, splits=2)               # holes after eliminate, ragged states
    yield "trained_25", 25, O.tables()


@pytest.mark.parametrize("name,dim,tables", list(model_cases()), ids=lambda v: v if isinstance(v, str) else "")
def test_device_tables_equal_host_tables(name, dim, tables):
    from speechrecognition_b200 import capi
    ctx = capi.Context(0)
    args = packed(tables, dim)
    ctx.set_option("host_tables", 1)
    ctx.set_model(*args)
    want = {k: ctx.debug_table(k) for k in TABLES}
    want_form = (ctx.get_counter("tc_form"), ctx.get_counter("tc_ksteps"), ctx.get_counter("scorer"))
    ctx.set_option("host_tables", 0)
    ctx.set_model(*args)
    assert (ctx.get_counter("tc_form"), ctx.get_counter("tc_ksteps"), ctx.get_counter("scorer")) == want_form
    sizes = {}
    for k in TABLES:
        got = ctx.debug_table(k)
        sizes[k] = got.size
        assert got.size == want[k].size, (name, k, got.size, want[k].size)
        assert same_table(k, got, want[k]), (name, k, int(np.flatnonzero(got != want[k])[0]))
    assert sizes["meanR"] > 0 and sizes["invU"] > 0
    ctx.close()


def fold_reference(tables, dim, cnt, sx, sxx, pooling):
    """The per-density statistics folded into the reference's accumulator slots in its order (Mixtures.cpp:333-357 with
    the 1e-4 seed of :243) -> table dict for the host MixtureModel."""
    t = {k: np.array(v, copy=True) for k, v in tables.items()}
    mi, vi = t["mean_idx"].astype(np.int64), t["var_idx"].astype(np.int64)
    macc = np.zeros_like(t["mean_acc"]).reshape(-1, dim)
    vacc = np.full_like(t["var_acc"], 1e-4).reshape(-1, dim)
    mw, vw = np.zeros_like(t["mean_wacc"]), np.zeros_like(t["var_wacc"])
    for j in range(len(mi)):
        mw[mi[j]] += cnt[j]; vw[vi[j]] += cnt[j]
        macc[mi[j]] += sx[j]; vacc[vi[j]] += sxx[j]
    t["mean_acc"], t["var_acc"], t["mean_wacc"], t["var_wacc"] = macc.reshape(-1), vacc.reshape(-1), mw, vw
    return t


@pytest.mark.parametrize("pooling", ["mixture", "none", "global"])
def test_mstep_matches_host_finalize(pooling):
    """sgmm_mstep against MixtureModel::finalize (host C++, the reference's operation order) on the SAME statistics:
    means, variances, inverse variances and weights bit-identical (IEEE divisions / products), norm and log weight
    within 4 ulp (the device's log against glibc's); and the tables sgmm_mstep builds equal the ones sgmm_set_model
    builds from the parameters it produced."""
    from speechrecognition_b200 import capi, hostapi
    rng = np.random.default_rng(17)
    dim, dens = 25, 12
    lex = synth.Lexicon()
    tables = synth.random_model(rng, dim, lex.num_states, dens, pooling)
    corp = synth.make_corpus(rng, lex, tables, dim, 60)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
    ctx.set_tdp(*helpers.TDP, lex.silence_state)
    ctx.align()
    cnt, sx, sxx, _ = ctx.accumulate()
    c2, x2, xx2, _ = ctx.get_statistics()
    assert np.array_equal(cnt, c2) and np.array_equal(sx, x2) and np.array_equal(sxx, xx2)
    ctx.mstep(pooling)
    got = ctx.get_model(want_var=True)
    H = hostapi.HostModel(dim, lex.num_states, pooling, host_only=True)
    H.load_tables(fold_reference(tables, dim, cnt, sx, sxx, pooling))
    H.finalize()
    t = H.tables()
    mi, vi = t["mean_idx"].astype(np.int64), t["var_idx"].astype(np.int64)
    live = cnt > 0
    assert live.sum() > 0.9 * len(cnt)
    eq = lambda a, b: np.array_equal(a, b, equal_nan=True)
    assert eq(got["mean"], t["means"].reshape(-1, dim)[mi])
    assert eq(got["var"], t["vars"].reshape(-1, dim)[vi])
    assert eq(got["inv_var"], t["vars_inv"].reshape(-1, dim)[vi])
    assert eq(got["weight"], t["mean_w"][mi])
    ulp = lambda a, b: np.nanmax(np.abs(a - b) / np.maximum(np.spacing(np.abs(b)), 1e-300))
    d_norm, d_logw = ulp(got["norm"], t["norm"][vi]), ulp(got["log_weight"][live], t["mean_wlog"][mi][live])
    print(f"mstep[{pooling}]: norm differs by <= {d_norm:.1f} ulp, log weight by <= {d_logw:.1f} ulp from glibc's")
    assert d_norm <= 4 and d_logw <= 4
    # the tables behind the new model = what sgmm_set_model derives from these very parameters
    mine = {k: ctx.debug_table(k) for k in TABLES}
    ctx2 = capi.Context(0)
    ctx2.set_option("host_tables", 1)
    ctx2.set_model(dim, tables["mix_offsets"], got["mean"], got["inv_var"], got["norm"], got["log_weight"])
    for k in TABLES:
        if k in ("var_of", "invU") and pooling == "none":
            continue   # (set_model folds equal consecutive rows; the layouts agree unless two densities coincide)
        assert same_table(k, mine[k], ctx2.debug_table(k)), k
    # and the next alignment with the device-built model equals the one with the uploaded copy
    a1, c1 = ctx.align()
    ctx2.set_features(corp["x"]); ctx2.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
    ctx2.set_tdp(*helpers.TDP, lex.silence_state)
    a2, c2_ = ctx2.align()
    assert np.array_equal(a1["state"], a2["state"]) and np.array_equal(c1, c2_)
    ctx.close(); ctx2.close()


def test_set_model_is_transactional():
    """A rejected sgmm_set_model leaves the context holding the previous model, fully usable."""
    from speechrecognition_b200 import capi
    rng = np.random.default_rng(9)
    dim = 25
    lex = synth.Lexicon()
    tables = synth.random_model(rng, dim, lex.num_states, 8, "mixture")
    corp = synth.make_corpus(rng, lex, tables, dim, 10)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
    ctx.set_tdp(*helpers.TDP, lex.silence_state)
    a0, c0 = ctx.align()
    s0, m0 = ctx.score_all(0, 100)
    args = list(packed(tables, dim))
    bad_off = np.array(args[1], copy=True); bad_off[5] = bad_off[6] + 3          # not monotone
    with pytest.raises(capi.SgmmError):
        ctx.set_model(dim, bad_off, *args[2:])
    huge = np.array([0, 70000], np.uint32)                                        # more than 65535 densities in a state
    z = np.zeros((70000, dim))
    with pytest.raises(capi.SgmmError):
        ctx.set_model(dim, huge, z, z + 1.0, np.zeros(70000), np.zeros(70000))
    D = int(tables["mix_offsets"][-1])
    with pytest.raises(capi.SgmmError):
        ctx.set_model(1, tables["mix_offsets"], np.zeros((D, 1)), np.ones((D, 1)), np.zeros(D), np.zeros(D))   # dim out of range
    ctx.dim, ctx.n_states, ctx.D = dim, lex.num_states, int(tables["mix_offsets"][-1])
    a1, c1 = ctx.align()
    s1, m1 = ctx.score_all(0, 100)
    assert np.array_equal(a0["state"], a1["state"]) and np.array_equal(c0, c1)
    assert np.array_equal(s0, s1) and np.array_equal(m0, m1)
    ctx.close()


def test_corpus_change_invalidates_cached_estep():
    """sgmm_align (corpus A) -> sgmm_set_corpus (corpus B, same frame total, other layout) -> sgmm_accumulate must not
    read A's emission matrix with B's offsets (advisor finding): the E-step re-scores, results equal the oracle's."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi
    rng = np.random.default_rng(4)
    dim = 25
    lex = synth.Lexicon()
    tables = synth.random_model(rng, dim, lex.num_states, 4, "none")
    corp = synth.make_corpus(rng, lex, tables, dim, 12)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
    ctx.set_tdp(*helpers.TDP, lex.silence_state)
    ali, _ = ctx.align()
    # corpus B: the same frames cut into other utterances (first two merged) with other automata
    fo = np.concatenate([[0], corp["frame_off"][2:]]).astype(np.uint64)
    auts = [lex.segment_automaton(rng.integers(1, 12, size=2)) for _ in range(len(fo) - 1)]
    ao = np.concatenate([[0], np.cumsum([len(a) for a in auts])]).astype(np.uint32)
    ctx.set_corpus(fo, np.concatenate(auts), ao)
    ctx.set_alignment(ali)
    cnt, sx, sxx, ssum = ctx.accumulate(first_pass=False)
    O = Oracle(dim, lex.num_states, "none", True, helpers.TDP, lex.silence_state)
    O.load_tables(tables)
    want_dens = O.accumulate(ali["state"].astype(np.uint32), corp["x"])
    dens, _ = ctx.last_frame_results()
    assert np.array_equal(dens, want_dens)
    t = O.tables()
    assert np.array_equal(cnt, t["mean_wacc"][t["mean_idx"]])
    ctx.close()


def test_one_frame_utterance_cost():
    """T == 1: align_sequence_full never enters its frame loop and returns +inf (Alignment.cpp:62-63,142)."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi
    rng = np.random.default_rng(2)
    dim = 25
    lex = synth.Lexicon()
    tables = synth.random_model(rng, dim, lex.num_states, 2, "none")
    x = rng.standard_normal((41, dim)).astype(np.float32)
    fo = np.array([0, 1, 41], np.uint64)
    auts = [np.array([0], np.uint32), lex.segment_automaton([3])]
    ao = np.array([0, 1, 1 + len(auts[1])], np.uint32)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(x)
    ctx.set_corpus(fo, np.concatenate(auts), ao)
    ctx.set_tdp(*helpers.TDP, lex.silence_state)
    ali, costs = ctx.align()
    O = Oracle(dim, lex.num_states, "none", True, helpers.TDP, lex.silence_state)
    O.load_tables(tables)
    c0, st0, _ = O.align_full(x[:1], auts[0])
    c1, st1, _ = O.align_full(x[1:], auts[1])
    assert costs[0] == c0 == np.inf and costs[1] == c1
    assert np.array_equal(ali["state"][:1].astype(np.uint32), st0) and np.array_equal(ali["state"][1:].astype(np.uint32), st1)
    ctx.close()
