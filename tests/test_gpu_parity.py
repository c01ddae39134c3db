"""-m gpu: the CUDA path (through the C-ABI) against the oracle on the same seeded inputs.

Bars (BASELINE.json north_star): alignments, density argmins, counts bit-exact; scores, sum_x, sum_xx within
1e-5 relative.  This implementation evaluates the reference's f64 operation order on the device, so the
tests below demand MORE: scores and path costs bit-identical, statistics to 1e-12 relative.
"""
import numpy as np
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu

STAT_RTOL = 1e-12   # north_star allows 1e-5


@pytest.fixture(scope="module")
def small25():
    O, corp, ali = helpers.trained_oracle(dim=25, n_utts=24, splits=2)
    ctx = helpers.gpu_context(O, corp, 25)
    yield O, corp, ali, ctx
    ctx.close()


@pytest.fixture(scope="module")
def small33():
    O, corp, ali = helpers.trained_oracle(dim=33, n_utts=16, splits=3, seed=11, min_obs=10, dens_gen=4)
    ctx = helpers.gpu_context(O, corp, 33)
    yield O, corp, ali, ctx
    ctx.close()


SCORERS = {"f64": 1, "ffma": 2, "tc": 3}


@pytest.mark.parametrize("scorer", ["f64", "ffma", "tc"])
@pytest.mark.parametrize("fx", ["small25", "small33"])
def test_score_all_bit_exact(fx, scorer, request):
    O, corp, ali, ctx = request.getfixturevalue(fx)
    ctx.set_option("scorer", SCORERS[scorer])
    n = 700
    want_s, want_a = O.min_score(corp["x"][:n])
    got_s, got_a = ctx.score_all(0, n)
    assert np.array_equal(got_a, want_a)
    assert np.array_equal(got_s, want_s)          # bit-identical f64
    # ragged range that is not a multiple of the 32-frame warp tile
    got_s2, got_a2 = ctx.score_all(37, 37 + 45)
    assert np.array_equal(got_s2, want_s[37:82]) and np.array_equal(got_a2, want_a[37:82])


@pytest.mark.parametrize("scorer", ["f64", "ffma", "tc"])
@pytest.mark.parametrize("fx", ["small25", "small33"])
def test_align_full_bit_exact(fx, scorer, request):
    O, corp, ali, ctx = request.getfixturevalue(fx)
    ctx.set_option("scorer", SCORERS[scorer])
    want_states, want_costs = O.align_batch(corp["x"], corp["frame_off"], corp["automata"], corp["aut_off"], threads=8)
    got, costs = ctx.align()
    assert np.array_equal(got["state"].astype(np.uint32), want_states)
    assert np.all(got["count"] == 1) and np.all(got["weight"] == 1.0)   # Alignment.cpp:131-133
    assert np.array_equal(costs, want_costs)      # f64 path costs, bit-identical


def test_align_destination_convention(small25):
    O, corp, ali, ctx = small25
    from speechrecognition_b200 import capi
    got, costs = ctx.align(capi.TDP_DESTINATION)
    for u in range(len(corp["frame_off"]) - 1):
        f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
        aut = corp["automata"][corp["aut_off"][u]:corp["aut_off"][u + 1]]
        c, st, _ = O.align_pruned(corp["x"][f0:f1], aut, 1e300)
        assert c == costs[u]
        assert np.array_equal(got["state"][f0:f1].astype(np.uint32), st)
    ctx.align()  # leave the source-convention alignment resident for the following tests


@pytest.mark.parametrize("threshold", [0.0, 5.0, 20.0, 40.0, 150.0])
@pytest.mark.parametrize("fx", ["small25", "small33"])
def test_align_pruned_beam(fx, threshold, request):
    """Aligner::align_sequence_pruned with a finite beam (Alignment.cpp:227-239): paths, costs and — when the beam
    kills the path to the last position — the "highest position alive" end, all as the reference has them."""
    O, corp, ali, ctx = request.getfixturevalue(fx)
    from speechrecognition_b200 import capi
    ctx.set_beam(threshold)
    got, costs = ctx.align(capi.TDP_DESTINATION)
    ctx.set_beam()
    short = 0
    for u in range(len(corp["frame_off"]) - 1):
        f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
        aut = corp["automata"][corp["aut_off"][u]:corp["aut_off"][u + 1]]
        c, st, pos = O.align_pruned(corp["x"][f0:f1], aut, threshold)
        assert c == costs[u], u
        assert np.array_equal(got["state"][f0:f1].astype(np.uint32), st), u
        short += int(pos[-1] != len(aut) - 1)
    if threshold <= 5.0:
        assert short > 0                   # the tight beams did strand some paths before the last position
    # the beam is a property of the destination convention only: the full aligner ignores it
    ctx.set_beam(threshold)
    again, costs2 = ctx.align()
    ctx.set_beam()
    ref, costs_ref = ctx.align()
    assert np.array_equal(again, ref) and np.array_equal(costs2, costs_ref)


def test_alignment_golden():
    """Committed outputs of the unmodified reference (tests/golden/trajectory_none.npz, tools/make_golden.py): scores,
    full and pruned alignments with their costs, and the accumulators of the following E-step."""
    import os
    from oracle.binding import TABLES
    from speechrecognition_b200 import capi, synth
    G = os.path.join(os.path.dirname(__file__), "golden")
    c = np.load(os.path.join(G, "corpus.npz"))
    rec = np.load(os.path.join(G, "trajectory_none.npz"))
    dim = int(c["dim"])
    lex = synth.Lexicon()
    t1 = {k: rec["t1_" + k] for k in TABLES}
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, t1)
    ctx.set_features(c["x"])
    ctx.set_corpus(c["frame_off"], c["automata"], c["aut_off"])
    ctx.set_tdp(*[float(v) for v in c["tdp"]], lex.silence_state)
    sc, am = ctx.score_all(0, 64)
    assert np.array_equal(sc, rec["scores64"]) and np.array_equal(am, rec["argmin64"])
    f0, f1 = int(c["frame_off"][2]), int(c["frame_off"][3])
    for thr, tag in ((40.0, "p40"), (1e300, "pinf")):
        ctx.set_beam(thr)
        got, costs = ctx.align(capi.TDP_DESTINATION)
        assert costs[2] == float(rec["cost_" + tag])
        assert np.array_equal(got["state"][f0:f1].astype(np.uint32), rec["ali_" + tag])
    ctx.set_beam()
    got, costs = ctx.align()
    assert np.array_equal(got["state"].astype(np.uint32), rec["ali_full"])
    assert np.array_equal(costs, rec["costs_full"])
    cnt, sx, sxx, _ = ctx.accumulate(first_pass=False)
    t2 = {k: rec["t2_" + k] for k in TABLES}
    want_cnt, want_sx = helpers.packed_stats(t2, dim)
    assert np.array_equal(cnt, want_cnt)
    np.testing.assert_allclose(sx, want_sx, rtol=STAT_RTOL, atol=1e-9)
    want_sxx = t2["var_acc"].reshape(-1, dim)[t2["var_idx"].astype(np.int64)] - 1e-4
    np.testing.assert_allclose(sxx, want_sxx, rtol=1e-9, atol=1e-7)
    ctx.close()


@pytest.mark.parametrize("fx", ["small25", "small33"])
def test_accumulate_matches_oracle(fx, request):
    O, corp, ali, ctx = request.getfixturevalue(fx)
    dim = O.dim
    ctx.set_alignment(ali)
    cnt, sx, sxx, ssum = ctx.accumulate(first_pass=False)
    dens, fscore = ctx.last_frame_results()
    # oracle: same E-step on a copy of the model
    from oracle.binding import Oracle
    P = Oracle(dim, O.n_mix, "none", True, O.tdp, O.silence)
    P.load_tables(O.tables())
    want_dens = P.accumulate(ali, corp["x"])
    t = P.tables()
    want_cnt, want_sx = helpers.packed_stats(t, dim)
    assert np.array_equal(dens, want_dens)                       # density argmin per frame: bit-exact
    assert np.array_equal(cnt, want_cnt)                         # counts: exact
    # frames aligned to a state whose densities were all eliminated are dropped, as in the reference
    assert cnt.sum() == np.sum(np.diff(t["mix_offsets"])[ali] > 0)
    np.testing.assert_allclose(sx, want_sx, rtol=STAT_RTOL, atol=1e-9)
    vi = t["var_idx"].astype(np.int64)
    want_sxx = t["var_acc"].reshape(-1, dim)[vi] - 1e-4          # reference seeds 1e-4 (Mixtures.cpp:243)
    np.testing.assert_allclose(sxx, want_sxx, rtol=1e-9, atol=1e-7)
    want_scores, _ = O.min_score(corp["x"][:300])
    assert np.array_equal(fscore[:300], want_scores[np.arange(300), ali[:300]])
    am = O.am_score(ali, corp["x"])
    assert abs(ssum / len(ali) - am) <= 1e-12 * abs(am)
    assert abs(ctx.am_score_sum() - ssum) == 0.0                 # deterministic reduction


def test_accumulate_first_pass_and_determinism(small25):
    O, corp, ali, ctx = small25
    ctx.set_alignment(ali)
    c1 = ctx.accumulate(first_pass=True)
    c2 = ctx.accumulate(first_pass=True)
    for a, b in zip(c1[:3], c2[:3]):
        assert np.array_equal(a, b)                              # bit-reproducible
    off = O.tables()["mix_offsets"]
    live = np.diff(off) > 0                                      # states that still have a density
    want = np.bincount(ali, minlength=O.n_mix).astype(np.float64)
    assert np.array_equal(c1[0][off[:-1][live]], want[live])     # everything lands in density 0 of its state
    assert c1[0].sum() == want[live].sum()
    a3 = ctx.accumulate(first_pass=False)
    a4 = ctx.accumulate(first_pass=False)
    for a, b in zip(a3[:3], a4[:3]):
        assert np.array_equal(a, b)


def test_edge_shapes():
    """Tiny and ragged utterances: S < 32 (two positions per lane), T not a multiple of the 16-frame tile,
    T == S (only the diagonal path), long automaton (S > 256), single-frame tail tiles."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(3)
    dim, n_states = 6, 40
    t = synth.random_model(rng, dim, n_states, 3, "none")
    O = Oracle(dim, n_states, "none", True, (2.0, 0.5, 7.0), 0)
    O.load_tables(t)
    shapes = [(5, 3), (17, 17), (33, 9), (16, 2), (100, 70), (640, 300), (31, 1), (2, 2), (48, 33)]
    xs, auts, f_off, a_off = [], [], [0], [0]
    for T, S in shapes:
        aut = rng.integers(0, n_states, size=S).astype(np.uint32)
        aut[rng.integers(0, S)] = 0   # make sure the silence state shows up
        st = aut[synth.sample_path(rng, T, S)]
        xs.append(synth.emit_frames(rng, t, dim, st)); auts.append(aut)
        f_off.append(f_off[-1] + T); a_off.append(a_off[-1] + S)
    x = np.concatenate(xs); automata = np.concatenate(auts)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, t); ctx.set_features(x); ctx.set_corpus(f_off, automata, a_off)
    ctx.set_tdp(2.0, 0.5, 7.0, 0)
    got, costs = ctx.align()
    want_states, want_costs = O.align_batch(x, f_off, automata, a_off)
    assert np.array_equal(costs, want_costs)
    assert np.array_equal(got["state"].astype(np.uint32), want_states)
    ctx.close()


def test_errors_are_loud():
    from speechrecognition_b200 import capi
    ctx = capi.Context(0)
    with pytest.raises(capi.SgmmError):
        ctx.align()
    with pytest.raises(capi.SgmmError):
        capi.Context(99)
    ctx.close()


def test_ffma_scorer_ties_and_near_ties():
    """The f32 filter must never decide a near-tie: duplicated densities (exact f64 ties -> the reference keeps
    the FIRST, Mixtures.cpp:705), densities 1e-9 apart, and far-apart ones, all in one model."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(17)
    dim, n_states, nd = 33, 12, 16
    t = synth.random_model(rng, dim, n_states, nd, "none")
    means = t["means"].reshape(-1, dim); vinv = t["vars_inv"].reshape(-1, dim)
    vars_ = t["vars"].reshape(-1, dim)
    for s in range(n_states):
        b = s * nd
        means[b + 5] = means[b + 2]; vinv[b + 5] = vinv[b + 2]; vars_[b + 5] = vars_[b + 2]      # exact duplicate
        t["norm"][b + 5] = t["norm"][b + 2]; t["mean_wlog"][b + 5] = t["mean_wlog"][b + 2]
        means[b + 9] = means[b + 3] * (1 + 1e-9); vinv[b + 9] = vinv[b + 3]; vars_[b + 9] = vars_[b + 3]
        t["norm"][b + 9] = t["norm"][b + 3]; t["mean_wlog"][b + 9] = t["mean_wlog"][b + 3]
    O = Oracle(dim, n_states, "none", True, (3.0, 0.0, 30.0), 0)
    O.load_tables(t)
    # frames drawn AT the duplicated / near-duplicated densities so that they are the winners
    st = np.repeat(np.arange(n_states), 40)
    x = synth.emit_frames(rng, t, dim, st)
    x[::4] = means[st[::4] * nd + 2] + 0.01 * rng.standard_normal((len(st[::4]), dim))
    x[1::4] = means[st[1::4] * nd + 3] + 0.01 * rng.standard_normal((len(st[1::4]), dim))
    x = x.astype(np.float32)
    want_s, want_a = O.min_score(x)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, t); ctx.set_features(x)
    ctx.set_option("scorer", 2)
    assert ctx.get_counter("scorer") == 2
    ctx.get_counter("fallbacks")
    got_s, got_a = ctx.score_all()
    assert np.array_equal(got_a, want_a) and np.array_equal(got_s, want_s)
    fb = ctx.get_counter("fallbacks")
    assert fb >= len(st) // 2                      # the planted ties did take the f64 rescan path
    assert fb < 0.2 * got_s.size                   # ... and only they did
    ctx.close()


@pytest.mark.parametrize("halves", [1, 2])
@pytest.mark.parametrize("dim,nd,pooling", [(33, 200, "mixture"), (25, 40, "none"), (33, 256, "global"), (12, 17, "none"),
                                            (48, 64, "mixture"), (48, 30, "none"), (25, 70, "mixture"), (32, 150, "global")])
def test_tensor_core_scorer_exact(dim, nd, pooling, halves):
    """tcgen05 3xTF32 GEMM + certified refinement: dense scores/argmins and batched alignments bit-identical to
    the oracle, in the pooled (K = d+1) and the full (K = 2d+1) operand forms, with ragged state sizes."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(23)
    lex = synth.Lexicon()
    t = synth.random_model(rng, dim, lex.num_states, nd, pooling)
    # make the state sizes ragged: drop a few densities of every third state
    if nd > 20:
        keep = np.ones(len(t["mean_idx"]), bool)
        for s in range(0, lex.num_states, 3):
            keep[s * nd + nd - 7: s * nd + nd] = False
        sizes = np.array([keep[s * nd:(s + 1) * nd].sum() for s in range(lex.num_states)])
        t["mean_idx"], t["var_idx"] = t["mean_idx"][keep], t["var_idx"][keep]
        t["mix_offsets"] = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint32)
    corp = synth.make_corpus(rng, lex, t, dim, 5)
    tdp = (3.0, 0.0, 30.0)
    O = Oracle(dim, lex.num_states, pooling, True, tdp, lex.silence_state)
    O.load_tables(t)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, t); ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(*tdp, lex.silence_state)
    ctx.set_option("scorer", 3)
    ctx.set_option("tc_halves", halves)           # 128 or 256 frames per GEMM CTA
    assert ctx.get_counter("scorer") == 3
    n = 300
    want_s, want_a = O.min_score(corp["x"][:n])
    ctx.get_counter("fallbacks")
    got_s, got_a = ctx.score_all(0, n)
    fb = ctx.get_counter("fallbacks")
    assert np.array_equal(got_a, want_a)
    assert np.array_equal(got_s, want_s)
    assert fb < 0.25 * got_s.size, f"{fb} of {got_s.size} cells fell back to the f64 rescan"
    ctx.set_option("tc_rescan", 1)                 # plain full-f64 rescan instead of the two-stage one: same bits
    got_s1, got_a1 = ctx.score_all(0, n)
    assert np.array_equal(got_a1, want_a) and np.array_equal(got_s1, want_s)
    assert ctx.get_counter("fallbacks") == fb
    ctx.set_option("tc_rescan", 0)
    want_states, want_costs = O.align_batch(corp["x"], corp["frame_off"], corp["automata"], corp["aut_off"], threads=8)
    got, costs = ctx.align()
    assert np.array_equal(costs, want_costs)
    assert np.array_equal(got["state"].astype(np.uint32), want_states)
    # both refinement kernels (frame-major; state-major with the state's table in shared memory): same bits
    for mode in (1, 2, 0):
        ctx.set_option("tc_refine", mode)
        again, costs2 = ctx.align()
        assert np.array_equal(again, got) and np.array_equal(costs2, costs), mode
    ctx.close()


@pytest.mark.parametrize("scale", [1.0, 3e-4, 700.0])
def test_fp16_operand_form_matches_tf32_and_oracle(scale):
    """The wide GEMM's fp16 hi/lo operand form (option tc_f16; rows and states scaled by powers of two) against the
    tf32 form and the oracle: dense scores / argmins and alignments bit-identical, for features far below and far
    above unit scale, a frame with a huge feature (rescanned in f64) and planted near-duplicate densities."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(31)
    dim, nd = 33, 200
    lex = synth.Lexicon()
    t = synth.random_model(rng, dim, lex.num_states, nd, "mixture")
    t["means"] *= scale                             # the model lives at the features' scale
    t["vars"] *= scale ** 2
    t["vars_inv"] /= scale ** 2
    t["norm"] += dim * np.log(scale)
    means = t["means"].reshape(-1, dim)
    for k, rel in ((1, 0.0), (3, 1e-9), (5, 1e-6), (7, 1e-4)):     # near-duplicates of density 0 of state 0
        means[k] = means[0] * (1.0 + rel)
    corp = synth.make_corpus(rng, lex, t, dim, 6)
    x = corp["x"].copy()
    x[17, 4] = 3.0e6 * scale                        # beyond the fp16 row scaling: the frame becomes NaN -> rescanned
    tdp = (3.0, 0.0, 30.0)
    O = Oracle(dim, lex.num_states, "mixture", True, tdp, lex.silence_state)
    O.load_tables(t)
    n = 400
    want_s, want_a = O.min_score(x[:n])
    want_states, want_costs = O.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"], threads=8)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, t); ctx.set_features(x)
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(*tdp, lex.silence_state)
    ctx.set_option("scorer", 3)
    fbs = []
    # fourth: sixteen epilogue warps per unit (option tc_epi16; certified-only verdicts); fifth: the default epilogue
    # without pair verdicts (option tc_pairs = 0)
    for f16, epi16, pairs in ((1, 0, 1), (0, 0, 1), (1, 0, 1), (1, 1, 1), (1, 0, 0)):
        ctx.set_option("tc_f16", f16)
        ctx.set_option("tc_epi16", epi16)
        ctx.set_option("tc_pairs", pairs)
        ctx.get_counter("fallbacks")
        got_s, got_a = ctx.score_all(0, n)
        fbs.append(ctx.get_counter("fallbacks"))
        assert np.array_equal(got_a, want_a), f16
        assert np.array_equal(got_s, want_s), f16
        got, costs = ctx.align()
        assert np.array_equal(got["state"].astype(np.uint32), want_states), f16
        assert np.array_equal(costs, want_costs), f16
    assert fbs[0] == fbs[2] and fbs[0] < 0.1 * want_s.size, fbs      # the fp16 form certifies nearly everything, too
    assert fbs[3] <= 1.2 * fbs[4] + 16, fbs                          # finer column parts: at most a few more uncertain cells
    assert fbs[0] <= fbs[4], fbs                                     # pair verdicts only ever take cells OUT of the rescan queue
    ctx.close()


@pytest.mark.parametrize("dim", [33, 32, 25])
def test_state_major_refinement_short_utterances(dim):
    """Utterances of 3..40 frames: a warp's 32 cells of one state then span several utterances (bulk-copy staging
    takes at most two, the rest goes through the element-wise path); near-ties planted so that the rescan queue is used."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(5)
    n_states, nd = 12, 64
    t = synth.random_model(rng, dim, n_states, nd, "mixture")
    means = t["means"].reshape(-1, dim)
    means[1] = means[0]                                    # exact duplicate density: a tie the filter cannot certify
    auts, xs, f_off, a_off = [], [], [0], [0]
    for u in range(150):
        S = int(rng.integers(1, 4))
        aut = rng.integers(0, n_states, size=S).astype(np.uint32)
        T = int(rng.integers(max(3, S), 41))
        st = aut[synth.sample_path(rng, T, S)]
        xs.append(synth.emit_frames(rng, t, dim, st)); auts.append(aut)
        f_off.append(f_off[-1] + T); a_off.append(a_off[-1] + S)
    x = np.concatenate(xs).astype(np.float32)
    frame_off, aut_off, automata = np.array(f_off, np.uint64), np.array(a_off, np.uint32), np.concatenate(auts)
    O = Oracle(dim, n_states, "mixture", True, (3.0, 0.0, 30.0), 0)
    O.load_tables(t)
    want_states, want_costs = O.align_batch(x, frame_off, automata, aut_off, threads=8)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, t); ctx.set_features(x)
    ctx.set_corpus(frame_off, automata, aut_off); ctx.set_tdp(3.0, 0.0, 30.0, 0)
    ctx.set_option("scorer", 3)
    for mode, rescan in ((2, 0), (1, 0), (2, 1)):
        ctx.set_option("tc_refine", mode)
        ctx.set_option("tc_rescan", rescan)
        got, costs = ctx.align()
        assert np.array_equal(got["state"].astype(np.uint32), want_states), mode
        assert np.array_equal(costs, want_costs), mode
        cnt, sx, sxx, ssum = ctx.accumulate(first_pass=False)
        dens, _ = ctx.last_frame_results()
        P = Oracle(dim, n_states, "mixture", True, (3.0, 0.0, 30.0), 0)
        P.load_tables(t)
        assert np.array_equal(dens, P.accumulate(want_states, x)), mode     # cached argmins of the aligned cells
    ctx.close()


@pytest.mark.parametrize("scorer", ["f64", "ffma", "tc"])
def test_accumulate_after_align_uses_cached_scores(small25, scorer):
    """align -> accumulate with an unchanged model takes the lookup path (the alignment pass already scored the
    aligned cells); it must agree bit for bit with the recomputing path and with the oracle."""
    O, corp, ali, ctx = small25
    ctx.set_option("scorer", SCORERS[scorer])
    got, _ = ctx.align()
    states = got["state"].astype(np.uint32)
    ctx.set_option("rescore_cache", 1)
    a = ctx.accumulate(first_pass=False)
    da, sa = ctx.last_frame_results()
    ctx.set_option("rescore_cache", 0)
    b = ctx.accumulate(first_pass=False)
    db, sb = ctx.last_frame_results()
    ctx.set_option("rescore_cache", 1)
    assert np.array_equal(da, db) and np.array_equal(sa, sb)
    for u, v in zip(a[:3], b[:3]):
        assert np.array_equal(u, v)
    assert a[3] == b[3]
    from oracle.binding import Oracle
    P = Oracle(O.dim, O.n_mix, "none", True, O.tdp, O.silence)
    P.load_tables(O.tables())
    want_dens = P.accumulate(states, corp["x"])
    assert np.array_equal(da, want_dens)
    # a new model invalidates the cache: the recomputing path must be taken (and still be right)
    ctx.set_model_tables(O.dim, O.tables())
    c = ctx.accumulate(first_pass=False)
    for u, v in zip(a[:3], c[:3]):
        assert np.array_equal(u, v)


@pytest.mark.parametrize("scorer", ["f64", "ffma", "tc"])
def test_async_upload_same_results(scorer):
    """Option async_upload: sgmm_set_features queues chunked copies from pinned memory and returns; the GEMM scorer
    consumes chunk by chunk, every other consumer waits for the last chunk.  Alignment, costs and statistics
    must be bit-identical to the synchronous upload."""
    from speechrecognition_b200 import capi
    O, corp, ali = helpers.trained_oracle(dim=33, n_utts=40, splits=3, seed=5, min_obs=10, dens_gen=4)
    ctx = helpers.gpu_context(O, corp, 33)
    ctx.set_option("scorer", SCORERS[scorer])
    want_ali, want_costs = ctx.align()
    want = ctx.accumulate(first_pass=False)
    x_pin = capi.pinned_empty(corp["x"].shape, np.float32)
    x_pin[...] = corp["x"]
    ctx.set_option("async_upload", 1)
    for _ in range(2):                       # second round: the chunks overwrite a buffer that is already resident
        ctx.set_features(x_pin)
        got_ali, got_costs = ctx.align()
        got = ctx.accumulate(first_pass=False)
        assert np.array_equal(got_ali, want_ali) and np.array_equal(got_costs, want_costs)
        for u, v in zip(got[:3], want[:3]):
            assert np.array_equal(u, v)
        assert got[3] == want[3]
    # upload followed directly by a consumer that is not the scorer
    ctx.set_features(x_pin)
    ctx.set_alignment(want_ali)
    again = ctx.accumulate(first_pass=False)
    for u, v in zip(again[:3], want[:3]):
        assert np.array_equal(u, v)
    ctx.set_features(x_pin)
    s_async, a_async = ctx.score_all(0, 64)
    ctx.set_option("async_upload", 0)
    ctx.set_features(x_pin)
    s_sync, a_sync = ctx.score_all(0, 64)
    assert np.array_equal(s_async, s_sync) and np.array_equal(a_async, a_sync)
    ctx.close()


@pytest.mark.parametrize("scorer", ["f64", "tc"])
@pytest.mark.parametrize("scale,word_penalty,am_threshold", [(1.0, 10.0, 20.0), (1.0, 3.0, 1e30), (0.05, 10.0, 60.0)])
def test_recognize_matches_oracle(scale, word_penalty, am_threshold, scorer):
    """BASELINE config 4: batched word-loop recognition (Recognizer.cpp:103-232) — recognised word sequences and the
    final word-end score identical to the oracle for the default beam, no beam, and negative emission scores."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    O0, corp, ali = helpers.trained_oracle(dim=25, n_utts=24, splits=2)
    x = (corp["x"] * scale).astype(np.float32)
    lex = synth.Lexicon()
    O = O0
    if scale != 1.0:                       # retrain on the scaled features: tiny variances, negative scores
        O = Oracle(25, lex.num_states, "none", True, O0.tdp, O0.silence)
        O.accumulate(ali, x, first_pass=True); O.finalize()
    ctx = capi.Context(0)
    ctx.set_model_tables(25, O.tables())
    ctx.set_features(x)
    ctx.set_tdp(*O.tdp, O.silence)
    ctx.set_option("scorer", SCORERS[scorer])
    ctx.set_lexicon(lex.automata, lex.silence)
    got, scores = ctx.recognize(corp["frame_off"], word_penalty, am_threshold)
    word_off = np.concatenate([[0], np.cumsum([len(a) for a in lex.automata])]).astype(np.uint32)
    word_states = np.concatenate(lex.automata).astype(np.uint32)
    n_words = 0
    for u in range(len(corp["frame_off"]) - 1):
        f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
        want, want_score = O.recognize(x[f0:f1], word_off, word_states, lex.silence, word_penalty, am_threshold)
        assert np.array_equal(got[u], want), (u, got[u], want)
        assert scores[u] == want_score, (u, scores[u], want_score)
        n_words += len(want)
    assert n_words > 0
    if scale < 1.0:
        s, _ = ctx.score_all(0, 64)
        assert (s < 0).any()               # the negative-score branch was exercised
    ctx.close()


@pytest.mark.parametrize("case", ["k2", "k5_sil_last", "k8_words32", "general_300"])
@pytest.mark.parametrize("negative", [False, True])
def test_recognize_register_kernel_random_lexicons(case, negative):
    """k_recognize_reg (lexicons of <= 256 table slots and <= 32 words, every slots-per-lane count) against the general
    kernel (option rec_reg = 0) and the oracle: random word lengths, one-state words, silence anywhere in the table,
    emission scores of both signs (the per-slot exact walk next to the collapsed word entries)."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng({"k2": 1, "k5_sil_last": 2, "k8_words32": 3, "general_300": 4}[case] + 10 * negative)
    if case == "k2":
        words, sil = [(1, 1)] + [(int(rng.integers(1, 5)), 2) for _ in range(6)], 0
    elif case == "k5_sil_last":
        words = [(int(rng.integers(2, 9)), int(rng.integers(1, 4))) for _ in range(9)] + [(1, 1)]
        sil = len(words) - 1
    elif case == "k8_words32":
        words, sil = [(1, 1)] + [(int(rng.integers(1, 5)), 2) for _ in range(31)], 0
    else:
        words, sil = [(1, 1)] + [(int(rng.integers(8, 14)), 2) for _ in range(14)], 0
    lex = synth.Lexicon(words, sil)
    slots = sum(max(n * r, 2) for n, r in words)
    assert (slots > 256) == (case == "general_300")
    dim = 12
    tables = synth.random_model(rng, dim, lex.num_states, 2, "none")
    if negative:                           # tiny variances: the emission scores along the true path are below zero
        tables["vars"] = tables["vars"] * 0.004
        tables["vars_inv"] = 1.0 / tables["vars"]
        tables["norm"] = (dim * np.log(2 * np.pi) + np.log(tables["vars"].reshape(-1, dim)).sum(1)) / 2
    corp = synth.make_corpus(rng, lex, tables, dim, 40, t_choices=([60, 150, 400], [0.6, 0.3, 0.1]), words_per_utt=(1, 3))
    O = Oracle(dim, lex.num_states, "none", True, (3.0, 0.0, 30.0), lex.silence_state)
    O.load_tables(tables)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(corp["x"])
    ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
    ctx.set_lexicon(lex.automata, lex.silence)
    s, _ = ctx.score_all(0, 64)
    assert (s < 0).any() or not negative
    word_off = np.concatenate([[0], np.cumsum([len(a) for a in lex.automata])]).astype(np.uint32)
    word_states = np.concatenate(lex.automata).astype(np.uint32)
    for wp, thr in ((10.0, 25.0), (0.0, 1e30), (4.0, 3.0)):
        ctx.set_option("rec_reg", 1)
        got, scores = ctx.recognize(corp["frame_off"], wp, thr)
        ctx.set_option("rec_reg", 0)
        got0, scores0 = ctx.recognize(corp["frame_off"], wp, thr)
        assert np.array_equal(scores, scores0)
        for u in range(40):
            assert np.array_equal(got[u], got0[u]), (u, wp, thr)
        for u in range(0, 40, 4):
            f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
            want, want_score = O.recognize(corp["x"][f0:f1], word_off, word_states, lex.silence, wp, thr)
            assert np.array_equal(got[u], want), (u, wp, thr, got[u], want)
            assert scores[u] == want_score or (np.isinf(scores[u]) and np.isinf(want_score))
    ctx.close()


@pytest.mark.parametrize("words,sil", [([(1, 1)], 0), ([(1, 1), (1, 1)], 1), ([(2, 1), (1, 2), (3, 3)], 0xffffffff),
                                       ([(1, 1), (5, 2), (4, 2)], 0)])
def test_recognize_edge_cases(words, sil):
    """Utterances of 1 .. 5 frames (shorter than any word), a lexicon of one word, one without silence, a zero and a
    tiny beam: both search kernels against the oracle (hypotheses and final scores, +inf included)."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(len(words) * 7 + 1)
    lex = synth.Lexicon(words, 0 if sil == 0xffffffff else sil)
    dim = 8
    tables = synth.random_model(rng, dim, lex.num_states, 2, "none")
    lens = [1, 2, 3, 4, 5, 17, 40, 1, 33, 2]
    x = rng.normal(0.0, 1.5, size=(sum(lens), dim)).astype(np.float32)
    frame_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    sil_state = lex.silence_state if sil != 0xffffffff else 0
    O = Oracle(dim, lex.num_states, "none", True, (3.0, 0.0, 30.0), sil_state)
    O.load_tables(tables)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(x)
    ctx.set_tdp(3.0, 0.0, 30.0, sil_state)
    ctx.set_lexicon(lex.automata, sil)
    word_off = np.concatenate([[0], np.cumsum([len(a) for a in lex.automata])]).astype(np.uint32)
    word_states = np.concatenate(lex.automata).astype(np.uint32)
    for wp, thr in ((10.0, 20.0), (0.0, 0.0), (2.5, 1e-3), (1.0, 1e30)):
        res = {}
        for reg in (1, 0):
            ctx.set_option("rec_reg", reg)
            res[reg] = ctx.recognize(frame_off, wp, thr)
        assert np.array_equal(res[1][1], res[0][1]), (wp, thr, res[1][1], res[0][1])
        for u in range(len(lens)):
            assert np.array_equal(res[1][0][u], res[0][0][u]), (u, wp, thr)
            f0, f1 = int(frame_off[u]), int(frame_off[u + 1])
            want, want_score = O.recognize(x[f0:f1], word_off, word_states, sil, wp, thr)
            assert np.array_equal(res[1][0][u], want), (u, wp, thr, res[1][0][u], want)
            assert res[1][1][u] == want_score, (u, wp, thr, res[1][1][u], want_score)
    ctx.close()


def test_recognize_in_several_runs():
    """sgmm_recognize cuts the batch into runs of whole utterances (dense scores of one run at a time, the longest
    utterance of a run launched first): the same hypotheses and scores whatever the run length."""
    from speechrecognition_b200 import capi, synth
    O, corp, _ = helpers.trained_oracle(dim=25, n_utts=24, splits=1)
    lex = synth.Lexicon()
    ctx = capi.Context(0)
    ctx.set_model_tables(25, O.tables())
    ctx.set_features(corp["x"])
    ctx.set_tdp(*O.tdp, O.silence)
    ctx.set_lexicon(lex.automata, lex.silence)
    want, want_scores = ctx.recognize(corp["frame_off"], 10.0, 20.0)
    assert sum(len(w) for w in want) > 0
    for run in (1, 700, 2500):
        for reg in (1, 0):
            ctx.set_option("rec_run_frames", run)
            ctx.set_option("rec_reg", reg)
            got, scores = ctx.recognize(corp["frame_off"], 10.0, 20.0)
            assert np.array_equal(scores, want_scores), (run, reg)
            assert len(got) == len(want) and all(np.array_equal(a, b) for a, b in zip(got, want)), (run, reg)
    # right behind a chunked upload the dense filter is launched per chunk (whole 256-frame tasks below the chunk's end)
    ctx.set_option("rec_reg", 1)
    ctx.set_option("async_upload", 1)
    for run, chunks in ((0, 16), (2500, 5), (0, 3)):
        ctx.set_option("rec_run_frames", run)
        ctx.set_option("upload_chunks", chunks)
        ctx.set_features(corp["x"])
        got, scores = ctx.recognize(corp["frame_off"], 10.0, 20.0)
        assert np.array_equal(scores, want_scores), (run, chunks)
        assert all(np.array_equal(a, b) for a, b in zip(got, want)), (run, chunks)
    ctx.close()


def test_recognize_golden():
    """Committed outputs of the unmodified reference Recognizer (tests/golden/recognize.npz)."""
    import os
    from oracle.binding import Oracle, TABLES
    from speechrecognition_b200 import capi, synth
    G = os.path.join(os.path.dirname(__file__), "golden")
    c = np.load(os.path.join(G, "corpus.npz"))
    rec = np.load(os.path.join(G, "trajectory_none.npz"))
    g = np.load(os.path.join(G, "recognize.npz"))
    dim = int(c["dim"])
    lex = synth.Lexicon()
    O = Oracle(dim, lex.num_states, "none", True, tuple(c["tdp"]), lex.silence_state)
    O.load_tables({k: rec["t2_" + k] for k in TABLES})
    O.finalize()
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, O.tables())
    ctx.set_features(c["x"])
    ctx.set_tdp(*[float(v) for v in c["tdp"]], lex.silence_state)
    ctx.set_lexicon(lex.automata, lex.silence)
    for ci, (wp, thr) in enumerate(g["configs"]):
        got, _ = ctx.recognize(c["frame_off"], float(wp), float(thr))
        words, off = g["words%d" % ci], g["off%d" % ci]
        for u in range(len(c["frame_off"]) - 1):
            assert np.array_equal(got[u], words[off[u]:off[u + 1]]), (ci, u)
    ctx.close()


def test_stress_shape_config5():
    """BASELINE config 5 shape: 4501 tied states x 64 densities, d = 48, per-state pooled variances; utterances are
    chains of 50 x 3 random states.  A few utterances against the oracle: alignment, path costs, argmin densities
    and accumulator counts bit-exact; u32 density indices (288 064 densities > 65 535)."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(synth.SEED)
    n_states, nd, dim = 4501, 64, 48
    tables = synth.random_model(rng, dim, n_states, nd, "mixture")
    corp = synth.make_chain_corpus(rng, tables, dim, n_states, 6, T=400)
    O = Oracle(dim, n_states, "mixture", True, (3.0, 0.0, 30.0), 0)
    O.load_tables(tables)
    want_states, want_costs = O.align_batch(corp["x"], corp["frame_off"], corp["automata"], corp["aut_off"], threads=8)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
    ctx.set_tdp(3.0, 0.0, 30.0, 0)
    assert ctx.get_counter("scorer") == 3            # the tensor-core scorer takes this shape (7 k-steps)
    got, costs = ctx.align()
    assert np.array_equal(got["state"], want_states)
    assert np.array_equal(costs, want_costs)
    cnt, sx, sxx, ssum = ctx.accumulate(first_pass=False)
    dens, score = ctx.last_frame_results()
    want_dens = O.accumulate(want_states, corp["x"])
    assert np.array_equal(dens, want_dens)
    t = O.tables()
    assert np.array_equal(cnt, t["mean_wacc"][t["mean_idx"].astype(np.int64)])
    assert cnt.sum() == len(want_states) and len(cnt) == n_states * nd
    np.testing.assert_allclose(sx, t["mean_acc"].reshape(-1, dim)[t["mean_idx"].astype(np.int64)], rtol=STAT_RTOL, atol=0)
    ctx.close()


def test_align_accumulate_equals_the_two_calls(small33):
    """sgmm_align_accumulate (one call, the alignment copy overlapping the E-step) == sgmm_align + sgmm_accumulate."""
    from speechrecognition_b200 import capi
    O, corp, ali, ctx = small33
    got, costs = ctx.align()
    cnt, sx, sxx, ssum = ctx.accumulate(first_pass=False)
    D, d, F, U = ctx.D, ctx.dim, ctx.n_frames, len(corp["frame_off"]) - 1
    a2 = capi.pinned_empty((F,), capi.ALI_DTYPE)
    c2 = capi.pinned_empty((U,), np.float64)
    st = (capi.pinned_empty((D,), np.float64), capi.pinned_empty((D, d), np.float64), capi.pinned_empty((D, d), np.float64))
    for _ in range(2):
        a2[...] = 0; c2[...] = 0
        r = ctx.align_accumulate(a2, c2, st)
        assert np.array_equal(a2["state"], got["state"]) and np.array_equal(a2["count"], got["count"])
        assert np.array_equal(c2, costs)
        assert np.array_equal(r[0], cnt) and np.array_equal(r[1], sx) and np.array_equal(r[2], sxx) and r[3] == ssum


@pytest.mark.parametrize("beam", [float("inf"), 60.0, 15.0])
def test_multiwarp_viterbi_long_utterances(beam):
    """k_viterbi_mw (utterances of >= 384 frames: positions split over four warps, one CTA barrier per frame) against the
    oracle and against the one-warp kernel (option vit_mw = 0): states and f64 path costs bit-identical in the full aligner
    (source convention) and in the pruned one (destination convention, finite beams included)."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(77)
    dim = 25
    lex = synth.Lexicon()
    tables = synth.random_model(rng, dim, lex.num_states, 3, "none")
    corp = synth.make_corpus(rng, lex, tables, dim, 14, t_choices=([500, 701, 1399, 279], [0.4, 0.2, 0.2, 0.2]), words_per_utt=(3, 7))
    # one very long automaton (S > 256: four positions per lane over four warps)
    O = Oracle(dim, lex.num_states, "none", True, helpers.TDP, lex.silence_state)
    O.load_tables(tables)
    res = {}
    for mw in (1, 0):
        ctx = capi.Context(0)
        ctx.set_option("vit_mw", mw)
        ctx.set_model_tables(dim, tables)
        ctx.set_features(corp["x"])
        ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
        ctx.set_tdp(*helpers.TDP, lex.silence_state)
        full = ctx.align(capi.TDP_SOURCE)
        ctx.set_beam(beam)
        pruned = ctx.align(capi.TDP_DESTINATION)
        res[mw] = (full[0]["state"].copy(), full[1].copy(), pruned[0]["state"].copy(), pruned[1].copy())
        ctx.close()
    for a, b in zip(res[1], res[0]):
        assert np.array_equal(a, b)
    want_states, want_costs = O.align_batch(corp["x"], corp["frame_off"], corp["automata"], corp["aut_off"], threads=8)
    assert np.array_equal(res[1][0].astype(np.uint32), want_states) and np.array_equal(res[1][1], want_costs)
    thr = 1e300 if beam == float("inf") else beam
    for u in range(len(corp["frame_off"]) - 1):
        f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
        aut = corp["automata"][corp["aut_off"][u]:corp["aut_off"][u + 1]]
        c, st, _ = O.align_pruned(corp["x"][f0:f1], aut, thr)
        assert c == res[1][3][u], (u, f1 - f0, len(aut))
        assert np.array_equal(res[1][2][f0:f1].astype(np.uint32), st), (u, f1 - f0, len(aut))


def test_pair_verdicts_planted_near_ties():
    """Pair verdicts of the tensor filter (option tc_pairs): near-duplicate densities planted so that the two candidates
    of a cell sit in the same 64-column sub-part, in different sub-parts of one column part, in different column parts and
    in different tiles; exact duplicates (the LOWER index must win, as in the reference's scan) and a triple (more than
    two candidates: rescanned).  Dense scores / argmins (frame-major refinement) and alignments (state-major refinement)
    equal the oracle's with and without the option, and the option takes cells out of the rescan queue."""
    from oracle.binding import Oracle
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(77)
    dim, nd = 33, 300                                  # 300 densities: two tiles (256 + 44 columns) per state
    lex = synth.Lexicon()
    t = synth.random_model(rng, dim, lex.num_states, nd, "mixture")
    means = t["means"].reshape(-1, dim)
    off = t["mix_offsets"]
    plant = {0: ((0, 1, 1e-9),),                       # same sub-part
             1: ((3, 100, 1e-9),),                     # sub-parts 0 and 1 of column part 0
             2: ((5, 200, 0.0),),                      # column parts 0 and 1, EXACT duplicate: index 5 wins
             3: ((7, 280, 1e-10),),                    # tiles 0 and 1
             4: ((2, 70, 0.0), (2, 140, 0.0)),         # a triple of exact duplicates: uncertain, rescanned, index 2 wins
             5: ((63, 64, 1e-9),)}                     # across the sub-part boundary
    for s, pairs in plant.items():
        for a, b, rel in pairs:
            means[off[s] + b] = means[off[s] + a] * (1.0 + rel)
            t["mean_wlog"][off[s] + b] = t["mean_wlog"][off[s] + a]
            t["mean_w"][off[s] + b] = t["mean_w"][off[s] + a]
    # frames sitting ON the planted densities, so that those are the winners
    corp = synth.make_corpus(rng, lex, t, dim, 8)
    x = corp["x"].copy()
    for i, (s, pairs) in enumerate(plant.items()):
        for r in range(6):
            x[40 * i + r] = (means[off[s] + pairs[0][0]] + 1e-3 * rng.standard_normal(dim)).astype(np.float32)
    O = Oracle(dim, lex.num_states, "mixture", True, helpers.TDP, lex.silence_state)
    O.load_tables(t)
    n = 400
    want_s, want_a = O.min_score(x[:n])
    want_states, want_costs = O.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"], threads=8)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, t); ctx.set_features(x)
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(*helpers.TDP, lex.silence_state)
    ctx.set_option("scorer", 3)
    assert ctx.get_counter("tc_form") == 1
    fb = {}
    for pairs in (1, 0):
        ctx.set_option("tc_pairs", pairs)
        ctx.get_counter("fallbacks")
        got_s, got_a = ctx.score_all(0, n)
        assert np.array_equal(got_a, want_a) and np.array_equal(got_s, want_s), pairs
        got, costs = ctx.align()
        fb[pairs] = ctx.get_counter("fallbacks")
        assert np.array_equal(got["state"].astype(np.uint32), want_states) and np.array_equal(costs, want_costs), pairs
    # the planted winners really are ties the filter cannot certify ...
    assert fb[0] >= 30, fb
    # ... and pair verdicts resolve most of them in place (the triple and the NaN-free leftovers stay)
    assert fb[1] < fb[0], fb
    for i, (s, pairs) in enumerate(plant.items()):
        assert (want_a[40 * i:40 * i + 6, s] == min(p[0] for p in pairs)).all() or pairs[0][2] != 0.0
    ctx.close()
