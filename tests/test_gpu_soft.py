"""-m gpu: the NON-max-approx branch through the C-ABI against the oracle — sum_score
(Mixtures.cpp:719-728) in dense scoring, alignment and the AM score; the posterior-weighted E-step (:300-357).

Density scores are bit-identical to the reference; exp / log are the device's f64 functions (<= 1 ulp from
glibc's), so these comparisons use rtol = 1e-12 (north_star allows 1e-5) and the alignments must be equal.
The oracle's own non-max-approx branch is pinned bit for bit against the unmodified reference in
tests/test_oracle_vs_reference.py::test_sum_score_and_soft_accumulation_live."""
import numpy as np
import pytest

from oracle.binding import Oracle
from tests import helpers

pytestmark = pytest.mark.gpu

RTOL = 1e-12


def sum_model(pooling, dim, seed):
    O, corp, ali = helpers.trained_oracle(dim=dim, n_utts=12, splits=2, pooling=pooling, seed=seed)
    P = Oracle(dim, O.n_mix, pooling, False, O.tdp, O.silence)      # same parameters, score() = sum_score
    P.load_tables(O.tables())
    ctx = helpers.gpu_context(P, corp, dim)
    ctx.set_option("max_approx", 0)
    return P, corp, ali, ctx


@pytest.fixture(scope="module", params=[("none", 25, 3), ("mixture", 33, 5)])
def soft(request):
    P, corp, ali, ctx = sum_model(*request.param)
    yield P, corp, ali, ctx, request.param[0]
    ctx.close()


def test_sum_scores_dense(soft):
    P, corp, ali, ctx, _ = soft
    scores, argmin = ctx.score_all(0, 400)
    want = P.score(corp["x"][:400])
    assert np.array_equal(np.isinf(scores), np.isinf(want))       # emptied mixtures: -log(0) = +inf on both sides
    np.testing.assert_allclose(scores, want, rtol=RTOL)
    assert not argmin.any()
    # and the max-approx answer is something else
    ctx.set_option("max_approx", 1)
    mins, _ = ctx.score_all(0, 400)
    ctx.set_option("max_approx", 0)
    fin = np.isfinite(want) & (want < 600.0)     # beyond that exp(-score) is subnormal and -log of it is coarse
    assert np.all(mins[fin] >= scores[fin] - 1e-9) and np.any(mins[fin] > scores[fin] + 1e-6)


def test_pruned_sum_equals_full_sum(soft):
    """The f32-pruned kernel (k_sum_pruned: densities certainly more than 45 above the state's minimum are left
    out) against the plain f64 sum over every density: at most a rounding apart."""
    P, corp, ali, ctx, _ = soft
    ctx.set_option("sum_prune", 0)
    full, _ = ctx.score_all(0, 600)
    ctx.set_option("sum_prune", 1)
    pruned, _ = ctx.score_all(0, 600)
    assert np.array_equal(np.isinf(full), np.isinf(pruned))
    fin = np.isfinite(full)
    assert np.max(np.abs(pruned[fin] - full[fin]) / np.abs(full[fin])) <= 4.5e-16
    assert np.mean(pruned[fin] == full[fin]) > 0.99


def test_align_with_sum_scores(soft):
    P, corp, ali, ctx, _ = soft
    got, costs = ctx.align()
    want_states, want_costs = P.align_batch(corp["x"], corp["frame_off"], corp["automata"], corp["aut_off"])
    assert np.array_equal(got["state"].astype(np.uint32), want_states)
    np.testing.assert_allclose(costs, want_costs, rtol=RTOL)
    # AM score through score() = sum_score (Training.cpp:585-612)
    am = P.am_score(want_states, corp["x"])
    assert abs(ctx.am_score_sum() / len(want_states) - am) <= RTOL * abs(am)


def test_accumulate_soft_matches_oracle(soft):
    P, corp, ali, ctx, pooling = soft
    dim = P.dim
    ctx.set_alignment(ali)
    cnt, sx, sxx, ssum = ctx.accumulate_soft()
    Q = Oracle(dim, P.n_mix, pooling, False, P.tdp, P.silence)
    Q.load_tables(P.tables())
    Q.accumulate(ali, corp["x"], False, False)
    t = Q.tables()
    want_cnt, want_sx = helpers.packed_stats(t, dim)
    np.testing.assert_allclose(cnt, want_cnt, rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(sx, want_sx, rtol=1e-11, atol=1e-9)
    # variance statistics: fold the per-density sums into the reference's (possibly pooled) slots; it seeds 1e-4
    vi = t["var_idx"].astype(np.int64)
    fold = np.zeros_like(t["var_acc"].reshape(-1, dim))
    np.add.at(fold, vi, sxx)
    used = np.zeros(len(fold), bool)
    used[vi] = True
    np.testing.assert_allclose(fold[used], t["var_acc"].reshape(-1, dim)[used] - 1e-4, rtol=1e-9, atol=1e-7)
    wfold = np.zeros(len(fold))
    np.add.at(wfold, vi, cnt)
    np.testing.assert_allclose(wfold[used], t["var_wacc"][used], rtol=1e-11, atol=1e-12)
    # posteriors of a frame sum to 1 (minus the skipped entries below 1e-8): total weight = frames aligned to states
    # that still have densities
    live = np.diff(t["mix_offsets"])[ali] > 0
    assert abs(cnt.sum() - live.sum()) < 1e-4
    # by-product: calc_am_score's numerator with sum scores
    am = P.am_score(ali, corp["x"])
    assert np.isinf(am) or abs(ssum / len(ali) - am) <= RTOL * abs(am)
    am2 = ctx.am_score_sum()              # the f32-pruned kernel: at most a rounding per frame from the full sums
    assert (np.isinf(ssum) and np.isinf(am2)) or abs(am2 - ssum) <= 1e-13 * abs(ssum)
    ctx.set_option("sum_prune", 0)
    assert ctx.am_score_sum() == ssum or np.isinf(ssum)      # the plain kernel adds the very same terms
    ctx.set_option("sum_prune", 1)
    # deterministic: same bits on a second run
    again = ctx.accumulate_soft()
    for a, b in zip((cnt, sx, sxx), again[:3]):
        assert np.array_equal(a, b)


def test_max_approx_estep_after_sum_alignment(soft):
    """accumulate(max_approx = true) on an alignment computed with sum scores must not look the (sum) scores up."""
    P, corp, ali, ctx, pooling = soft
    got, _ = ctx.align()
    states = got["state"].astype(np.uint32)
    cnt, sx, sxx, _ = ctx.accumulate(first_pass=False)
    Q = Oracle(P.dim, P.n_mix, pooling, True, P.tdp, P.silence)
    Q.load_tables(P.tables())
    want_dens = Q.accumulate(states, corp["x"], False, True)
    dens, _ = ctx.last_frame_results()
    assert np.array_equal(dens, want_dens)
    want_cnt, want_sx = helpers.packed_stats(Q.tables(), P.dim)
    assert np.array_equal(cnt, want_cnt)
    np.testing.assert_allclose(sx, want_sx, rtol=1e-12, atol=1e-9)


def test_soft_errors_are_loud():
    from speechrecognition_b200 import capi
    ctx = capi.Context(0)
    with pytest.raises(capi.SgmmError):
        ctx.accumulate_soft()             # no model / features / alignment
    ctx.close()


def test_soft_accumulation_many_states():
    """More than 256 states: the stable sort by aligned state takes two radix passes; odd dimension, 5 densities per
    state, a random alignment (every state receives frames)."""
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(12)
    n_states, nd, dim = 700, 5, 9
    tables = synth.random_model(rng, dim, n_states, nd, "none")
    corp = synth.make_chain_corpus(rng, tables, dim, n_states, 12, T=300, phones=30, spp=3)
    x = corp["x"]
    ali = rng.integers(0, n_states, size=len(x)).astype(np.uint32)
    P = Oracle(dim, n_states, "none", False, (3.0, 0.0, 30.0), 0)
    P.load_tables(tables)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(x)
    ctx.set_option("max_approx", 0)
    ctx.set_alignment(ali)
    cnt, sx, sxx, ssum = ctx.accumulate_soft()
    P.accumulate(ali, x, False, False)
    t = P.tables()
    want_cnt, want_sx = helpers.packed_stats(t, dim)
    np.testing.assert_allclose(cnt, want_cnt, rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(sx, want_sx, rtol=1e-11, atol=1e-9)
    vi = t["var_idx"].astype(np.int64)
    np.testing.assert_allclose(sxx, t["var_acc"].reshape(-1, dim)[vi] - 1e-4, rtol=1e-9, atol=1e-7)
    # a frame far from every density of its state: exp underflows, the normaliser is 0, the reference accumulates NaN
    x2 = x.copy()
    x2[5] = 1e4
    ctx.set_features(x2)
    ctx.set_alignment(ali)
    cnt2, _, _, ssum2 = ctx.accumulate_soft()
    Q = Oracle(dim, n_states, "none", False, (3.0, 0.0, 30.0), 0)
    Q.load_tables(tables)
    Q.accumulate(ali, x2, False, False)
    want2, _ = helpers.packed_stats(Q.tables(), dim)
    assert np.array_equal(np.isnan(cnt2), np.isnan(want2)) and np.isnan(cnt2).sum() == nd
    assert np.isinf(ssum2)
    ctx.close()


@pytest.mark.parametrize("dim,nd", [(33, 200), (32, 256), (25, 70)])
def test_sum_scores_on_the_tensor_path(dim, nd):
    """Option sum_tc (default on): the gathered alignment pass with sum scores takes the tensor filter's candidate masks
    (densities within 45 of a cell's best score, a certified superset) and the state-major kernel in SUM form instead of
    k_sum_pruned's f32 stage.  Same criterion, same f64 evaluation, same summation order: the two paths must agree to the
    last bit on alignments and to 1e-15 on path costs, and both with the oracle (1e-12).  Includes a frame with a huge
    feature (its row of the filter is NaN: every density is a candidate) and planted duplicate densities."""
    from speechrecognition_b200 import capi, synth
    rng = np.random.default_rng(100 + dim)
    lex = synth.Lexicon()
    t = synth.random_model(rng, dim, lex.num_states, nd, "mixture")
    means = t["means"].reshape(-1, dim)
    off = t["mix_offsets"]
    means[off[2] + 9] = means[off[2] + 1]                 # duplicates: both within 45 of each other, always
    means[off[3] + nd - 1] = means[off[3]] * (1 + 1e-7)
    corp = synth.make_corpus(rng, lex, t, dim, 20)
    x = corp["x"].copy()
    x[11, 3] = 3.0e6                                       # beyond the fp16 row scaling
    P = Oracle(dim, lex.num_states, "mixture", False, helpers.TDP, lex.silence_state)
    P.load_tables(t)
    want_states, want_costs = P.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"], threads=8)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, t); ctx.set_features(x)
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(*helpers.TDP, lex.silence_state)
    ctx.set_option("scorer", 3)
    ctx.set_option("max_approx", 0)
    got = {}
    for tc in (1, 0):
        ctx.set_option("sum_tc", tc)
        ali, costs = ctx.align()
        st = ctx.stage_ms()
        got[tc] = (ali["state"].copy(), costs.copy(), st["score_filter"])
        fin = np.isfinite(want_costs)
        assert np.array_equal(ali["state"].astype(np.uint32), want_states), tc
        np.testing.assert_allclose(costs[fin], want_costs[fin], rtol=RTOL)
        am = P.am_score(want_states, x)
        assert abs(ctx.am_score_sum() / len(want_states) - am) <= RTOL * abs(am) or not np.isfinite(am)
    # the tensor filter ran in the first pass only — and not at all for the narrow model (70 densities per state take
    # the 128-column tiles, which have no mask mode: the option falls back to k_sum_pruned by itself)
    assert (got[1][2] > 0.0) == (nd > 128) and got[0][2] == 0.0
    assert np.array_equal(got[1][0], got[0][0])
    fin = np.isfinite(got[0][1])
    np.testing.assert_allclose(got[1][1][fin], got[0][1][fin], rtol=1e-15)
    ctx.close()
