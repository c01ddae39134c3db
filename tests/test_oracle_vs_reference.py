"""CPU, build container only: the oracle against the live reference (oracle/_ref/libsietill_ref.so) on fresh
seeded inputs — every table after every EM operation, alignments, path costs, scores: bit-identical."""
import numpy as np
import pytest

from oracle import binding
from oracle.binding import Oracle, Reference, TABLES
from speechrecognition_b200 import synth

pytestmark = pytest.mark.skipif(not binding.have_ref(), reason="oracle/_ref not built (needs /root/reference)")


def same(a, b, tag):
    for k in TABLES:
        assert a[k].shape == b[k].shape and np.array_equal(a[k], b[k], equal_nan=True), (tag, k)


@pytest.mark.parametrize("pooling", ["none", "mixture", "global"])
def test_em_and_alignment_live(pooling):
    rng = np.random.default_rng(1)
    lex = synth.Lexicon()
    d = 25
    gen = synth.random_model(rng, d, lex.num_states, 2, "none")
    corp = synth.make_corpus(rng, lex, gen, d, 24)
    tdp = (3.0, 0.0, 30.0)
    O = Oracle(d, 106, pooling, True, tdp, 0)
    R = Reference(d, 106, pooling, True, tdp, 0)
    same(O.tables(), R.tables(), "init")
    ali = synth.linear_alignment(corp["frame_off"], corp["automata"], corp["aut_off"])
    x = corp["x"]
    O.accumulate(ali, x, True); R.accumulate(ali, x, True)
    same(O.tables(), R.tables(), "acc0")
    O.finalize(); R.finalize()
    same(O.tables(), R.tables(), "fin0")
    for it in range(2):
        if it > 0:
            O.split(30); R.split(30); same(O.tables(), R.tables(), "split")
            O.accumulate(ali, x); R.accumulate(ali, x); O.finalize(); R.finalize()
            O.eliminate(15); R.eliminate(15); same(O.tables(), R.tables(), "elim")
            O.accumulate(ali, x); R.accumulate(ali, x); O.finalize(); R.finalize()
            same(O.tables(), R.tables(), "fin")
            assert O.num_densities() == R.num_densities()
        so, co = O.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"], threads=4)
        sr, cr = R.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"])
        assert np.array_equal(so, sr) and np.array_equal(co, cr)
        f0, f1 = int(corp["frame_off"][3]), int(corp["frame_off"][4])
        aut = corp["automata"][corp["aut_off"][3]:corp["aut_off"][4]]
        for thr in (50.0, 1e9):
            c1, s1, _ = O.align_pruned(x[f0:f1], aut, thr)
            c2, s2 = R.align_pruned(x[f0:f1], aut, thr)
            assert c1 == c2 and np.array_equal(s1, s2)
        ali = so
        O.accumulate(ali, x); R.accumulate(ali, x); O.finalize(); R.finalize()
        same(O.tables(), R.tables(), "fin%d" % it)
        assert O.am_score(ali, x) == R.am_score(ali, x)
        s_o, a_o = O.min_score(x[:200]); s_r, a_r = R.min_score(x[:200])
        assert np.array_equal(s_o, s_r) and np.array_equal(a_o, a_r)


@pytest.mark.parametrize("scale,word_penalty,am_threshold", [(1.0, 10.0, 20.0), (1.0, 3.0, 1e30), (0.05, 10.0, 60.0)])
def test_recognize_live(scale, word_penalty, am_threshold):
    """orc_recognize against the unmodified Recognizer::recognizeSequence_pruned (Recognizer.cpp:103-232): default
    beam, no beam, and features scaled so that emission scores are NEGATIVE (the pre-emission rejection test of
    :142-145 then drops candidates a plain minimum would keep)."""
    rng = np.random.default_rng(5)
    lex = synth.Lexicon()
    d = 25
    gen = synth.random_model(rng, d, lex.num_states, 2, "none")
    corp = synth.make_corpus(rng, lex, gen, d, 16)
    x = (corp["x"] * scale).astype(np.float32)
    tdp = (3.0, 0.0, 30.0)
    O = Oracle(d, 106, "none", True, tdp, 0)
    R = Reference(d, 106, "none", True, tdp, 0, extra_config={"am-threshold": float(am_threshold),
                                                             "word-penalty": float(word_penalty)})
    ali = synth.linear_alignment(corp["frame_off"], corp["automata"], corp["aut_off"])
    for M in (O, R):
        M.accumulate(ali, x, True); M.finalize()
    ali, _ = O.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"], threads=4)
    for M in (O, R):
        M.accumulate(ali, x); M.finalize()
    word_off = np.concatenate([[0], np.cumsum([len(a) for a in lex.automata])]).astype(np.uint32)
    word_states = np.concatenate(lex.automata).astype(np.uint32)
    n_nonempty = 0
    for u in range(8):
        f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
        got, score = O.recognize(x[f0:f1], word_off, word_states, lex.silence, word_penalty, am_threshold)
        want = R.recognize(x[f0:f1])
        assert np.array_equal(got, want), (u, got, want)
        n_nonempty += len(want) > 0
    assert n_nonempty >= 4


def test_edit_distance_live():
    """The host mirror's editDistance (csrc/host/Recognizer.cpp) against Recognizer::editDistance (Recognizer.cpp:332-388)
    on random word sequences, including the reference's column-0 behaviour for leading insertions."""
    from speechrecognition_b200 import hostapi
    R = Reference(9, 106, "none", True, (3.0, 0.0, 30.0), 0)
    rng = np.random.default_rng(3)
    cases = [([1, 2], [5, 6, 1, 2]), ([], [4, 4, 4]), ([1, 2, 3], []), ([], [])]
    for _ in range(300):
        nr, nh = int(rng.integers(0, 9)), int(rng.integers(0, 9))
        cases.append((rng.integers(1, 5, nr).tolist(), rng.integers(1, 5, nh).tolist()))
    for ref, hyp in cases:
        assert hostapi.edit_distance(ref, hyp) == R.edit_distance(ref, hyp), (ref, hyp)


@pytest.mark.parametrize("approx", [True, False])
def test_linear_segmentation_live(approx):
    """The host Trainer's initial segmentation (csrc/host/Training.cpp) against the reference's
    linear_segmentation_approximation / linear_segmentation_running_sums (Training.cpp:350-546): boundaries and the
    frame -> state mapping, on utterances whose first component is a noisy silence / speech / silence energy contour."""
    from speechrecognition_b200 import hostapi
    R = Reference(9, 106, "none", True, (3.0, 0.0, 30.0), 0)
    rng = np.random.default_rng(17)
    lex = synth.Lexicon()
    for u in range(12):
        T = int(rng.integers(40, 260))
        b1, b2 = sorted(rng.integers(5, T - 5, 2))
        x = rng.normal(0, 1, (T, 9)).astype(np.float32)
        x[:, 0] = rng.normal(0.2, 0.3, T)
        x[b1:b2, 0] += rng.uniform(1.0, 4.0)
        aut = lex.segment_automaton(rng.integers(1, 12, int(rng.integers(1, 4))))
        got_states, got_b = hostapi.linear_segmentation(x, aut, approx)
        want_states, want_b = R.linear_segmentation(x, aut, approx)
        assert got_b == want_b, (u, got_b, want_b)
        assert np.array_equal(got_states, want_states), u


@pytest.mark.parametrize("pooling", ["none", "mixture"])
def test_sum_score_and_soft_accumulation_live(pooling):
    """models built with max_approx == false — sum_score (Mixtures.cpp:719-728) in scoring,
    alignment and the AM score, posterior-weighted accumulate (:300-357).  Oracle == reference bit for bit
    (same libm)."""
    rng = np.random.default_rng(11)
    lex = synth.Lexicon()
    d = 25
    gen = synth.random_model(rng, d, lex.num_states, 2, "none")
    corp = synth.make_corpus(rng, lex, gen, d, 16)
    tdp = (3.0, 0.0, 30.0)
    O = Oracle(d, 106, pooling, False, tdp, 0)
    R = Reference(d, 106, pooling, False, tdp, 0)
    ali = synth.linear_alignment(corp["frame_off"], corp["automata"], corp["aut_off"])
    x = corp["x"]
    O.accumulate(ali, x, True); R.accumulate(ali, x, True)
    O.finalize(); R.finalize()
    for it in range(2):
        O.split(30); R.split(30)
        O.accumulate(ali, x, False, False); R.accumulate(ali, x, False, False)
        same(O.tables(), R.tables(), "soft-acc")
        O.finalize(); R.finalize()
        same(O.tables(), R.tables(), "soft-fin")
        so, sr = O.score(x[:200]), R.score(x[:200])
        assert np.array_equal(so, sr, equal_nan=True)
        a_o, c_o = O.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"])
        a_r, c_r = R.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"])
        assert np.array_equal(a_o, a_r) and np.array_equal(c_o, c_r)
        assert O.am_score(a_o, x) == R.am_score(a_r, x)
        ali = a_o
    assert O.num_densities() > 150


@pytest.mark.parametrize("nf,n1,n2,step,norm,emax", [(12, 12, 1, 3, True, True), (12, 12, 1, 3, False, True),
                                                      (7, 7, 7, 1, True, False), (20, 4, 2, 5, True, True)])
def test_feature_processing_live(nf, n1, n2, step, norm, emax):
    """SignalAnalysis::process_features: oracle == unmodified reference, bit for bit, on fresh inputs."""
    rng = np.random.default_rng(nf * 100 + step)
    nt = nf + n1 + n2
    mean = rng.uniform(-5, 5, nt) if norm else None
    std = rng.uniform(0.2, 9.0, nt) if norm else None
    for T in (step + 1, step + 2, 50, 333):
        raw = (rng.standard_normal((T, nf)) * 7.0 + 3.0).astype(np.float32)
        a = binding.oracle_process_features(raw, n1, n2, step, mean, std, emax)
        b = binding.reference_process_features(raw, n1, n2, step, mean, std, emax)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
