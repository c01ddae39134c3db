"""-m gpu: the C++ host classes (MixtureModel / Aligner / Trainer, csrc/host) driving the CUDA path, against
the oracle doing the same steps on the CPU.  These read like the reference's own usage (SieTill.cpp:107-113,
Training.cpp:51-235): construct the model, train, look at the AM-score trajectory and the alignment."""
import numpy as np
import pytest

from oracle.binding import Oracle
from speechrecognition_b200 import hostapi, synth
from tests import helpers

pytestmark = pytest.mark.gpu

TDP = (3.0, 0.0, 30.0)


def words_of(corp):
    w = np.concatenate(corp["words"]).astype(np.uint64)
    off = np.concatenate([[0], np.cumsum([len(a) for a in corp["words"]])]).astype(np.uint64)
    return w, off


def oracle_train(O, corp, min_obs, num_splits, num_aligns, num_estimates, pruned=False, initial=None):
    """Trainer::train (Training.cpp:114-226) spelled out with the oracle's primitives."""
    x = corp["x"]
    ali = initial if initial is not None else synth.linear_alignment(corp["frame_off"], corp["automata"], corp["aut_off"])
    stats = []
    O.accumulate(ali, x, first_pass=True); O.finalize()
    stats.append((-1, 0, 0, O.am_score(ali, x)))
    for i in range(num_splits + 1):
        if i > 0:
            O.split(2 * min_obs)
            O.accumulate(ali, x); O.finalize()
            O.eliminate(min_obs)
            O.accumulate(ali, x); O.finalize()
            stats.append((i, -1, 0, O.am_score(ali, x)))
        for j in range(num_aligns):
            if pruned:
                parts = []
                for u in range(len(corp["frame_off"]) - 1):
                    f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
                    aut = corp["automata"][int(corp["aut_off"][u]):int(corp["aut_off"][u + 1])]
                    parts.append(O.align_pruned(x[f0:f1], aut, 1e300 if pruned is True else float(pruned))[1])
                ali = np.concatenate(parts)
            else:
                ali, _ = O.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"], threads=8)
            for k in range(1 if i == 0 else num_estimates):
                O.accumulate(ali, x); O.finalize()
                stats.append((i, j, k, O.am_score(ali, x)))
    return np.array(stats), ali


@pytest.mark.parametrize("pooling,pruned,segmentation", [("none", False, "given"), ("mixture", False, "given"),
                                                         ("global", False, "given"), ("none", True, "given"),
                                                         ("none", 50.0, "given"), ("none", 12.0, "given"),
                                                         ("none", False, "reference")])
def test_trainer_trajectory_matches_oracle(pooling, pruned, segmentation, tmp_path):
    rng = np.random.default_rng(21)
    dim, min_obs = 12, 12
    lex = synth.Lexicon()
    gen = synth.random_model(rng, dim, lex.num_states, 2, "none")
    corp = synth.make_corpus(rng, lex, gen, dim, 40, t_choices=([90, 140, 200], [0.5, 0.3, 0.2]), words_per_utt=(1, 3))
    O = Oracle(dim, lex.num_states, pooling, True, TDP, lex.silence_state)
    if segmentation == "given":      # an initial alignment handed to Trainer::train
        init = synth.linear_alignment(corp["frame_off"], corp["automata"], corp["aut_off"])
    else:                            # the Trainer's own initial segmentation (Training.cpp:78-99, pinned against the
        parts = []                   # reference in tests/test_oracle_vs_reference.py), energy contour on c0
        for u in range(len(corp["frame_off"]) - 1):
            f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
            aut = corp["automata"][int(corp["aut_off"][u]):int(corp["aut_off"][u + 1])]
            parts.append(hostapi.linear_segmentation(corp["x"][f0:f1], aut, approx=True)[0])
        init = np.concatenate(parts)
    want_stats, want_ali = oracle_train(O, corp, min_obs, 2, 2, 2, pruned, initial=init)

    w, wo = words_of(corp)
    stats, ali, model, timers = hostapi.train(dim, pooling, corp["x"], corp["frame_off"], w, wo, TDP, min_obs, 2, 2, 2,
                                              alignment_pruning=bool(pruned), initial=init if segmentation == "given" else None,
                                              pruning_threshold=float("inf") if isinstance(pruned, bool) else pruned,
                                              mixture_path=str(tmp_path) + "/", alignment_path=str(tmp_path) + "/")
    # same lines in the same order as the reference's training-stats file (Training.cpp:126,158,205)
    assert np.array_equal(stats[:, :3], want_stats[:, :3])
    np.testing.assert_allclose(stats[:, 3], want_stats[:, 3], rtol=1e-10)
    assert np.array_equal(ali["state"].astype(np.uint32), want_ali)
    got, want = model.tables(), O.tables()
    for k in ("mix_offsets", "mean_idx", "var_idx", "mean_refs", "var_refs", "mean_wacc", "var_wacc"):
        assert np.array_equal(got[k], want[k]), k          # topology and counts: exact
    for k in ("means", "mean_acc", "var_acc", "vars_inv", "norm", "mean_wlog"):
        np.testing.assert_allclose(got[k], want[k], rtol=1e-9, atol=1e-12, equal_nan=True, err_msg=k)
    # artefacts the trainer leaves behind (Training.cpp:131-136,185-193,215-224)
    for name in ("lin.mix", "0.mix", "1.mix", "2.mix", "0-0.dump", "2-1.dump"):
        assert (tmp_path / name).exists(), name
    from speechrecognition_b200 import mixio
    dumped, ma = mixio.read_alignment(tmp_path / "2-1.dump")
    assert ma == 1 and np.array_equal(dumped["state"], ali["state"])
    reread = hostapi.HostModel(dim, lex.num_states, pooling, host_only=True, load_from=str(tmp_path / "2.mix"))
    assert reread.num_densities() == model.num_densities() == O.num_densities()
    assert timers["align"] > 0 and timers["estimate"] > 0


@pytest.mark.parametrize("pooling", ["mixture", "none"])
def test_trainer_without_max_approx_matches_oracle(pooling):
    """SURVEY §8 f4 through the host classes: MixtureModel / Trainer built with max_approx == false — sum scores in
    the alignment and the AM score (Mixtures.cpp:719-744), posterior-weighted E-step (:300-357)."""
    rng = np.random.default_rng(33)
    dim, min_obs = 12, 12
    lex = synth.Lexicon()
    gen = synth.random_model(rng, dim, lex.num_states, 2, "none")
    corp = synth.make_corpus(rng, lex, gen, dim, 40, t_choices=([90, 140, 200], [0.5, 0.3, 0.2]), words_per_utt=(1, 3))
    O = Oracle(dim, lex.num_states, pooling, False, TDP, lex.silence_state)
    init = synth.linear_alignment(corp["frame_off"], corp["automata"], corp["aut_off"])
    want_stats, want_ali = oracle_train(O, corp, min_obs, 2, 2, 2, False, initial=init)
    w, wo = words_of(corp)
    stats, ali, model, _ = hostapi.train(dim, pooling, corp["x"], corp["frame_off"], w, wo, TDP, min_obs, 2, 2, 2,
                                         initial=init, max_approx=False)
    assert np.array_equal(stats[:, :3], want_stats[:, :3])
    np.testing.assert_allclose(stats[:, 3], want_stats[:, 3], rtol=1e-10)
    assert np.array_equal(ali["state"].astype(np.uint32), want_ali)
    got, want = model.tables(), O.tables()
    for k in ("mix_offsets", "mean_idx", "var_idx", "mean_refs", "var_refs"):
        assert np.array_equal(got[k], want[k]), k          # topology: exact
    for k in ("mean_wacc", "var_wacc", "means", "mean_acc", "var_acc", "vars_inv", "norm", "mean_wlog"):
        np.testing.assert_allclose(got[k], want[k], rtol=1e-9, atol=1e-12, equal_nan=True, err_msg=k)
    # FeatureScorer::score of such a model is the sum score, prepared or one frame at a time
    x = corp["x"][:40]
    want_sc = O.score(x)
    np.testing.assert_allclose(model.scores(x, prepared=True), want_sc, rtol=1e-12)
    np.testing.assert_allclose(model.scores(x[:3], prepared=False), want_sc[:3], rtol=1e-12)


@pytest.fixture(scope="module")
def trained():
    O, corp, ali = helpers.trained_oracle(dim=25, n_utts=12, splits=2)
    H = hostapi.HostModel(25, O.n_mix, "none")
    H.load_tables(O.tables())
    yield O, corp, ali, H
    H.close()


def test_feature_scorer_idiom(trained):
    """prepare_sequence + score()/min_score() (FeatureScorer.hpp:12-16) and the unprepared one-frame path."""
    O, corp, ali, H = trained
    x = corp["x"][:150]
    want_s, want_a = O.min_score(x)
    got_s, got_a = H.score_sequence(x, prepared=True)
    assert np.array_equal(got_s, want_s) and np.array_equal(got_a, want_a)
    got_s, got_a = H.score_sequence(x[:6], prepared=False)
    assert np.array_equal(got_s, want_s[:6]) and np.array_equal(got_a, want_a[:6])


def test_per_utterance_aligner_signatures(trained):
    """Aligner::align_sequence_full / align_sequence_pruned keep their reference signatures (Alignment.hpp:46-55)."""
    O, corp, ali, H = trained
    for u in (0, 5):
        f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
        aut = corp["automata"][int(corp["aut_off"][u]):int(corp["aut_off"][u + 1])]
        want_cost, want_states, _ = O.align_full(corp["x"][f0:f1], aut)
        cost, got = H.align_sequence(TDP, O.silence, corp["x"][f0:f1], aut)
        assert cost == want_cost and np.array_equal(got["state"].astype(np.uint32), want_states)
        assert np.all(got["count"] == 1) and np.all(got["weight"] == 1.0)
        want_cost, want_states, _ = O.align_pruned(corp["x"][f0:f1], aut, 1e300)
        cost, got = H.align_sequence(TDP, O.silence, corp["x"][f0:f1], aut, pruned=True, threshold=1e300)
        assert cost == want_cost and np.array_equal(got["state"].astype(np.uint32), want_states)
        for thr in (4.0, 35.0):
            want_cost, want_states, _ = O.align_pruned(corp["x"][f0:f1], aut, thr)
            cost, got = H.align_sequence(TDP, O.silence, corp["x"][f0:f1], aut, pruned=True, threshold=thr)
            assert cost == want_cost and np.array_equal(got["state"].astype(np.uint32), want_states)


def test_model_accumulate_and_edit_detection(trained):
    """MixtureModel::accumulate folds the GPU statistics into the reference's slots; editing the host
    alignment in place between two calls is noticed (checksum) and re-uploaded."""
    O, corp, ali, H = trained
    x = corp["x"]
    a = hostapi.as_alignment(ali)
    O.load_tables(H.tables())
    ssum = H.accumulate(a, x)
    am = O.am_score(ali, x)
    O.accumulate(ali, x)
    got, want = H.tables(), O.tables()
    assert np.array_equal(got["mean_wacc"], want["mean_wacc"]) and np.array_equal(got["var_wacc"], want["var_wacc"])
    np.testing.assert_allclose(got["mean_acc"], want["mean_acc"], rtol=1e-12, atol=1e-9)
    np.testing.assert_allclose(got["var_acc"], want["var_acc"], rtol=1e-12, atol=1e-9)
    assert abs(ssum / len(ali) - am) <= 1e-12 * abs(am)
    # edit in place: move the first utterance's first frames to another state of its automaton
    ali2 = ali.copy()
    ali2[3:9] = ali2[20]
    a["state"][:] = ali2.astype(np.uint16)
    H.accumulate(a, x)
    O.accumulate(ali2, x)
    assert np.array_equal(H.tables()["mean_wacc"], O.tables()["mean_wacc"])


def test_recognizer_matches_oracle(trained):
    """Recognizer::recognize through the host class (one batched GPU search) == orc_recognize per utterance, and the
    WER bookkeeping (edit distance incl. the reference's column-0 behaviour) is consistent with the hypotheses."""
    O, corp, ali, H = trained
    lex = synth.Lexicon()
    w, wo = words_of(corp)
    word_off = np.concatenate([[0], np.cumsum([len(a) for a in lex.automata])]).astype(np.uint32)
    word_states = np.concatenate(lex.automata).astype(np.uint32)
    for wp, thr in ((10.0, 20.0), (4.0, 1e30)):
        hyps, stats = hostapi.recognize(H, corp["x"], corp["frame_off"], w, wo, TDP, wp, thr)
        errs = 0
        for u in range(len(corp["frame_off"]) - 1):
            f0, f1 = int(corp["frame_off"][u]), int(corp["frame_off"][u + 1])
            want, _ = O.recognize(corp["x"][f0:f1], word_off, word_states, lex.silence, wp, thr)
            assert np.array_equal(hyps[u], want), (u, hyps[u], want)
            errs += hostapi.edit_distance(corp["words"][u], hyps[u])[0]
        assert stats["errors"] == errs and stats["ref_words"] == len(w)
        assert stats["wer"] == pytest.approx(100.0 * errs / len(w))
