"""CPU: the oracle (oracle/sietill_oracle.c) against the committed fixtures that tools/make_golden.py produced
with the UNMODIFIED reference (oracle/_ref).  Everything is compared bit for bit (NaNs included: states that
never received a frame have 0/0 means in the reference, Mixtures.cpp:391-394)."""
import os

import numpy as np
import pytest

from oracle.binding import Oracle, TABLES
from speechrecognition_b200 import synth

G = os.path.join(os.path.dirname(__file__), "golden")


def same_tables(got, rec, prefix):
    for k in TABLES:
        want = rec[prefix + k]
        assert got[k].shape == want.shape, (prefix, k)
        assert np.array_equal(got[k], want, equal_nan=True), (prefix, k)


@pytest.mark.parametrize("pooling", ["none", "mixture", "global"])
def test_training_trajectory_matches_reference(pooling):
    c = np.load(os.path.join(G, "corpus.npz"))
    rec = np.load(os.path.join(G, "trajectory_%s.npz" % pooling))
    x, dim, min_obs = c["x"], int(c["dim"]), int(c["min_obs"])
    lex = synth.Lexicon()
    O = Oracle(dim, lex.num_states, pooling, True, tuple(c["tdp"]), lex.silence_state)
    ali = synth.linear_alignment(c["frame_off"], c["automata"], c["aut_off"])
    assert np.array_equal(ali, rec["ali_linear"])
    O.accumulate(ali, x, first_pass=True); O.finalize()
    same_tables(O.tables(), rec, "t0_")
    assert O.am_score(ali, x) == float(rec["am0"])
    O.split(2 * min_obs)
    same_tables(O.tables(), rec, "tsplit_")
    O.accumulate(ali, x); O.finalize()
    O.eliminate(min_obs)
    O.accumulate(ali, x); O.finalize()
    same_tables(O.tables(), rec, "t1_")
    st, costs = O.align_batch(x, c["frame_off"], c["automata"], c["aut_off"])
    assert np.array_equal(st, rec["ali_full"])
    assert np.array_equal(costs, rec["costs_full"])
    f0, f1 = int(c["frame_off"][2]), int(c["frame_off"][3])
    aut = c["automata"][c["aut_off"][2]:c["aut_off"][3]]
    for thr, tag in ((40.0, "p40"), (1e300, "pinf")):
        cost, s, _ = O.align_pruned(x[f0:f1], aut, thr)
        assert cost == float(rec["cost_" + tag]) and np.array_equal(s, rec["ali_" + tag])
    sc, am = O.min_score(x[:64])
    assert np.array_equal(sc, rec["scores64"]) and np.array_equal(am, rec["argmin64"])
    O.accumulate(st, x)
    same_tables(O.tables(), rec, "t2_")
    O.finalize()
    assert O.am_score(st, x) == float(rec["am2"])


def test_tdp_model():
    """TdpModel::score, TdpModel.cpp:19-29"""
    from oracle.binding import Oracle
    import ctypes as C
    O = Oracle(2, 2)
    f = O.L.orc_tdp_score
    f.restype = C.c_double
    args = lambda s, j: (C.c_double(3.0), C.c_double(0.5), C.c_double(30.0), C.c_uint32(0), C.c_uint32(s), C.c_uint32(j))
    assert f(*args(0, 0)) == 0.5 and f(*args(0, 2)) == 0.5       # silence always pays tdp_forward
    assert f(*args(5, 0)) == 3.0 and f(*args(5, 1)) == 0.5 and f(*args(5, 2)) == 30.0
    assert f(*args(5, 3)) == float("inf")


def test_unreachable_and_degenerate_alignments():
    rng = np.random.default_rng(0)
    t = synth.random_model(rng, 4, 6, 2, "none")
    O = Oracle(4, 6, "none", True, (1.0, 0.0, 2.0), 0)
    O.load_tables(t)
    x = rng.standard_normal((1, 4)).astype(np.float32)
    cost, st, pos = O.align_full(x, [0, 1, 2])
    assert cost == float("inf")                                   # T = 1 returns +inf (Alignment.cpp:143)
    x = rng.standard_normal((5, 4)).astype(np.float32)
    cost, st, pos = O.align_full(x, [3])
    assert np.all(st == 3) and np.isfinite(cost)                  # single position: loops only
    cost, st, pos = O.align_full(x[:3], [1, 2, 3, 4, 5])         # S == 2T-1: skipping every frame is the only path
    assert np.array_equal(pos, [0, 2, 4]) and np.isfinite(cost)


def test_recognition_matches_reference():
    """orc_recognize against the committed outputs of the unmodified Recognizer (tests/golden/recognize.npz,
    tools/make_golden.py recognition()): final "none" model, every golden utterance, three beam / penalty settings."""
    c = np.load(os.path.join(G, "corpus.npz"))
    rec = np.load(os.path.join(G, "trajectory_none.npz"))
    g = np.load(os.path.join(G, "recognize.npz"))
    dim = int(c["dim"])
    lex = synth.Lexicon()
    O = Oracle(dim, lex.num_states, "none", True, tuple(c["tdp"]), lex.silence_state)
    O.load_tables({k: rec["t2_" + k] for k in TABLES})
    O.finalize()
    word_off = np.concatenate([[0], np.cumsum([len(a) for a in lex.automata])]).astype(np.uint32)
    word_states = np.concatenate(lex.automata).astype(np.uint32)
    total = 0
    for ci, (wp, thr) in enumerate(g["configs"]):
        words, off = g["words%d" % ci], g["off%d" % ci]
        for u in range(len(c["frame_off"]) - 1):
            x = c["x"][int(c["frame_off"][u]):int(c["frame_off"][u + 1])]
            got, _ = O.recognize(x, word_off, word_states, lex.silence, float(wp), float(thr))
            assert np.array_equal(got, words[off[u]:off[u + 1]]), (ci, u)
            total += len(got)
    assert total > 0


def _feature_cases():
    g = np.load(os.path.join(G, "features.npz"))
    for ci, (nf, n1, n2, step, norm, emax) in enumerate(g["cases"]):
        mean = g["mean%d" % ci] if norm else None
        std = g["std%d" % ci] if norm else None
        yield g["raw%d" % ci], g["off%d" % ci], g["out%d" % ci], int(n1), int(n2), int(step), mean, std, bool(emax)


def test_feature_processing_matches_reference():
    """SignalAnalysis::process_features (deltas, mean / variance and energy-max normalisation) —
    the oracle against the committed outputs of the unmodified reference, bit for bit."""
    from oracle.binding import oracle_process_features
    n = 0
    for raw, off, want, n1, n2, step, mean, std, emax in _feature_cases():
        for u in range(len(off) - 1):
            a, b = int(off[u]), int(off[u + 1])
            got = oracle_process_features(raw[a:b], n1, n2, step, mean, std, emax)
            assert np.array_equal(got.view(np.uint32), want[a:b].view(np.uint32))
            n += 1
    assert n == 20
