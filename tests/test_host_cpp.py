"""CPU: the C++ host classes (speechrecognition_b200/csrc/host) that do NOT need a GPU — the M-step
(finalize / split / eliminate), the .mix format, lexicon/automata, TDP model, utterance sharding — against
the golden fixtures minted from the unmodified reference and against the oracle.  The E-step between the
M-steps is done by the oracle here (it is the checker); on a GPU box tests/test_host_gpu.py runs the same
trajectory with the CUDA E-step."""
import os

import numpy as np
import pytest

from oracle.binding import Oracle, TABLES
from speechrecognition_b200 import hostapi, synth

G = os.path.join(os.path.dirname(__file__), "golden")


def same_tables(got, want, prefix="", keys=TABLES):
    for k in keys:
        w = want[prefix + k]
        assert got[k].shape == w.shape, (prefix, k)
        assert np.array_equal(got[k], w, equal_nan=True), (prefix, k)


@pytest.mark.parametrize("pooling", ["none", "mixture", "global"])
def test_m_step_matches_reference_fixtures(pooling, tmp_path):
    """finalize / split / eliminate / write of the C++ MixtureModel reproduce, bit for bit, the tables the
    unmodified reference produced (tests/golden/trajectory_*.npz, Mixtures.cpp:374-576,845-878)."""
    c = np.load(os.path.join(G, "corpus.npz"))
    rec = np.load(os.path.join(G, "trajectory_%s.npz" % pooling))
    x, dim, min_obs = c["x"], int(c["dim"]), int(c["min_obs"])
    lex = synth.Lexicon()
    O = Oracle(dim, lex.num_states, pooling, True, tuple(c["tdp"]), lex.silence_state)   # E-step stand-in
    H = hostapi.HostModel(dim, lex.num_states, pooling, host_only=True)
    ali = rec["ali_linear"]

    def e_step(first=False):
        O.load_tables(H.tables())
        O.accumulate(ali, x, first_pass=first)
        H.load_tables(O.tables())

    e_step(True); H.finalize()
    same_tables(H.tables(), rec, "t0_")
    H.split(2 * min_obs)
    same_tables(H.tables(), rec, "tsplit_")
    e_step(); H.finalize()
    H.eliminate(min_obs)
    e_step(); H.finalize()
    same_tables(H.tables(), rec, "t1_")
    assert H.num_densities() == int(np.count_nonzero(rec["t1_mean_refs"]))
    # accumulators after the E-step on the Viterbi alignment -> the .mix file the reference wrote
    H.load_tables({k: rec["t2_" + k] for k in TABLES})
    out = tmp_path / "host.mix"
    H.write(out)
    assert out.read_bytes() == open(os.path.join(G, "mix_%s.mix" % pooling), "rb").read()


@pytest.mark.parametrize("pooling", ["none", "mixture", "global"])
def test_mix_read_regenerates_parameters(pooling, tmp_path):
    """read() = load statistics + topology, then finalize (Mixtures.cpp:748-843); write(read(f)) == f."""
    c = np.load(os.path.join(G, "corpus.npz"))
    rec = np.load(os.path.join(G, "trajectory_%s.npz" % pooling))
    dim = int(c["dim"])
    lex = synth.Lexicon()
    path = os.path.join(G, "mix_%s.mix" % pooling)
    H = hostapi.HostModel(dim, lex.num_states, pooling, host_only=True, load_from=path)
    again = tmp_path / "again.mix"
    H.write(again)
    if pooling != "mixture":
        assert again.read_bytes() == open(path, "rb").read()
    else:
        # read() recounts the references from the densities (Mixtures.cpp:752-800), so the variance slots of
        # mixtures that eliminate() emptied — still written by the trainer, whose refs stay 1 under pooling
        # (:566-569) — are dropped by a load/store cycle; after that the file is a fixed point.
        assert len(again.read_bytes()) < os.path.getsize(path)
        H2 = hostapi.HostModel(dim, lex.num_states, pooling, host_only=True, load_from=str(again))
        H2.write(tmp_path / "again2.mix")
        assert (tmp_path / "again2.mix").read_bytes() == again.read_bytes()
        return
    # parameters = finalize of the t2 statistics (indices compacted: dead slots are not stored)
    O = Oracle(dim, lex.num_states, pooling)
    O.load_tables({k: rec["t2_" + k] for k in TABLES})
    O.finalize()
    want, got = O.tables(), H.tables()
    live_m = np.nonzero(want["mean_refs"])[0]
    assert np.array_equal(got["means"].reshape(-1, dim), want["means"].reshape(-1, dim)[live_m], equal_nan=True)
    assert np.array_equal(got["mean_wlog"], want["mean_wlog"][live_m], equal_nan=True)
    live_v = np.nonzero(want["var_refs"])[0]
    assert np.array_equal(got["vars_inv"].reshape(-1, dim), want["vars_inv"].reshape(-1, dim)[live_v], equal_nan=True)
    assert np.array_equal(got["norm"], want["norm"][live_v], equal_nan=True)
    assert np.array_equal(got["mix_offsets"], want["mix_offsets"])


def test_mix_read_errors_are_loud(tmp_path):
    lex = synth.Lexicon()
    bad = tmp_path / "bad.mix"
    bad.write_bytes(b"NOTAMIX\x00" + b"\x00" * 64)
    with pytest.raises(hostapi.HostError, match="Invalid magic header"):
        hostapi.HostModel(9, lex.num_states, "none", host_only=True, load_from=str(bad))
    good = open(os.path.join(G, "mix_none.mix"), "rb").read()
    (tmp_path / "trunc.mix").write_bytes(good[:100])
    with pytest.raises(hostapi.HostError, match="Error reading"):
        hostapi.HostModel(9, lex.num_states, "none", host_only=True, load_from=str(tmp_path / "trunc.mix"))
    with pytest.raises(hostapi.HostError, match="Invalid dimension"):
        hostapi.HostModel(10, lex.num_states, "none", host_only=True, load_from=os.path.join(G, "mix_none.mix"))


def test_host_only_model_refuses_to_compute():
    """No CPU fallback: a model built without a GPU context cannot score, align or accumulate."""
    H = hostapi.HostModel(9, 106, "none", host_only=True)
    with pytest.raises(hostapi.HostError, match="no GPU context"):
        H.accumulate(np.zeros(4, np.uint32), np.zeros((4, 9), np.float32), first_pass=True)
    with pytest.raises(hostapi.HostError, match="no GPU context"):
        H.score_sequence(np.zeros((2, 9), np.float32))


def test_model_without_gpu_is_a_loud_error():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(hostapi.HostError, match="no CUDA device|sgmm_create"):
        hostapi.HostModel(9, 106, "none")


def test_lexicon_automata_and_tdp():
    lex = synth.Lexicon()
    rng = np.random.default_rng(3)
    for _ in range(5):
        w = rng.integers(1, 12, size=int(rng.integers(1, 8)))
        assert np.array_equal(hostapi.segment_automaton(w), lex.segment_automaton(w).astype(np.uint16))
    assert len(hostapi.segment_automaton([5])) == 1 + 24 + 1            # "fuenf": 12 states x 2 (Lexicon.cpp:76)
    tdp = (3.0, 0.5, 30.0)
    O = Oracle(2, 2)
    for state in (0, 7):
        for jump in (0, 1, 2, 3):
            want = O.L.orc_tdp_score
            want.restype = __import__("ctypes").c_double
            import ctypes as C
            w = want(C.c_double(tdp[0]), C.c_double(tdp[1]), C.c_double(tdp[2]), C.c_uint32(0), C.c_uint32(state), C.c_uint32(jump))
            assert hostapi.tdp_score(tdp, 0, state, jump) == w
    assert hostapi.tdp_score(tdp, 0, 0, 2) == 0.5 and hostapi.tdp_score(tdp, 0, 7, 3) == float("inf")


def test_shard_utterances_is_a_balanced_partition():
    rng = np.random.default_rng(1)
    T = rng.choice([279, 500, 701, 1399], size=1000, p=[0.903, 0.072, 0.005, 0.02])
    fo = np.concatenate([[0], np.cumsum(T)]).astype(np.uint64)
    for world in (1, 2, 3, 4, 8):
        b = hostapi.shard_utterances(fo, world)
        assert b[0] == 0 and b[-1] == 1000 and np.all(np.diff(b.astype(np.int64)) >= 0)
        frames = np.diff(fo[b.astype(np.int64)].astype(np.int64))
        assert frames.sum() == fo[-1]
        assert frames.max() - frames.min() <= 1399          # within one utterance of each other
    # more ranks than utterances: the empty shards are at the end, nothing is lost
    b = hostapi.shard_utterances(np.array([0, 10, 20], np.uint64), 4)
    assert b[0] == 0 and b[-1] == 2 and np.all(np.diff(b.astype(np.int64)) >= 0)
