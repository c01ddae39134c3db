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
