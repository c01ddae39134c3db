"""Shared builders for the parity tests: seeded synthetic corpora and oracle-trained models."""
import numpy as np

from oracle.binding import Oracle
from speechrecognition_b200 import synth

TDP = (3.0, 0.0, 30.0)


def trained_oracle(dim=25, n_utts=24, splits=2, pooling="none", seed=7, min_obs=15, dens_gen=2):
    """A small sietill-shaped corpus and an Oracle model trained on it the way Trainer::train does
    (Training.cpp:114-226: first-pass accumulate, then split / eliminate / realign rounds), so the model
    carries everything the reference produces: holes after eliminate, un-halved weights after split."""
    rng = np.random.default_rng(seed)
    lex = synth.Lexicon()
    gen = synth.random_model(rng, dim, lex.num_states, dens_gen, "none")
    corp = synth.make_corpus(rng, lex, gen, dim, n_utts)
    O = Oracle(dim, lex.num_states, pooling, True, TDP, lex.silence_state)
    ali = synth.linear_alignment(corp["frame_off"], corp["automata"], corp["aut_off"])
    x = corp["x"]
    O.accumulate(ali, x, first_pass=True)
    O.finalize()
    for i in range(splits + 1):
        if i > 0:
            O.split(2 * min_obs)
            O.accumulate(ali, x); O.finalize()
            O.eliminate(min_obs)
            O.accumulate(ali, x); O.finalize()
        ali, _ = O.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"], threads=8)
        O.accumulate(ali, x); O.finalize()
    return O, corp, ali


def gpu_context(O, corp, dim):
    from speechrecognition_b200 import capi
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, O.tables())
    ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
    ctx.set_tdp(*O.tdp, O.silence)
    return ctx


def packed_stats(tables, dim):
    """Reference accumulators (mean_acc/var_acc/wacc indexed by mean_idx/var_idx) -> per-density packed
    (count, sum_x) in state-major order, as sgmm_accumulate reports them."""
    mi = tables["mean_idx"].astype(np.int64)
    return tables["mean_wacc"][mi], tables["mean_acc"].reshape(-1, dim)[mi]
