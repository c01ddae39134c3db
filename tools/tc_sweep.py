"""Time the K1 stages (filter / refine) of one alignment pass for the GEMM-kernel options.  GPU box only."""
import sys
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

utts = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(1), lex, tables, 33, utts)
ctx = capi.Context(0)
ctx.set_model_tables(33, tables); ctx.set_features(corp["x"])
ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
for halves, dbg in ((2, 0),):
    ctx.set_option("tc_halves", halves); ctx.set_option("tc_debug", dbg)
    for _ in range(3):
        ctx.align(want_alignment=False, want_costs=False)
    ms = [ctx.stage_ms() for _ in range(3) if ctx.align(want_alignment=False, want_costs=False) or True]
    print("halves", halves, "debug", dbg, "frames", len(corp["x"]), {k: round(float(np.mean([m[k] for m in ms])), 3) for k in ("score_filter", "score_refine", "viterbi")})
