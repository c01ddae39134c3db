"""A/B of the multi-warp Viterbi kernel (option vit_mw, read by sgmm_set_corpus) at the bench shape: stage time.
usage: python tools/vit_ab.py [utts]"""
import sys
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

utts = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, utts)
ctx = capi.Context(0)
ctx.set_model_tables(33, tables); ctx.set_features(corp["x"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
ref = None
for mw in [int(v) for v in sys.argv[2:]] or (0, 1, 0, 1):
    ctx.set_option("vit_mw", mw)
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
    for _ in range(3):
        ctx.align(want_alignment=False, want_costs=False)
    v = []
    for _ in range(5):
        ali, costs = ctx.align()
        v.append(ctx.stage_ms()["viterbi"])
    if ref is None:
        ref = (ali.copy(), costs.copy())
    same = np.array_equal(ali["state"], ref[0]["state"]) and np.array_equal(costs, ref[1])
    print("vit_mw=%d viterbi %.3f ms (min %.3f)  same alignment+costs: %s" % (mw, float(np.mean(v)), min(v), same), flush=True)
