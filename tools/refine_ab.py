"""A/B of refinement options at the bench shape: stage times with an option toggled.
usage: python tools/refine_ab.py option v0 v1 ..."""
import sys
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

opt = sys.argv[1] if len(sys.argv) > 1 else "tc_civ"
vals = [int(v) for v in sys.argv[2:]] or [0, 1, 0, 1]
utts = 4096
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, utts)
ctx = capi.Context(0)
ctx.set_model_tables(33, tables); ctx.set_features(corp["x"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
ref = None
for v in vals:
    ctx.set_option(opt, v)
    for _ in range(3):
        ctx.align(want_alignment=False, want_costs=False)
    f, r, vit = [], [], []
    for _ in range(5):
        ali, costs = ctx.align()
        st = ctx.stage_ms()
        f.append(st["score_filter"]); r.append(st["score_refine"]); vit.append(st["viterbi"])
    if ref is None:
        ref = (ali.copy(), costs.copy())
    same = np.array_equal(ali["state"], ref[0]["state"]) and np.array_equal(costs, ref[1])
    print("%s=%d filter %.3f refine %.3f (min %.3f) viterbi %.3f  same alignment+costs: %s" % (opt, v, np.mean(f), np.mean(r), min(r), np.mean(vit), same), flush=True)
