"""A/B of the packed K remainder of the fp16 operand form (option tc_rem, read by sgmm_set_model) at the bench shape.
usage: python tools/rem_ab.py [utts]"""
import sys
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

utts = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, utts)
ctx = capi.Context(0)
ctx.set_features(corp["x"])
ref = None
for rem in (0, 1, 0, 1):
    ctx.set_option("tc_rem", rem)
    ctx.set_model_tables(33, tables)
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
    for _ in range(3):
        ctx.align(want_alignment=False, want_costs=False)
    ctx.get_counter("fallbacks")
    f, r = [], []
    for _ in range(5):
        ali, costs = ctx.align()
        st = ctx.stage_ms()
        f.append(st["score_filter"]); r.append(st["score_refine"])
    fb = ctx.get_counter("fallbacks") / 5
    if ref is None:
        ref = (ali.copy(), costs.copy())
    same = np.array_equal(ali["state"], ref[0]["state"]) and np.array_equal(costs, ref[1])
    print("tc_rem=%d (in use: %d terms) filter %.3f ms (min %.3f)  refine %.3f  uncertain cells/step %.0f  same alignment+costs: %s"
          % (rem, ctx.get_counter("tc_rem"), float(np.mean(f)), min(f), float(np.mean(r)), fb, same), flush=True)
