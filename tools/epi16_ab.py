"""A/B of the GEMM epilogue organisation (option tc_epi16) at the bench shape: filter time, uncertain cells.
usage: python tools/epi16_ab.py [utts]"""
import sys
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

utts = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, utts)
ctx = capi.Context(0)
ctx.set_model_tables(33, tables); ctx.set_features(corp["x"])
ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
ref = None
for epi16 in (0, 1, 0, 1):
    ctx.set_option("tc_epi16", epi16)
    for _ in range(3):
        ctx.align(want_alignment=False, want_costs=False)
    ctx.get_counter("fallbacks")
    f = []
    for _ in range(5):
        ali, costs = ctx.align()
        f.append(ctx.stage_ms()["score_filter"])
    fb = ctx.get_counter("fallbacks") / 5
    if ref is None:
        ref = (ali.copy(), costs.copy())
    same = np.array_equal(ali["state"], ref[0]["state"]) and np.array_equal(costs, ref[1])
    print("tc_epi16=%d filter %.3f ms (min %.3f)  uncertain cells/step %.0f  same alignment+costs: %s"
          % (epi16, float(np.mean(f)), min(f), fb, same), flush=True)
