"""(Needs a library built with the measurement paths: make -C speechrecognition_b200/csrc clean all EXTRA_NVFLAGS="-DSGMM_TC_TWOPASS_REF=1 -DSGMM_TC_PROF=1";
they are compiled out by default because dead code in the filter kernel costs it 3 %.)
A/B of the GEMM epilogue's TMEM traffic at the bench shape: tc_debug 0 = one pass, 64-column sub-parts (default),
256 = one pass, 32-column sub-parts, 128 = the two-pass scan.  Filter time, uncertain cells, identical results.
usage: python tools/epi_onepass_ab.py [utts] [modes...]"""
import sys
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

utts = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
# mode = tc_debug + 10000 * tc_epi16
modes = [int(v) for v in sys.argv[2:]] or [128, 0, 10128, 10000, 128, 0, 10128, 10000]
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, utts)
ctx = capi.Context(0)
ctx.set_model_tables(33, tables); ctx.set_features(corp["x"])
ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
ref = None
for mode in modes:
    ctx.set_option("tc_debug", mode % 10000)
    ctx.set_option("tc_epi16", mode // 10000)
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
    print("mode=%5d filter %.3f ms (min %.3f)  refine %.3f  uncertain cells/step %.0f  same alignment+costs: %s"
          % (mode, float(np.mean(f)), min(f), float(np.mean(r)), fb, same), flush=True)
