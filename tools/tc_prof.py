"""(Needs a library built with the measurement paths: make -C speechrecognition_b200/csrc clean all EXTRA_NVFLAGS="-DSGMM_TC_PROF=1";
they are compiled out by default because dead code in the filter kernel costs it 3 %.)
Cycle profile of k_tc_gemm (option tc_debug & 32): where the MMA-issuing warp and the epilogue warps wait.
usage: python tools/tc_prof.py [utts]"""
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
ctx.set_option("scorer", 3)
modes = [int(v) for v in sys.argv[2:]] or [32, 33, 160, 161]   # +128: two-pass epilogue, +1: no epilogue math
for dbg, f16 in [(m, 1) for m in modes]:
    ctx.set_option("tc_debug", dbg % 10000)
    ctx.set_option("tc_epi16", dbg // 10000)
    ctx.set_option("tc_f16", f16)
    for _ in range(3):
        ctx.align(want_alignment=False, want_costs=False)
    [ctx.get_counter("tc_prof%d" % i) for i in range(1, 13)]
    ctx.align(want_alignment=False, want_costs=False)
    st = ctx.stage_ms()
    c = [0] + [ctx.get_counter("tc_prof%d" % i) for i in range(1, 13)]
    n = max(1, c[7])
    print("f16=%d " % f16, end="")
    print("debug=%5d filter %.3f ms | CTAs %d  MMA-warp life %.0f cyc (%.2f GHz)  wait A %.0f  wait B %.0f  wait epilogue %.0f  "
          "MMAs %d -> %.1f cyc/MMA busy | epilogue: wait %.0f/%.0f  life %.0f/%.0f (groups %d)"
          % (dbg, st["score_filter"], c[7], c[1] / n, c[1] / max(1, c[2]), c[3] / n, c[4] / n, c[5] / n, c[6] / n,
             (c[1] - c[3] - c[4] - c[5]) / max(1, c[6]), 2 * c[8] / max(1, c[12]), 2 * c[9] / max(1, c[12]),
             2 * c[10] / max(1, c[12]), 2 * c[11] / max(1, c[12]), c[12]))
