"""A/B of GEMM launch options at the bench shape: filter time per option setting.
usage: python tools/gemm_ab.py key=value[,key=value] ..."""
import sys
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, 4096)
base = {"tc_halves": 0, "tc_epi16": 0, "tc_f16": -1, "tc_pack": 1}
for spec in ["" ] + sys.argv[1:]:
    opts = dict(base)
    for kv in filter(None, spec.split(",")):
        k, v = kv.split("=")
        opts[k] = int(v)
    ctx = capi.Context(0)
    for k, v in opts.items():
        ctx.set_option(k, v)
    ctx.set_model_tables(33, tables); ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
    for _ in range(3):
        ctx.align(want_alignment=False, want_costs=False)
    f, tot = [], []
    for _ in range(5):
        ctx.align(want_alignment=False, want_costs=False)
        st = ctx.stage_ms()
        f.append(st["score_filter"]); tot.append(st["score"])
    print("%-32s filter %.3f ms  K1 stage %.3f ms" % (spec or "(default)", float(np.mean(f)), float(np.mean(tot))), flush=True)
    ctx.close()
