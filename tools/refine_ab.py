"""A/B of the refinement kernels on the bench workload: frame-major (tc_refine 1) vs state-major (2).
Prints per-stage device times and checks that both give identical bits."""
import sys
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

utts = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, utts)
ctx = capi.Context(0)
ctx.set_model_tables(33, tables)
ctx.set_features(corp["x"])
ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
res = {}
modes = [int(a) for a in sys.argv[2:]] or [1, 2, 1, 2]
for mode in modes:
    ctx.set_option("tc_refine", mode % 10)
    ctx.set_option("tc_rescan", (mode // 10) % 10)
    ctx.set_option("tc_f16", 0 if mode >= 100 else 1)
    ms = []
    for it in range(6):
        ali, costs = ctx.align()
        if it >= 2:
            ms.append(ctx.stage_ms())
    keys = ms[0].keys()
    print("tc_refine", mode, {k: round(float(np.mean([m[k] for m in ms])), 3) for k in keys}, flush=True)
    res[mode] = (ali.copy(), costs.copy())
ks = list(res)
if len(ks) > 1:
    for k in ks[1:]:
        assert np.array_equal(res[ks[0]][0], res[k][0]) and np.array_equal(res[ks[0]][1], res[k][1])
    print("identical alignments and costs")
