"""Sum scores (option max_approx = 0) of the gathered alignment pass: the tensor path (option sum_tc = 1: the filter's
candidate masks + the state-major kernel in SUM form) against k_sum_pruned (sum_tc = 0) at the bench shape, and both
against the oracle's sum scores on a sample.  usage: python tools/sum_ab.py [utts]"""
import sys, json
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth
from oracle.binding import Oracle

utts = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, utts)
ctx = capi.Context(0)
ctx.set_model_tables(33, tables); ctx.set_features(corp["x"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
ctx.set_option("max_approx", 0)
res = {}
for tc in (1, 0):
    ctx.set_option("sum_tc", tc)
    ctx.align(want_alignment=False, want_costs=False)
    ms = []
    for _ in range(2):
        ctx.timer_start(); ali, costs = ctx.align(); ms.append(ctx.timer_stop())
    st = ctx.stage_ms()
    res[tc] = (min(ms), ali["state"].copy(), costs.copy(), st["score"], st["score_filter"], st["score_refine"])
    print("sum_tc=%d align %.2f ms (score %.2f: filter %.2f, refine %.2f)" % (tc, min(ms), st["score"], st["score_filter"], st["score_refine"]), flush=True)
same = np.array_equal(res[0][1], res[1][1])
rel = np.abs(res[0][2] - res[1][2]) / np.abs(res[0][2])
print("alignments equal:", same, " path costs rel diff max %.3e" % rel.max())
# oracle on the first 24 utterances
O = Oracle(33, lex.num_states, "mixture", False, (3.0, 0.0, 30.0), lex.silence_state)
O.load_tables(tables)
n = 24
fo, ao = corp["frame_off"], corp["aut_off"]
ws, wc = O.align_batch(corp["x"][:fo[n]], fo[:n + 1], corp["automata"][:ao[n]], ao[:n + 1], threads=8)
for tc in (1, 0):
    print("sum_tc=%d vs oracle (first %d utterances): states equal %s, costs rel diff max %.3e" % (
        tc, n, np.array_equal(res[tc][1][:fo[n]].astype(np.uint32), ws), np.max(np.abs(res[tc][2][:n] - wc) / np.abs(wc))))
