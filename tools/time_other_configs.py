"""Times BASELINE configs[3] (recognition pass) and configs[4] (rwth-shaped stress: 4501 states x 64 densities, d = 48,
1 M frames) on one B200.  Prints one JSON line.  usage: python tools/time_other_configs.py"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speechrecognition_b200 import capi, synth  # noqa: E402

out = {}


def timed(ctx, fn, reps=3):
    fn()
    best = 1e30
    for _ in range(reps):
        ctx.timer_start()
        fn()
        best = min(best, ctx.timer_stop())
    return best


# ---- config 4: recognition over the sietill digit lexicon, the bench model (106 x 256, d = 33), 4096 utterances
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, 4096)
F = int(corp["frame_off"][-1])
ctx = capi.Context(0)
ctx.set_model_tables(33, tables)
ctx.set_features(corp["x"])
ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
ctx.set_lexicon(lex.automata, lex.silence)
ms = timed(ctx, lambda: ctx.recognize(corp["frame_off"], 10.0, 20.0), reps=2)
st = ctx.stage_ms()
hyp, _ = ctx.recognize(corp["frame_off"], 10.0, 20.0)
ref_words = corp["words"]
nref = sum(len(w) for w in ref_words)
# word accuracy on the synthetic corpus (the frames were sampled from this very model): a sanity number, not a metric
same = sum(int(len(h) == len(r) and np.array_equal(h, r)) for h, r in zip(hyp, ref_words))
out["config4_recognition"] = {"utterances": 4096, "frames": F, "ms": ms, "frames_per_s": F / (ms / 1e3),
                              "score_ms": st["score"], "search_ms": st["viterbi"], "d2h_ms": st["d2h"],
                              "sentences_exact": same, "reference_words": nref}
ctx.close()

# ---- config 5: 4501 tied states x 64 densities, d = 48, 2000 utterances x 500 frames = 1 M frames
rng = np.random.default_rng(synth.SEED)
n_states, nd, dim = 4501, 64, 48
tables = synth.random_model(rng, dim, n_states, nd, "mixture")
corp = synth.make_chain_corpus(rng, tables, dim, n_states, 2000, T=500)
F = int(corp["frame_off"][-1])
ctx = capi.Context(0)
ctx.set_model_tables(dim, tables)
ctx.set_features(corp["x"])
ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
ctx.set_tdp(3.0, 0.0, 30.0, 0)
ms_a = timed(ctx, lambda: ctx.align(want_alignment=False, want_costs=False))
st = ctx.stage_ms()
ms_e = timed(ctx, lambda: ctx.accumulate(first_pass=False, want=False))
st2 = ctx.stage_ms()
# densities the gathered pass scores: distinct states per utterance x 64 x frames
dens = 0
for u in range(2000):
    a = corp["automata"][int(corp["aut_off"][u]):int(corp["aut_off"][u + 1])]
    dens += len(np.unique(a)) * nd * int(corp["frame_off"][u + 1] - corp["frame_off"][u])
out["config5_stress"] = {"states": n_states, "densities_per_state": nd, "dim": dim, "frames": F, "utterances": 2000,
                         "scorer": int(ctx.get_counter("scorer")), "align_ms": ms_a, "accumulate_ms": ms_e,
                         "frames_per_s_align_plus_accumulate": F / ((ms_a + ms_e) / 1e3),
                         "score_ms": st["score"], "score_filter_ms": st["score_filter"], "score_refine_ms": st["score_refine"],
                         "viterbi_ms": st["viterbi"], "sort_ms": st2["sort"], "reduce_ms": st2["reduce"],
                         "gathered_tflops_algorithmic": 2.0 * (dim + 1) * dens / (st["score_filter"] / 1e3) / 1e12}
ctx.close()
print(json.dumps(out))
