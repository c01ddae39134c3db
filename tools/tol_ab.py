"""The tolerance mode (option "tolerance", experiment, default off) against the exact mode at the bench shape:
step time (align + accumulate + mstep) and the mismatch list — utterances whose alignment differs, frames that differ,
relative path-cost differences, count / sum differences of the accumulated statistics.
usage: python tools/tol_ab.py [utts] [seed offset]"""
import sys, json
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

utts = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
so = int(sys.argv[2]) if len(sys.argv) > 2 else 0
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED + 100 * so), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1 + 100 * so), lex, tables, 33, utts)
ctx = capi.Context(0)
ctx.set_option("vit_mw", 0)
ctx.set_model_tables(33, tables); ctx.set_features(corp["x"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
fo = corp["frame_off"]
res = {}
for mode in (0, 1):
    ctx.set_option("tolerance", mode)
    for _ in range(3):
        ctx.align(want_alignment=False, want_costs=False); ctx.accumulate(first_pass=False)
    tt = []
    for _ in range(5):
        ali, costs = ctx.align()
        st = dict(ctx.stage_ms())
        cnt, sx, sxx, ssum = ctx.accumulate(first_pass=False)
        st2 = ctx.stage_ms()
        tt.append((st["score_filter"], st["score_refine"], st["viterbi"], st2["rescore"], st2["sort"] + st2["reduce"]))
    m = np.mean(tt, axis=0)
    res[mode] = dict(ali=ali["state"].copy(), costs=costs.copy(), cnt=cnt.copy(), sx=sx.copy(), ssum=ssum,
                     ms=dict(filter=m[0], refine_or_finish=m[1], viterbi=m[2], rescore=m[3], accumulate=m[4], total=float(m.sum())))
    print("tolerance=%d  " % mode + "  ".join("%s %.3f" % kv for kv in res[mode]["ms"].items()), flush=True)
a, b = res[0], res[1]
diff_frames = a["ali"] != b["ali"]
utt_diff = [u for u in range(utts) if diff_frames[fo[u]:fo[u + 1]].any()]
rel = np.abs(b["costs"] - a["costs"]) / np.abs(a["costs"])
out = {
    "seed_offset": so, "utterances": utts, "frames": int(fo[-1]),
    "ms_exact": a["ms"], "ms_tolerance": b["ms"],
    "utterances_with_different_alignment": len(utt_diff), "frames_with_different_state": int(diff_frames.sum()),
    "path_cost_rel_diff_max": float(rel.max()), "path_cost_rel_diff_median": float(np.median(rel)),
    "count_abs_diff_total": float(np.abs(a["cnt"] - b["cnt"]).sum()), "densities_with_different_count": int((a["cnt"] != b["cnt"]).sum()),
    "sum_x_rel_diff_max": float(np.max(np.abs(a["sx"] - b["sx"]) / np.maximum(np.abs(a["sx"]), 1e-300))),
    "am_score_sum_rel_diff": float(abs(a["ssum"] - b["ssum"]) / abs(a["ssum"])),
}
# the mismatching utterances: where, how many frames, and how far the two path costs are apart
out["mismatches"] = [{"utt": int(u), "frames": int(fo[u + 1] - fo[u]), "frames_different": int(diff_frames[fo[u]:fo[u + 1]].sum()),
                      "cost_exact": float(a["costs"][u]), "cost_tolerance": float(b["costs"][u])} for u in utt_diff[:50]]
print(json.dumps(out))
