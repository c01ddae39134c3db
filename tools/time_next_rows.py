"""Times the SURVEY §8(f) rows at the bench shape on one B200 (106 states x 256 densities, d = 33, 4096 utterances):
sum-score alignment pass, posterior-weighted accumulate, device feature processing.  Prints one JSON line.
usage: python tools/time_next_rows.py [--utts 4096]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speechrecognition_b200 import capi, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utts", type=int, default=4096)
    ap.add_argument("--dens", type=int, default=256)
    ap.add_argument("--dim", type=int, default=33)
    a = ap.parse_args()
    lex = synth.Lexicon()
    tables = synth.random_model(np.random.default_rng(synth.SEED), a.dim, lex.num_states, a.dens, "mixture")
    corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, a.dim, a.utts)
    F = int(corp["frame_off"][-1])
    ctx = capi.Context(0)
    ctx.set_model_tables(a.dim, tables)
    ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
    ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
    out = {"frames": F, "utterances": a.utts, "densities_per_state": a.dens, "dim": a.dim}

    def timed(fn, reps=3):
        fn()
        best = 1e30
        for _ in range(reps):
            ctx.timer_start()
            fn()
            best = min(best, ctx.timer_stop())
        return best

    out["align_max_approx_ms"] = timed(lambda: ctx.align(want_alignment=False, want_costs=False))
    out["accumulate_max_approx_ms"] = timed(lambda: ctx.accumulate(first_pass=False, want=False))
    ctx.set_option("max_approx", 0)
    out["align_sum_scores_ms"] = timed(lambda: ctx.align(want_alignment=False, want_costs=False), reps=2)
    st = ctx.stage_ms()
    out["align_sum_scores_score_ms"] = st["score"]
    t0 = time.time()
    out["accumulate_soft_ms"] = timed(lambda: ctx.accumulate_soft(), reps=2)
    st = ctx.stage_ms()
    out["accumulate_soft_stages"] = {k: st[k] for k in ("sort", "reduce", "rescore", "d2h")}
    out["am_score_sum_scores_ms"] = timed(lambda: ctx.am_score_sum(), reps=2)
    ctx.set_option("max_approx", 1)
    # feature processing: 12 file features -> 25, same utterance lengths
    raw = np.random.default_rng(1).standard_normal((F, 12)).astype(np.float32)
    mean, std = np.zeros(25), np.ones(25)
    raw_pin = capi.pinned_empty(raw.shape, np.float32)
    raw_pin[...] = raw
    out["process_features_ms"] = timed(lambda: ctx.process_features(raw_pin, corp["frame_off"], 12, 1, 3, mean, std, True, want=False))
    st = ctx.stage_ms()
    out["process_features_h2d_ms"] = st["h2d"]
    out["process_features_bytes"] = int(F * (12 + 25) * 4)
    print(json.dumps(out))
    ctx.close()


if __name__ == "__main__":
    main()
