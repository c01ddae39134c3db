"""Stress shape (4501 states x 64 densities, d = 48): stage times of the alignment pass under scorer options (gpurun).
usage: stress_ab.py [n_utts] -- option sets are listed below; the alignment of every set is compared with the first."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from speechrecognition_b200 import capi, synth

n_utts = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
n_states, nd, dim = 4501, 64, 48
rng = np.random.default_rng(synth.SEED)
tables = synth.random_model(rng, dim, n_states, nd, "mixture")
corp = synth.make_chain_corpus(np.random.default_rng(synth.SEED + 201), tables, dim, n_states, n_utts, T=500)
ref = None
for name, opts in (("default", {}), ("wide_f16", {"tc_wide": 1}), ("wide_tf32", {"tc_wide": 1, "tc_f16": 0})):
    ctx = capi.Context(0)
    for k, v in opts.items():
        ctx.set_option(k, v)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])
    ctx.set_tdp(3.0, 0.0, 30.0, 0)
    for _ in range(2):
        ali, costs = ctx.align()
    st = []
    for _ in range(3):
        ctx.align(want_alignment=False, want_costs=False)
        st.append(ctx.stage_ms())
    same = None
    if ref is None:
        ref = (ali["state"].copy(), np.array(costs))
    else:
        same = bool(np.array_equal(ali["state"], ref[0]) and np.array_equal(costs, ref[1]))
    print(json.dumps({"case": name, "form": ctx.get_counter("tc_form"), "ksteps": ctx.get_counter("tc_ksteps") if hasattr(ctx, "get_counter") else None,
                      "filter_ms": float(np.mean([s["score_filter"] for s in st])),
                      "refine_ms": float(np.mean([s["score_refine"] for s in st])),
                      "rescans": ctx.get_counter("fallbacks"), "same_as_default": same}), flush=True)
    ctx.close()
