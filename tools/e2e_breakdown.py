"""Host-side wall-clock split of the end-to-end step (set_features -> align -> accumulate) for upload options."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from speechrecognition_b200 import capi, synth

utts = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
corp = synth.make_corpus(np.random.default_rng(synth.SEED + 1), lex, tables, 33, utts)
F, U = len(corp["x"]), utts
ctx = capi.Context(0)
ctx.set_model_tables(33, tables)
ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"]); ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
x_pin = capi.pinned_empty(corp["x"].shape, np.float32); x_pin[...] = corp["x"]
ali_pin = capi.pinned_empty((F,), capi.ALI_DTYPE); cost_pin = capi.pinned_empty((U,), np.float64)
D = int(tables["mix_offsets"][-1])
stat_pin = (capi.pinned_empty((D,), np.float64), capi.pinned_empty((D, 33), np.float64), capi.pinned_empty((D, 33), np.float64))
for mode in ("sync", 1, 4, 8, 10, 12, 16):
    ctx.set_option("async_upload", 0 if mode == "sync" else 1)
    if mode != "sync":
        ctx.set_option("upload_chunks", mode)
    ts = []
    for it in range(6):
        t0 = time.perf_counter(); ctx.set_features(x_pin)
        t1 = time.perf_counter(); ctx.align(out=ali_pin, out_costs=cost_pin)
        t2 = time.perf_counter(); ctx.accumulate(first_pass=False, out=stat_pin)
        t3 = time.perf_counter()
        if it >= 2:
            ts.append((t1 - t0, t2 - t1, t3 - t2, t3 - t0))
    m = np.mean(ts, axis=0) * 1e3
    print("upload %-4s: set_features %.2f  align %.2f  accumulate %.2f  total %.2f ms" % (mode, *m))
