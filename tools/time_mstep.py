#!/usr/bin/env python
"""Wall-clock of the per-iteration model update at the benchmark shape (106 x 256 densities, d = 33, pooled):
host finalize (csrc/host MixtureModel::finalize) and sgmm_set_model (K5 table build + upload), best of 5."""
import json
import sys
import time
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speechrecognition_b200 import capi, hostapi, synth  # noqa: E402

dens = int(sys.argv[1]) if len(sys.argv) > 1 else 256
lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, dens, "mixture")
H = hostapi.HostModel(33, lex.num_states, "mixture", host_only=True)
H.load_tables(tables)
tf = []
for _ in range(5):
    t0 = time.perf_counter(); H.finalize(); tf.append(time.perf_counter() - t0)
ctx = capi.Context(0)
mi, vi = tables["mean_idx"].astype(np.int64), tables["var_idx"].astype(np.int64)
args = (33, tables["mix_offsets"], tables["means"].reshape(-1, 33)[mi], tables["vars_inv"].reshape(-1, 33)[vi],
        tables["norm"][vi], tables["mean_wlog"][mi])
ts = []
for _ in range(5):
    t0 = time.perf_counter(); ctx.set_model(*args); ts.append(time.perf_counter() - t0)
print(json.dumps({"densities": int(tables["mix_offsets"][-1]), "host_finalize_ms": 1e3 * min(tf), "set_model_ms": 1e3 * min(ts),
                  "set_model_ms_all": [1e3 * t for t in ts]}))
