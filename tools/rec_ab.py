"""Search-kernel timing by utterance-length mix (gpurun): separates the issue-bound start of k_recognize from its tail."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from speechrecognition_b200 import capi, synth

lex = synth.Lexicon()
tables = synth.random_model(np.random.default_rng(synth.SEED), 33, lex.num_states, 256, "mixture")
for name, U, tc in (("all279", 4096, ([279], [1.0])), ("all1399_592", 592, ([1399], [1.0])), ("all1399_148", 148, ([1399], [1.0])),
                    ("bench", 4096, synth.SIETILL_T)):
    if len(sys.argv) > 1 and name not in sys.argv[1:]:
        continue
    corp = synth.make_corpus(np.random.default_rng(synth.SEED + 101), lex, tables, 33, U, t_choices=tc)
    ctx = capi.Context(0)
    ctx.set_model_tables(33, tables)
    ctx.set_features(corp["x"])
    ctx.set_tdp(3.0, 0.0, 30.0, lex.silence_state)
    ctx.set_lexicon(lex.automata, lex.silence)
    ctx.set_option("rec_reg", int(os.environ.get("REC_REG", "1")))
    for _ in range(2):
        ctx.recognize(corp["frame_off"], 10.0, 20.0)
    v = []
    for _ in range(3):
        ctx.recognize(corp["frame_off"], 10.0, 20.0)
        v.append(ctx.stage_ms()["viterbi"])
    F = int(corp["frame_off"][-1])
    print(json.dumps({"case": name, "utts": U, "frames": F, "search_ms": float(np.mean(v))}), flush=True)
    ctx.close()
