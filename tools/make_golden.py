"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built from /root/reference).

Run in the build container only:  python tools/make_golden.py
The reference has no golden vectors of its own for this path, so these fixtures are its
outputs on seeded synthetic inputs: a miniature Trainer::train trajectory per variance-pooling mode.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.binding import Reference, TABLES  # noqa: E402
from speechrecognition_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
TDP = (3.0, 0.0, 30.0)
DIM = 9
MIN_OBS = 6


def main():
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(20161107)
    lex = synth.Lexicon()
    gen = synth.random_model(rng, DIM, lex.num_states, 2, "none")
    corp = synth.make_corpus(rng, lex, gen, DIM, 10, t_choices=([70, 95, 130], [0.5, 0.3, 0.2]), words_per_utt=(1, 2))
    x = corp["x"]
    np.savez_compressed(os.path.join(OUT, "corpus.npz"), x=x, frame_off=corp["frame_off"], automata=corp["automata"],
                        aut_off=corp["aut_off"], tdp=np.array(TDP), dim=DIM, min_obs=MIN_OBS)
    for pooling in ("none", "mixture", "global"):
        R = Reference(DIM, lex.num_states, pooling, True, TDP, lex.silence_state)
        rec = {}
        ali = synth.linear_alignment(corp["frame_off"], corp["automata"], corp["aut_off"])
        rec["ali_linear"] = ali
        R.accumulate(ali, x, first_pass=True); R.finalize()
        for k, v in R.tables().items():
            rec["t0_" + k] = v
        rec["am0"] = R.am_score(ali, x)
        R.split(2 * MIN_OBS)
        for k, v in R.tables().items():
            rec["tsplit_" + k] = v
        R.accumulate(ali, x); R.finalize()
        R.eliminate(MIN_OBS)
        R.accumulate(ali, x); R.finalize()
        for k, v in R.tables().items():
            rec["t1_" + k] = v
        st, costs = R.align_batch(x, corp["frame_off"], corp["automata"], corp["aut_off"])
        rec["ali_full"], rec["costs_full"] = st, costs
        f0, f1 = int(corp["frame_off"][2]), int(corp["frame_off"][3])
        aut = corp["automata"][corp["aut_off"][2]:corp["aut_off"][3]]
        for thr, tag in ((40.0, "p40"), (1e300, "pinf")):
            c, s = R.align_pruned(x[f0:f1], aut, thr)
            rec["ali_" + tag], rec["cost_" + tag] = s, c
        sc, am = R.min_score(x[:64])
        rec["scores64"], rec["argmin64"] = sc, am
        R.accumulate(st, x)
        for k, v in R.tables().items():
            rec["t2_" + k] = v          # accumulators after the E-step on the Viterbi alignment (not finalized)
        R.finalize()
        rec["am2"] = R.am_score(st, x)
        path = os.path.join(OUT, "mix_%s.mix" % pooling)
        R.write(path)
        np.savez_compressed(os.path.join(OUT, "trajectory_%s.npz" % pooling), **rec)
        print(pooling, "densities", R.num_densities(), "am", rec["am0"], rec["am2"])


REC_CONFIGS = ((10.0, 20.0), (10.0, 1e30), (2.0, 45.0))   # (word-penalty, am-threshold); defaults first (Recognizer.cpp:31-32)


def recognition():
    """recognize.npz: Recognizer::recognizeSequence_pruned (Recognizer.cpp:103-232) on every golden utterance with the
    final model of the "none" trajectory (mix_none.mix), for three (word-penalty, am-threshold) settings."""
    c = np.load(os.path.join(OUT, "corpus.npz"))
    lex = synth.Lexicon()
    rec = {"configs": np.array(REC_CONFIGS)}
    for ci, (wp, thr) in enumerate(REC_CONFIGS):
        R = Reference(DIM, lex.num_states, "none", True, TDP, lex.silence_state, mix_path=os.path.join(OUT, "mix_none.mix"),
                      extra_config={"am-threshold": float(thr), "word-penalty": float(wp)})
        words, off = [], [0]
        for u in range(len(c["frame_off"]) - 1):
            w = R.recognize(c["x"][int(c["frame_off"][u]):int(c["frame_off"][u + 1])])
            words.extend(int(v) for v in w)
            off.append(len(words))
        rec["words%d" % ci], rec["off%d" % ci] = np.array(words, np.uint32), np.array(off, np.uint32)
        print("recognize", wp, thr, "->", off[-1], "words")
    np.savez_compressed(os.path.join(OUT, "recognize.npz"), **rec)


FEAT_CASES = [  # (n_file, n_first, n_second, deriv_step, normalise, energy_max_norm)
    (12, 12, 1, 3, True, True),      # the bundled digits setup (SignalAnalysis.cpp:53-56, Normalization.bin)
    (12, 12, 1, 3, False, False),
    (9, 5, 5, 2, True, False),
    (16, 16, 0, 1, False, True),
    (33, 0, 0, 3, True, True),
]


def features():
    """features.npz: SignalAnalysis::process_features (SignalAnalysis.cpp:379-399) of the unmodified reference on
    seeded raw frames, one utterance per call, for the parameter sets above."""
    from oracle.binding import reference_process_features
    rng = np.random.default_rng(20161107)
    rec = {"cases": np.array([[int(v) for v in c] for c in FEAT_CASES])}
    for ci, (nf, n1, n2, step, norm, emax) in enumerate(FEAT_CASES):
        lens = [step + 1, 37, 64, 5 + step]
        raw = (rng.standard_normal((sum(lens), nf)) * rng.uniform(0.5, 20.0, nf) + rng.uniform(-30, 30, nf)).astype(np.float32)
        nt = nf + n1 + n2
        mean = rng.uniform(-20, 20, nt) if norm else None
        std = rng.uniform(0.3, 15.0, nt) if norm else None
        off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        out = np.concatenate([reference_process_features(raw[int(off[u]):int(off[u + 1])], n1, n2, step, mean, std, emax)
                              for u in range(len(lens))])
        rec["raw%d" % ci], rec["off%d" % ci], rec["out%d" % ci] = raw, off, out
        if norm:
            rec["mean%d" % ci], rec["std%d" % ci] = mean, std
    np.savez_compressed(os.path.join(OUT, "features.npz"), **rec)
    print("features:", len(FEAT_CASES), "cases")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "recognize":
        recognition()
    elif len(sys.argv) > 1 and sys.argv[1] == "features":
        features()
    else:
        main()
        recognition()
        features()
