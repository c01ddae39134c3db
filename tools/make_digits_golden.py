#!/usr/bin/env python
"""BASELINE configs[0] on the REAL bundled data: pack a subset of the reference's German-digits corpus and mint, with the
UNMODIFIED reference binary (oracle/_ref/sietill_ref), what one training pass + recognition produce on it.

Run in the build container only (needs /root/reference):

    python tools/make_digits_golden.py                       # the committed fixture: 256 segments, every 51st
    python tools/make_digits_golden.py --first 1000 --local  # tests/golden/_local/digits_first1000.npz (git-ignored,
                                                             # travels to the GPU box with the snapshot)
    python tools/make_digits_golden.py --all --time-only     # whole corpus, CPU-reference timing only

Inputs (SURVEY.md §8(d) "cfg 1 inputs (real)"): src/sietill/corpora/corpus_train.json x data/new_features/*.mm2 (raw
12-dim f32 cepstra) x src/sietill/Normalization.bin; configuration = src/sietill/config/train_demo_corpus.config with
feature-path -> data/new_features/, num-splits 0, alignment-pruning false, pooling "none" (single Gaussian per state).
With num-splits 0, Trainer::train (Training.cpp:44-235) is: linear segmentation -> first-pass accumulate + finalize ->
lin.mix -> full-DP realignment -> 0-0.dump -> accumulate + finalize -> AM score -> 0.mix, i.e. "Viterbi alignment + 1 EM
iteration".  Recognizer::recognize (Recognizer.cpp:38-92) then decodes the same segments with 0.mix.

The pack holds what travels: the raw frames exactly as they lie in the .mm2 files, utterance offsets, the word ids of
the transcriptions, the normalization blobs, and the reference's outputs (alignments, .mix files, AM scores, recognised
word sequences, WER / SER counts, Timer buckets).
"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speechrecognition_b200 import mixio  # noqa: E402

REF = "/root/reference"
SIETILL = os.path.join(REF, "src", "sietill")
FEATURES = os.path.join(REF, "data", "new_features") + "/"
BIN = os.path.join(ROOT, "oracle", "_ref", "sietill_ref")
# Lexicon.cpp:70-85 (build_sietill_lexicon): word ids in insertion order
WORDS = ["[silence]", "eins", "zwei", "drei", "vier", "fuenf", "sechs", "sieben", "acht", "neun", "null", "zwo"]

TRAIN_CONFIG = {
    "verbosity": "noLog", "energy-max-norm": True, "pooling": "none", "max-approx": True,
    "write-linear-segmentation": False, "realign": True, "approx-linear-segmentation": True,
    "n-features-file": 12, "n-features-first": 12, "n-features-second": 1, "deriv-step": 3,
    "sample-rate": 8000, "window-shift": 10, "window-size": 25, "dft-length": 1024, "n-mel-features": 15,
    "tdp-loop": 20.0, "tdp-forward": 0.0, "tdp-skip": 20.0,
    "min-obs": 1, "num-splits": 0, "num-aligns": 1, "num-estimates": 1, "num-max-aligns": 1,
    "alignment-pruning": False, "pruning-threshold": 120.0,
    "am-threshold": 100.0, "word-penalty": 20.0, "pruned-search": True,
}


def lexicon_order():
    """word -> id as build_sietill_lexicon assigns them (checked against the source text)."""
    src = open(os.path.join(SIETILL, "Lexicon.cpp")).read()
    names = re.findall(r'add_word\(\s*"([^"]+)"', src)
    return {w: i for i, w in enumerate(names)}, names


def run(cmd, cwd, env=None):
    t0 = time.perf_counter()
    p = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True, env=env)
    if p.returncode != 0:
        sys.stderr.write(p.stderr[-4000:])
        raise SystemExit("command failed: " + " ".join(cmd))
    return p.stderr, time.perf_counter() - t0


def timers(stderr):
    out = {}
    for key, pat in (("linear_alignment", r"Linear alignment took ([0-9.e+-]+)"), ("estimate", r"Estimation\s+took ([0-9.e+-]+)"),
                     ("align", r"Alignment\s+took ([0-9.e+-]+)"), ("io", r"IO\s+took ([0-9.e+-]+)"),
                     ("score", r"Score comp\. took ([0-9.e+-]+)"), ("training", r"Training\s+took ([0-9.e+-]+)")):
        m = re.search(pat, stderr)
        if m:
            out[key] = float(m.group(1))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--first", type=int, default=0, help="the first N segments")
    ap.add_argument("--stride", type=int, default=51, help="otherwise every K-th segment ...")
    ap.add_argument("--count", type=int, default=256, help="... up to this many")
    ap.add_argument("--all", action="store_true")
    ap.add_argument("--local", action="store_true", help="write under tests/golden/_local (git-ignored)")
    ap.add_argument("--time-only", action="store_true")
    # Recognizer::recognize adds the per-segment error counts into shared totals inside its `omp parallel for` without
    # synchronisation (Recognizer.cpp:46-61: acc += ed; ref_total += ...; sentence_errors++): with several threads the
    # printed WER / SER totals lose updates and vary from run to run (13389/18815/144 vs 13390/18818/144 on the whole
    # corpus, two runs of the same binary).  The golden totals are therefore minted with ONE thread.
    ap.add_argument("--threads", type=int, default=1, help="OMP threads of the recognition run (0: all; >1 races)")
    a = ap.parse_args()

    segs = json.load(open(os.path.join(SIETILL, "corpora", "corpus_train.json")))["segments"]
    if a.all:
        pick, tag = segs, "all"
    elif a.first:
        pick, tag = segs[:a.first], "first%d" % a.first
    else:
        pick, tag = segs[::a.stride][:a.count], "stride%d_%d" % (a.stride, a.count)
    wid, names = lexicon_order()
    assert names == WORDS, names

    tmp = tempfile.mkdtemp(prefix="digits_")
    json.dump({"segments": pick}, open(os.path.join(tmp, "corpus.json"), "w"))
    cfg = dict(TRAIN_CONFIG)
    cfg.update({"action": "train", "corpus": os.path.join(tmp, "corpus.json"), "feature-path": FEATURES,
                "normalization-path": os.path.join(SIETILL, "Normalization.bin"), "mixture-path": tmp + "/",
                "alignment-path": tmp + "/", "training-stats-path": os.path.join(tmp, "stats.data"),
                "max-recognition-runs": len(pick)})
    json.dump(cfg, open(os.path.join(tmp, "train.json"), "w"))
    err, wall = run([BIN, os.path.join(tmp, "train.json"), "train"], tmp)
    tt = timers(err)
    print("train: %d segments, wall %.1f s, buckets %s" % (len(pick), wall, tt))
    cfg.update({"action": "recognize", "load-mixtures-from": os.path.join(tmp, "0.mix")})
    json.dump(cfg, open(os.path.join(tmp, "rec.json"), "w"))
    env = dict(os.environ)
    if a.threads:
        env["OMP_NUM_THREADS"] = str(a.threads)
    rerr, rwall = run([BIN, os.path.join(tmp, "rec.json"), "recognize"], tmp, env)
    m = re.search(r"^WER:\s*([0-9.]+)% \(S/I/D\) (\d+)/(\d+)/(\d+)\s*$", rerr, re.M)
    ser = float(re.search(r"^SER:\s*([0-9.]+)%", rerr, re.M).group(1))
    wer, sub, ins, dele = float(m.group(1)), int(m.group(2)), int(m.group(3)), int(m.group(4))
    rec_time = float(re.search(r"^Time: ([0-9.e+-]+) seconds", rerr, re.M).group(1))
    hyp = {}
    for line in rerr.splitlines():
        mm = re.match(r"^(\d+)/\d+ WER:.*?\| (.*?)\| (.*)$", line)
        if mm:
            hyp[int(mm.group(1)) - 1] = ([int(v) for v in mm.group(2).split()], [int(v) for v in mm.group(3).split()])
    assert len(hyp) == len(pick), (len(hyp), len(pick))
    print("recognize: WER %.4f%% (S/I/D %d/%d/%d) SER %.4f%%, search %.1f s, wall %.1f s" % (wer, sub, ins, dele, ser, rec_time, rwall))

    ali, max_aligns = mixio.read_alignment(os.path.join(tmp, "0-0.dump"))
    frames = len(ali)
    stats = np.loadtxt(os.path.join(tmp, "stats.data")).reshape(-1, 4)
    summary = {"tag": tag, "segments": len(pick), "frames": int(frames), "cores_recognize": a.threads or os.cpu_count(),
               "train_wall_s": wall, "train_buckets_s": tt, "recognize_search_s": rec_time, "recognize_wall_s": rwall,
               "wer_percent": wer, "ser_percent": ser, "sid": [sub, ins, dele], "am_scores": stats[:, 3].tolist(),
               "frames_per_s_align_1thread": frames / tt["align"] if tt.get("align") else None,
               "frames_per_s_train_1thread": frames / tt["training"] if tt.get("training") else None}
    print(json.dumps(summary))
    if a.time_only:
        return

    # ---- the pack: raw frames exactly as Corpus::read reads them (Corpus.cpp:95-101)
    raws, fo, words, wo = [], [0], [], [0]
    for s in pick:
        r = np.fromfile(FEATURES + s["name"] + ".mm2", dtype=np.float32).reshape(-1, 12)
        raws.append(r)
        fo.append(fo[-1] + len(r))
        words.extend(wid[w] for w in s["orth"].split())
        wo.append(len(words))
    raw = np.concatenate(raws)
    assert len(raw) == frames, (len(raw), frames)
    norm = np.fromfile(os.path.join(SIETILL, "Normalization.bin"), dtype=np.float64)
    assert norm.size == 50
    hyp_words, hyp_off = [], [0]
    for u in range(len(pick)):
        h, r = hyp[u]
        assert r == words[wo[u]:wo[u + 1]], (u, r)
        hyp_words.extend(h)
        hyp_off.append(len(hyp_words))
    # the reference's initial segmentation, through the harness on reference-processed features
    from oracle.binding import Reference, reference_process_features
    from speechrecognition_b200 import synth
    lex = synth.Lexicon()
    R = Reference(25, lex.num_states, "none", True, (20.0, 0.0, 20.0), lex.silence_state)
    lin = np.zeros(frames, np.uint16)
    xsum = np.zeros(len(pick), np.float64)
    xs = []
    for u in range(len(pick)):
        x = reference_process_features(raws[u], 12, 1, 3, norm[:25], norm[25:], True)
        aut = lex.segment_automaton(words[wo[u]:wo[u + 1]])
        st, _ = R.linear_segmentation(x, aut, approx=True)
        lin[fo[u]:fo[u + 1]] = st
        xsum[u] = float(x.astype(np.float64).sum())
        xs.append(x)
    xall = np.concatenate(xs)
    # the stats file holds 6 significant digits: the same AM scores (Training.cpp:585-612) at full precision, from the
    # models the reference wrote (read() re-runs finalize on the stored accumulators, Mixtures.cpp:748-878)
    am_full = []
    for mix, st in (("lin.mix", lin), ("0.mix", ali["state"])):
        Rm = Reference(25, lex.num_states, "none", True, (20.0, 0.0, 20.0), lex.silence_state, mix_path=os.path.join(tmp, mix))
        am_full.append(Rm.am_score(st.astype(np.uint32), xall))
    assert np.allclose(am_full, stats[:, 3], rtol=1e-5), (am_full, stats[:, 3])
    out_dir = os.path.join(ROOT, "tests", "golden", "_local" if a.local else "")
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, "digits_%s.npz" % tag)
    np.savez_compressed(
        path, raw=raw, frame_off=np.array(fo, np.uint64), words=np.array(words, np.uint32), word_off=np.array(wo, np.uint32),
        norm_mean=norm[:25], norm_std=norm[25:], names=np.array([s["name"] for s in pick]),
        config=json.dumps(TRAIN_CONFIG), ali_linear=lin, ali_0_0=ali["state"].astype(np.uint16), stats=stats,
        lin_mix=np.fromfile(os.path.join(tmp, "lin.mix"), dtype=np.uint8), mix0=np.fromfile(os.path.join(tmp, "0.mix"), dtype=np.uint8),
        hyp_words=np.array(hyp_words, np.uint32), hyp_off=np.array(hyp_off, np.uint32),
        wer=np.array([wer, ser, sub, ins, dele, sub + ins + dele, len(words)], np.float64), feature_sums=xsum,
        am_scores=np.array(am_full, np.float64),
        summary=json.dumps(summary))
    print("wrote", path, "%.1f MB" % (os.path.getsize(path) / 1e6))


if __name__ == "__main__":
    main()
