"""BASELINE configs[0] — the reference's bundled German-digits data (real cepstra, not synthetic): raw .mm2 frames ->
load-time feature processing -> linear segmentation -> first-pass estimate -> full-DP Viterbi realignment -> one EM
iteration -> recognition with the trained model.  Golden values were minted by the UNMODIFIED reference binary
(oracle/_ref/sietill_ref) with tools/make_digits_golden.py; the packs hold the raw frames as they lie in the files.

Bars (north_star): alignments / density choices / counts bit-exact; AM score 1e-10; WER after one EM iteration identical.
"""
import glob
import os
import tempfile

import numpy as np
import pytest

from speechrecognition_b200 import mixio, synth

HERE = os.path.dirname(os.path.abspath(__file__))
PACKS = sorted(glob.glob(os.path.join(HERE, "golden", "digits_*.npz")) +
               glob.glob(os.path.join(HERE, "golden", "_local", "digits_*.npz")))
TDP = (20.0, 0.0, 20.0)            # train_demo_corpus.config:28-30
WORD_PENALTY, AM_THRESHOLD = 20.0, 100.0


def load(path):
    g = np.load(path)
    return {k: g[k] for k in g.files}


def mix_tables(blob, tmp):
    p = os.path.join(tmp, "m.mix")
    blob.tofile(p)
    return mixio.read_mix(p)[1]


def test_packs_present():
    assert PACKS, "tests/golden/digits_*.npz missing (tools/make_digits_golden.py)"


@pytest.mark.parametrize("pack", PACKS[:1], ids=os.path.basename)
def test_oracle_port_on_real_data(pack):
    """-m "not gpu": the C restatement against the reference's outputs on real data — feature processing (sums of the
    processed utterances), the full-DP alignment with the reference's own lin.mix model, and the E-step counts of 0.mix."""
    from oracle.binding import Oracle, oracle_process_features
    g = load(pack)
    fo, wo = g["frame_off"].astype(np.int64), g["word_off"].astype(np.int64)
    lex = synth.Lexicon()
    n = min(48, len(fo) - 1)
    xs, auts = [], []
    for u in range(n):
        x = oracle_process_features(g["raw"][fo[u]:fo[u + 1]], 12, 1, 3, g["norm_mean"], g["norm_std"], True)
        assert float(x.astype(np.float64).sum()) == g["feature_sums"][u]
        xs.append(x)
        auts.append(lex.segment_automaton(g["words"][wo[u]:wo[u + 1]]))
    x = np.concatenate(xs)
    with tempfile.TemporaryDirectory() as tmp:
        t = mix_tables(g["lin_mix"], tmp)
    O = Oracle(25, lex.num_states, "none", True, TDP, lex.silence_state)
    d = 25
    M = len(t["mean_wacc"])
    tabs = {"means": np.zeros(M * d), "mean_acc": t["mean_acc"].reshape(-1), "mean_w": np.zeros(M), "mean_wlog": np.zeros(M),
            "mean_wacc": t["mean_wacc"], "mean_refs": np.ones(M, np.uint64), "vars": np.zeros(M * d), "vars_inv": np.zeros(M * d),
            "var_acc": t["var_acc"].reshape(-1), "var_wacc": t["var_wacc"], "var_refs": np.ones(M, np.uint64), "norm": np.zeros(M),
            "mix_offsets": t["mix_offsets"], "mean_idx": t["mean_idx"], "var_idx": t["var_idx"]}
    O.load_tables(tabs)
    O.finalize()                     # MixtureModel::read re-runs finalize on the stored statistics
    off = np.concatenate([[0], np.cumsum([len(v) for v in xs])]).astype(np.uint64)
    aoff = np.concatenate([[0], np.cumsum([len(a) for a in auts])]).astype(np.uint32)
    states, _ = O.align_batch(x, off, np.concatenate(auts), aoff, threads=8)
    assert np.array_equal(states, g["ali_0_0"][:len(states)].astype(np.uint32))


def digits_gpu(pack, tmp_path):
    from oracle.binding import Reference, have_ref
    from speechrecognition_b200 import capi, hostapi
    g = load(pack)
    fo, wo = g["frame_off"], g["word_off"]
    U, F = len(fo) - 1, int(fo[-1])
    lex = synth.Lexicon()
    # ---- load-time feature processing on the device (SignalAnalysis::process_features), raw 12-dim frames in
    ctx = capi.Context(0)
    x = ctx.process_features(g["raw"], fo, 12, 1, 3, g["norm_mean"], g["norm_std"], True)
    ctx.close()
    sums = np.array([float(x[int(fo[u]):int(fo[u + 1])].astype(np.float64).sum()) for u in range(U)])
    assert np.array_equal(sums, g["feature_sums"]), "processed features differ from the reference's"
    # ---- the Trainer's own linear segmentation (host code) against the reference's
    lin = np.zeros(F, np.uint16)
    for u in range(U):
        aut = lex.segment_automaton(g["words"][int(wo[u]):int(wo[u + 1])])
        lin[int(fo[u]):int(fo[u + 1])] = hostapi.linear_segmentation(x[int(fo[u]):int(fo[u + 1])], aut, approx=True)[0]
    assert np.array_equal(lin, g["ali_linear"]), "linear segmentation differs from the reference's"
    # ---- Trainer::train, num-splits 0: linear segmentation -> lin.mix -> realign (full DP) -> 1 EM iteration -> 0.mix
    out = str(tmp_path) + "/"
    stats, ali, model, timers = hostapi.train(25, "none", x, fo, g["words"].astype(np.uint64), wo.astype(np.uint64), TDP,
                                              min_obs=1, num_splits=0, num_aligns=1, num_estimates=1, alignment_pruning=False,
                                              initial=None, mixture_path=out, alignment_path=out)
    assert np.array_equal(stats[:, :3], g["stats"][:, :3])
    np.testing.assert_allclose(stats[:, 3], g["am_scores"], rtol=1e-10)           # AM score after lin.mix / after 0.mix
    np.testing.assert_allclose(stats[:, 3], g["stats"][:, 3], rtol=2e-5)          # (the stats file keeps 6 digits)
    assert np.array_equal(ali["state"], g["ali_0_0"]), "Viterbi alignment differs from the reference's 0-0.dump"
    dumped, _ = mixio.read_alignment(out + "0-0.dump")
    assert np.array_equal(dumped["state"], g["ali_0_0"]) and np.all(dumped["count"] == 1) and np.all(dumped["weight"] == 1.0)
    for name, key in (("lin.mix", "lin_mix"), ("0.mix", "mix0")):
        got = mixio.read_mix(out + name)[1]
        want = mix_tables(g[key], str(tmp_path))
        for k in ("mean_idx", "var_idx", "mix_offsets", "mean_wacc", "var_wacc"):
            assert np.array_equal(got[k], want[k]), (name, k)                    # topology and counts: exact
        np.testing.assert_allclose(got["mean_acc"], want["mean_acc"], rtol=1e-11, atol=1e-9, err_msg=name)
        np.testing.assert_allclose(got["var_acc"], want["var_acc"], rtol=1e-11, atol=1e-9, err_msg=name)
    # ---- recognition with the GPU-trained model: word sequences, WER / SER and S/I/D as the reference printed them
    hyps, rs = hostapi.recognize(model, x, fo, g["words"].astype(np.uint64), wo.astype(np.uint64), TDP, WORD_PENALTY, AM_THRESHOLD)
    ho = g["hyp_off"]
    for u in range(U):
        assert np.array_equal(hyps[u], g["hyp_words"][int(ho[u]):int(ho[u + 1])]), u
    wer, ser, sub, ins, dele, errs, nref = g["wer"]
    assert (rs["substitutions"], rs["insertions"], rs["deletions"], rs["errors"], rs["ref_words"]) == (sub, ins, dele, errs, nref)
    assert abs(rs["wer"] - wer) < 5e-5 and abs(rs["ser"] - ser) < 5e-5          # the reference prints 6 decimals
    # ---- the GPU-written 0.mix fed to the UNMODIFIED reference Recognizer gives the same hypotheses (hence the same WER)
    checked = 0
    if have_ref():
        R = Reference(25, lex.num_states, "none", True, TDP, lex.silence_state, mix_path=out + "0.mix",
                      extra_config={"am-threshold": AM_THRESHOLD, "word-penalty": WORD_PENALTY})
        for u in range(0, U, max(1, U // 64)):
            got = R.recognize(x[int(fo[u]):int(fo[u + 1])])
            assert np.array_equal(got, g["hyp_words"][int(ho[u]):int(ho[u + 1])]), u
            checked += 1
    return {"segments": U, "frames": F, "wer": rs["wer"], "ser": rs["ser"], "timers": timers, "ref_recognizer_utts": checked}


@pytest.mark.gpu
@pytest.mark.parametrize("pack", PACKS, ids=os.path.basename)
def test_digits_one_em_iteration_and_wer(pack, tmp_path):
    r = digits_gpu(pack, tmp_path)
    print("digits:", r)
