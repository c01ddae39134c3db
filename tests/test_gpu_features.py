"""-m gpu: sgmm_process_features (SignalAnalysis::process_features on the device) against the
oracle and the committed outputs of the unmodified reference: bit-exact f32."""
import os

import numpy as np
import pytest

from oracle.binding import oracle_process_features
from speechrecognition_b200 import capi
from tests import helpers

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_feature_processing_golden():
    g = np.load(os.path.join(G, "features.npz"))
    ctx = capi.Context(0)
    for ci, (nf, n1, n2, step, norm, emax) in enumerate(g["cases"]):
        mean = g["mean%d" % ci] if norm else None
        std = g["std%d" % ci] if norm else None
        got = ctx.process_features(g["raw%d" % ci], g["off%d" % ci], int(n1), int(n2), int(step), mean, std, bool(emax))
        assert np.array_equal(bits(got), bits(g["out%d" % ci])), ci
    ctx.close()


@pytest.mark.parametrize("norm,emax", [(True, True), (False, False), (True, False)])
def test_feature_processing_matches_oracle_and_stays_resident(norm, emax):
    """A sietill-shaped corpus (12 file features -> 25) processed on the device; the processed matrix is then the
    resident feature matrix: scoring it gives the same bits as uploading the oracle's output."""
    rng = np.random.default_rng(4)
    lens = rng.choice([279, 500, 701, 40, 4], size=60)
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    raw = (rng.standard_normal((int(off[-1]), 12)) * rng.uniform(1, 12, 12)).astype(np.float32)
    mean = rng.uniform(-3, 3, 25) if norm else None
    std = rng.uniform(0.5, 8.0, 25) if norm else None
    want = np.concatenate([oracle_process_features(raw[int(off[u]):int(off[u + 1])], 12, 1, 3, mean, std, emax)
                           for u in range(len(lens))])
    O, corp, ali = helpers.trained_oracle(dim=25, n_utts=8, splits=1)
    ctx = capi.Context(0)
    ctx.set_model_tables(25, O.tables())
    got = ctx.process_features(raw, off, 12, 1, 3, mean, std, emax)
    assert np.array_equal(bits(got), bits(want))
    s1, a1 = ctx.score_all(0, 500)
    ctx.set_features(want)
    s2, a2 = ctx.score_all(0, 500)
    assert np.array_equal(s1, s2) and np.array_equal(a1, a2)
    ctx.close()


def test_feature_processing_errors_are_loud():
    ctx = capi.Context(0)
    raw = np.zeros((10, 12), np.float32)
    with pytest.raises(capi.SgmmError):        # an utterance of deriv_step frames: the reference reads out of bounds
        ctx.process_features(raw, np.array([0, 3, 10], np.uint64), 12, 1, 3)
    with pytest.raises(capi.SgmmError):        # second derivatives of columns that have no first derivative
        ctx.process_features(raw, np.array([0, 10], np.uint64), 2, 5, 3)
    with pytest.raises(capi.SgmmError):
        ctx.process_features(raw, np.array([0, 9], np.uint64), 12, 1, 3)
    ctx.close()


def test_host_signal_analysis_with_normalization_file(tmp_path):
    """The host class SignalAnalysis (same configuration keys, the reference's normalization-file format: two blobs of
    n_features_total doubles, SignalAnalysis.cpp:363-375) driving the device pass, against the oracle."""
    from speechrecognition_b200 import hostapi
    rng = np.random.default_rng(9)
    lens = [120, 4, 279, 33]
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    raw = (rng.standard_normal((int(off[-1]), 12)) * 5.0 + 1.0).astype(np.float32)
    mean, std = rng.uniform(-2, 2, 25), rng.uniform(0.5, 4.0, 25)
    path = tmp_path / "Normalization.bin"
    path.write_bytes(mean.astype("<f8").tobytes() + std.astype("<f8").tobytes())
    want = np.concatenate([oracle_process_features(raw[int(off[u]):int(off[u + 1])], 12, 1, 3, mean, std, True)
                           for u in range(len(lens))])
    model = hostapi.HostModel(25, 106, "none")
    got = model.process_features(raw, off, 12, 1, 3, True, str(path))
    assert np.array_equal(bits(got), bits(want))
    got2 = model.process_features(raw, off, 12, 1, 3, False, "")          # no file: deltas only
    want2 = np.concatenate([oracle_process_features(raw[int(off[u]):int(off[u + 1])], 12, 1, 3, None, None, False)
                            for u in range(len(lens))])
    assert np.array_equal(bits(got2), bits(want2))
    with pytest.raises(RuntimeError):
        model.process_features(raw, off, 12, 1, 3, True, str(tmp_path / "missing.bin"))
    model.close()
