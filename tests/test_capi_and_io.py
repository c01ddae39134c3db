"""CPU: the C-ABI library loads and exports every symbol include/sietill_b200.h declares (no compute call),
fails loudly without a GPU, and the on-disk formats are byte-compatible with the reference's."""
import os
import re

import numpy as np
import pytest

from speechrecognition_b200 import capi, mixio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(os.path.dirname(__file__), "golden")


def header_symbols():
    src = open(os.path.join(ROOT, "include", "sietill_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sgmm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = capi.load()
    names = header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(capi.SYMBOLS) == names            # the Python binding's list tracks the header
    assert b"sm_100a" in lib.sgmm_version()


def test_every_option_is_documented_in_the_header():
    """Every key sgmm_set_option accepts is described in include/sietill_b200.h."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = open(os.path.join(root, "speechrecognition_b200", "csrc", "sgmm_api.cu")).read()
    body = src[src.index("int sgmm_set_option"):src.index("int sgmm_get_counter")]
    keys = re.findall(r'!strcmp\(key, "([a-z0-9_]+)"\)', body)
    header = open(os.path.join(root, "include", "sietill_b200.h")).read()
    assert len(keys) > 20
    missing = [k for k in keys if '"%s"' % k not in header]
    assert not missing, missing


def test_no_gpu_is_a_loud_error():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.SgmmError, match="no CUDA device"):
        capi.Context(0)


@pytest.mark.parametrize("pooling", ["none", "mixture", "global"])
def test_mix_file_is_byte_compatible(pooling, tmp_path):
    """Write the statistics of the golden trajectory with our writer: the bytes must equal the file the
    reference's MixtureModel::write produced (Mixtures.cpp:834-878)."""
    rec = np.load(os.path.join(G, "trajectory_%s.npz" % pooling))
    from oracle.binding import Oracle, TABLES
    from speechrecognition_b200 import synth
    c = np.load(os.path.join(G, "corpus.npz"))
    dim = int(c["dim"])
    # the .mix was written after the final accumulate+finalize; statistics = t2 tables
    t = {k: rec["t2_" + k] for k in TABLES}
    out = tmp_path / "ours.mix"
    mixio.write_mix(str(out), dim, t)
    want = open(os.path.join(G, "mix_%s.mix" % pooling), "rb").read()
    assert out.read_bytes() == want
    d2, r = mixio.read_mix(str(out))
    assert d2 == dim and r["mix_offsets"][-1] == len(t["mean_idx"])
    live = t["mean_refs"] > 0
    assert np.array_equal(r["mean_wacc"], t["mean_wacc"][live])


def test_alignment_dump_roundtrip(tmp_path):
    a = np.zeros(77, mixio.ALI_DTYPE)
    a["state"] = np.arange(77) % 106
    a["count"] = 1
    a["weight"] = 1.0
    p = tmp_path / "a.dump"
    mixio.write_alignment(str(p), a)
    raw = p.read_bytes()
    assert len(raw) == 16 + 8 * 77                   # u64 max_aligns, u64 frames, 8-byte items (Alignment.cpp:303-308)
    b, k = mixio.read_alignment(str(p))
    assert k == 1 and np.array_equal(a, b)


def test_lexicon_and_automata():
    from speechrecognition_b200 import synth
    lex = synth.Lexicon()
    assert lex.num_states == 106 and lex.silence_state == 0           # Lexicon.cpp:70-85
    aut = lex.segment_automaton([1, 5])
    assert len(aut) == 1 + 18 + 1 + 24 + 1                              # sil w sil w sil, states doubled
    assert aut[0] == 0 and aut[-1] == 0 and aut[1] == 1 and aut[2] == 1
