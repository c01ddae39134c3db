"""-m gpu, collected LAST: sgmm_score_sum_static — the allocation-free variant of the
score sum whose partials live in a file-scope __device__ array (csrc/k_noalloc.cu; no counterpart in the reference)
— against the existing entry point (sgmm_accumulate's score_sum, the
k_score_partial / k_score_final pair with its cudaMalloc'ed partials), through the C-ABI, on two calls in a row.

The two summation trees differ, so the bar is the statistics tolerance of the other GPU tests (1e-12 relative), not
bit-identity.  torch only provides the device buffers the entry point is handed (it allocates nothing itself)."""
import math

import numpy as np
import pytest

from speechrecognition_b200 import synth
from tests import helpers

pytestmark = pytest.mark.gpu

RTOL = 1e-12   # as tests/test_gpu_parity.py STAT_RTOL


def test_score_sum_static_matches_existing_entry_point():
    import torch
    from speechrecognition_b200 import capi

    rng = np.random.default_rng(23)
    dim = 25
    lex = synth.Lexicon()
    tables = synth.random_model(rng, dim, lex.num_states, 4, "none")
    corp = synth.make_corpus(rng, lex, tables, dim, 40)
    ctx = capi.Context(0)
    ctx.set_model_tables(dim, tables)
    ctx.set_features(corp["x"])
    ctx.set_corpus(corp["frame_off"], corp["automata"], corp["aut_off"])

    # two different alignments of the same corpus (Viterbi, then the uniform linear segmentation) -> two per-frame score
    # vectors and the existing entry point's sums
    ctx.set_tdp(*helpers.TDP, lex.silence_state)
    want, scores = [], []
    for which in ("viterbi", "linear"):
        if which == "viterbi":
            ctx.align()
        else:
            ctx.set_alignment(synth.linear_alignment(corp["frame_off"], corp["automata"], corp["aut_off"]))
        _, _, _, ssum = ctx.accumulate(first_pass=False)
        _, fscore = ctx.last_frame_results()
        want.append(ssum)
        scores.append(fscore)
    assert want[0] != want[1] and not np.array_equal(scores[0], scores[1])

    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev)          # an explicit, non-NULL cudaStream_t (NULL would mean the context's own stream)
    d_in = [torch.from_numpy(s).to(dev) for s in scores]
    # a long vector first, then the short ones: stale partials of an earlier call must never be read
    n_long = 3 * 4096 * 37 + 11
    long_host = rng.standard_normal(n_long) * 50.0
    d_long = torch.from_numpy(long_host).to(dev)
    d_out = torch.full((4,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    ctx.score_sum_static(0, 0, d_out[3:].data_ptr(), stream.cuda_stream)                       # warm-up: n = 0 -> 0.0
    ctx.score_sum_static(d_long.data_ptr(), n_long, d_out[2:].data_ptr(), stream.cuda_stream)
    # the two calls in a row that are checked against the existing entry point
    ctx.score_sum_static(d_in[0].data_ptr(), d_in[0].numel(), d_out[0:].data_ptr(), stream.cuda_stream)
    ctx.score_sum_static(d_in[1].data_ptr(), d_in[1].numel(), d_out[1:].data_ptr(), stream.cuda_stream)
    torch.cuda.synchronize()
    got = d_out.cpu().numpy()
    assert got[3] == 0.0
    for i in range(2):
        assert abs(got[i] - want[i]) <= RTOL * abs(want[i]), (i, got[i], want[i])
        assert abs(got[i] - math.fsum(scores[i])) <= RTOL * abs(want[i])
    assert abs(got[2] - math.fsum(long_host)) <= RTOL * float(np.abs(long_host).sum())
    # same input again: the tree depends on n alone -> the same bits
    ctx.score_sum_static(d_in[0].data_ptr(), d_in[0].numel(), d_out[2:].data_ptr(), stream.cuda_stream)
    torch.cuda.synchronize()
    assert d_out[2].item() == got[0]

    # sizes beyond the static scratch are rejected with a clear error, and nothing is launched
    launches = ctx.launch_count()
    with pytest.raises(capi.SgmmError, match="exceed the static scratch"):
        ctx.score_sum_static(d_long.data_ptr(), 4096 * 16384 + 1, d_out.data_ptr(), stream.cuda_stream)
    with pytest.raises(capi.SgmmError, match="null argument"):
        ctx.score_sum_static(0, 5, d_out.data_ptr(), stream.cuda_stream)
    assert ctx.launch_count() == launches
    # ... and the context is still usable afterwards
    _, _, _, again = ctx.accumulate(first_pass=False)
    assert again == want[1]
    ctx.close()
