"""-m gpu, needs >= 2 GPUs (skipped otherwise): two ranks, one B200 each, utterance shards, ONE ncclAllReduce of the
packed statistics inside sgmm_accumulate.  Every rank must end up with the statistics a single GPU
computes over the whole corpus: counts exactly, sums to 1e-12 (the reduction order differs), alignments of the shards
equal to the corresponding slices of the single-GPU alignment."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from tests import helpers

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from speechrecognition_b200 import capi, hostapi
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    O, corp, _ = helpers.trained_oracle(dim=33, n_utts=30, splits=2, seed=11, min_obs=10, dens_gen=4)
    fo, ao = corp["frame_off"], corp["aut_off"]
    b = hostapi.shard_utterances(fo, world).astype(np.int64)
    u0, u1 = int(b[rank]), int(b[rank + 1])
    f0, f1 = int(fo[u0]), int(fo[u1])
    a0, a1 = int(ao[u0]), int(ao[u1])
    ctx = capi.Context(rank)
    uid = torch.from_numpy(capi.Context.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)).cuda()
    dist.broadcast(uid, 0)
    ctx.comm_init(rank, world, uid.cpu().numpy())
    ctx.set_model_tables(33, O.tables())
    ctx.set_features(corp["x"][f0:f1])
    ctx.set_corpus(fo[u0:u1 + 1] - fo[u0], corp["automata"][a0:a1], ao[u0:u1 + 1] - ao[u0])
    ctx.set_tdp(*O.tdp, O.silence)
    ali, costs = ctx.align()
    cnt, sx, sxx, ssum = ctx.accumulate(first_pass=False)
    am = ctx.am_score_sum()
    # the posterior-weighted E-step goes through the same all-reduce (same alignment, sum scores)
    scnt, ssx, ssxx, sssum = ctx.accumulate_soft()
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), ali=ali["state"], costs=costs, cnt=cnt, sx=sx, sxx=sxx,
             ssum=ssum, am=am, f0=f0, f1=f1, u0=u0, u1=u1, scnt=scnt, ssx=ssx, ssxx=ssxx, sssum=sssum)
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_ranks_equal_one_gpu(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    O, corp, _ = helpers.trained_oracle(dim=33, n_utts=30, splits=2, seed=11, min_obs=10, dens_gen=4)
    ctx = helpers.gpu_context(O, corp, 33)
    ali, costs = ctx.align()
    cnt, sx, sxx, ssum = ctx.accumulate(first_pass=False)
    scnt, ssx, ssxx, sssum = ctx.accumulate_soft()
    ctx.close()
    r = [np.load(tmp_path / ("rank%d.npz" % k)) for k in range(world)]
    for k in range(world):
        assert np.array_equal(r[k]["ali"], ali["state"][int(r[k]["f0"]):int(r[k]["f1"])])
        assert np.array_equal(r[k]["costs"], costs[int(r[k]["u0"]):int(r[k]["u1"])])
        assert np.array_equal(r[k]["cnt"], cnt)                                   # counts: exact, on every rank
        np.testing.assert_allclose(r[k]["sx"], sx, rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(r[k]["sxx"], sxx, rtol=1e-12, atol=1e-12)
        assert float(r[k]["ssum"]) == pytest.approx(ssum, rel=1e-12)
        assert float(r[k]["am"]) == pytest.approx(ssum, rel=1e-12)
        np.testing.assert_allclose(r[k]["scnt"], scnt, rtol=1e-11, atol=1e-12)      # soft statistics, all-reduced
        np.testing.assert_allclose(r[k]["ssx"], ssx, rtol=1e-11, atol=1e-10)
        np.testing.assert_allclose(r[k]["ssxx"], ssxx, rtol=1e-11, atol=1e-10)
        assert float(r[k]["sssum"]) == pytest.approx(sssum, rel=1e-12)
    assert np.array_equal(r[0]["sx"], r[1]["sx"]) and np.array_equal(r[0]["sxx"], r[1]["sxx"])
    assert int(r[0]["f1"]) == int(r[1]["f0"]) and int(r[1]["f1"]) == len(corp["x"])
