"""CPU, world_size 2 over gloo: the N>1 host logic — utterance sharding (shard_utterances), per-rank partial
statistics that start from zero, ONE all-reduce of the packed [count | sum_x | sum_xx | score_sum] block,
the 1e-4 variance seed added once after the reduce — with the oracle standing in for the
CUDA E-step (this box has no GPU).  The reduced statistics must equal the single-process ones."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from speechrecognition_b200 import hostapi, synth
from tests import helpers


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _packed_partial(O, tables0, states, x, dim):
    """What sgmm_accumulate returns on one rank: per-density count / sum_x / sum_xx from ZERO + score sum."""
    O.load_tables(tables0)
    am = O.am_score(states, x) * len(states) if len(states) else 0.0
    O.accumulate(states, x)
    t = O.tables()
    mi, vi = t["mean_idx"].astype(np.int64), t["var_idx"].astype(np.int64)
    cnt = t["mean_wacc"][mi]
    sx = t["mean_acc"].reshape(-1, dim)[mi]
    sxx = t["var_acc"].reshape(-1, dim)[vi] - 1e-4          # NO_POOLING model: one variance slot per density
    return np.concatenate([cnt, sx.reshape(-1), sxx.reshape(-1), [am]])


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    O, corp, ali = helpers.trained_oracle(dim=12, n_utts=10, splits=1, seed=3, min_obs=8)
    tables0 = O.tables()
    fo = corp["frame_off"]
    b = hostapi.shard_utterances(fo, world).astype(np.int64)
    f0, f1 = int(fo[b[rank]]), int(fo[b[rank + 1]])
    part = _packed_partial(O, tables0, ali[f0:f1], corp["x"][f0:f1], 12)
    t = torch.from_numpy(part)
    dist.all_reduce(t)                                        # the one collective of the EM iteration
    np.save(os.path.join(out_dir, "rank%d.npy" % rank), t.numpy())
    np.save(os.path.join(out_dir, "bounds%d.npy" % rank), b)
    dist.destroy_process_group()


def test_sharded_accumulation_equals_single_process(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npy"), np.load(tmp_path / "rank1.npy")
    assert np.array_equal(r0, r1)                             # every rank holds the same reduced block
    assert np.array_equal(np.load(tmp_path / "bounds0.npy"), np.load(tmp_path / "bounds1.npy"))
    O, corp, ali = helpers.trained_oracle(dim=12, n_utts=10, splits=1, seed=3, min_obs=8)
    whole = _packed_partial(O, O.tables(), ali, corp["x"], 12)
    D = len(O.tables()["mean_idx"])
    assert np.array_equal(r0[:D], whole[:D])                  # counts: exact
    assert r0[:D].sum() == whole[:D].sum() <= len(ali)     # frames of emptied mixtures count nowhere
    np.testing.assert_allclose(r0[D:], whole[D:], rtol=1e-9, atol=1e-9)
