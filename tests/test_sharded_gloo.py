"""The multi-GPU exchange logic (conan_b200/sharded.py) on CPU: gloo, world_size 2 and 3.

The kernels are replaced by the CPU oracle through `local_pass`; what is tested is the protocol:
row exchange, exclusive scan of the hysteresis state across shards, and the all-reduce merges.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


class _Stats:
    pass


def _oracle_pass(levels, P, nlevels, lo, hi, eq, frame0=0, first_row=None, prev_row=None, status_in=None, out=None,
                 work=None):
    from oracle import oracle
    n = lambda t: None if t is None else t.numpy().view(np.uint16) if t.dtype == torch.int16 else t.numpy()
    o = oracle.pair_stats(n(levels)[:, :P], lo, hi, eq, frame0=frame0,
                          first_row=None if first_row is None else n(first_row)[:P],
                          prev_row=None if prev_row is None else n(prev_row)[:P],
                          status_in=None if status_in is None else n(status_in))
    st = _Stats()
    for k, v in o.items():
        setattr(st, k, torch.from_numpy(v.view(np.int16) if v.dtype == np.uint16 else v))
    return st


def _oracle_summary(levels, P, nlevels, lo, hi, eq):
    lv = (levels.numpy().view(np.uint16) if levels.dtype == torch.int16 else levels.numpy())[:, :P].astype(np.int64)
    if eq:
        hi = lo
    decisive = (lv < lo) | (lv >= hi)
    decided = decisive.any(axis=0)
    last = lv.shape[0] - 1 - np.argmax(decisive[::-1], axis=0)
    status = (lv[last, np.arange(P)] < lo) & decided
    return torch.from_numpy(decided.astype(np.uint8)), torch.from_numpy(status.astype(np.uint8))


def _worker(rank, world, port, nf, P, seed, tmp):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from conan_b200 import sharded
    rng = np.random.default_rng(seed)
    base = rng.uniform(0.2, 0.9, size=P)
    r = np.clip(base + np.cumsum(rng.normal(0, 0.03, size=(nf, P)), axis=0), 0, 1)
    r[:, ::7] = 0.55                               # pairs that never leave the buffer
    r[0, ::14] = 0.1                               # ... half of them switched on in frame 0
    levels = np.rint(r * 100).astype(np.uint16)
    bounds = np.linspace(0, nf, world + 1).astype(int)
    mine = torch.from_numpy(levels[bounds[rank]:bounds[rank + 1]].view(np.int16).copy())
    st = sharded.pair_stats_sharded(mine, P, 101, 50, 61, 0, frame0=int(bounds[rank]), local_pass=_oracle_pass,
                                    local_summary=_oracle_summary)
    res = {k: getattr(st, k).numpy().copy() for k in ("sum_k", "sum_k2", "n_int", "n_enc", "first_f", "last_f",
                                                     "status", "native", "nonnative", "ssd_prev", "ssd_first")}
    np.savez(os.path.join(tmp, "rank%d.npz" % rank), levels=levels, **res)
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,nf", [(2, 60), (3, 91)])
def test_sharded_stats_equal_single_run(orc, tmp_path, world, nf):
    P = 157
    mp.spawn(_worker, args=(world, _free_port(), nf, P, 3, str(tmp_path)), nprocs=world, join=True)
    ranks = [np.load(str(tmp_path / ("rank%d.npz" % r))) for r in range(world)]
    levels = ranks[0]["levels"]
    want = orc.pair_stats(levels, 50, 61, 0)
    assert want["n_enc"].max() > 1 and (want["n_int"][::14] == nf).all()   # the carried pairs stay on
    for r in ranks:      # per-pair results are identical on every rank and equal the single run
        for k in ("sum_k", "sum_k2", "n_int", "n_enc", "first_f", "last_f", "status"):
            assert np.array_equal(r[k], want[k]), k
    for k in ("native", "nonnative", "ssd_prev", "ssd_first"):   # per-frame results concatenate
        assert np.array_equal(np.concatenate([r[k] for r in ranks]), want[k]), k
