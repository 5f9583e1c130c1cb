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


# ---- stage 3 on shards: the collectives around the kernels, with numpy stand-ins for the kernels ----
class _CpuOps:
    """Stand-ins with the signatures of stage3.linregress_all / build_operand / gemm_3xtf32 (float64 on the CPU)."""

    def __init__(self, values):
        self.lut = np.asarray(values, np.float64)

    def _vals(self, levels):
        lv = levels.numpy().view(np.uint16) if levels.dtype == torch.int16 else levels.numpy()
        return self.lut[lv.astype(np.int64)]

    def linregress_all(self, levels, P, nlevels, d_lut, xs, shift):
        return {"mean": self._vals(levels)[:, :P].mean(axis=0)}

    def build_operand(self, levels, d_sel, nlevels, d_lut, d_mean, rows_are_frames, scale=1.0):
        x = (self._vals(levels)[:, d_sel.numpy()] - d_mean.numpy().astype(np.float64)[None, :]) * scale
        x = x if rows_are_frames else x.T
        hi = torch.from_numpy(np.ascontiguousarray(x))
        n2 = torch.from_numpy((x ** 2).sum(axis=1)) if rows_are_frames else None
        return hi, torch.zeros_like(hi), n2, 4.0           # a scale the caller has to undo (alpha)

    def gemm_3xtf32(self, a_hi, a_lo, b_hi, b_lo, K, epilogue=0, alpha=1.0, row_norm2=None, col_norm2=None, symmetric=False):
        acc = (a_hi + a_lo).double() @ (b_hi + b_lo).double().T * 16.0     # operands "scaled by 4" each
        if epilogue == 0:
            return acc * alpha
        return alpha * torch.sqrt(torch.clamp(16.0 * row_norm2[:, None] + 16.0 * col_norm2[None, :] - 2.0 * acc, min=0.0))


def _worker3(rank, world, port, nf, P, seed, tmp):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from conan_b200 import sharded
    rng = np.random.default_rng(seed)
    levels = rng.integers(0, 101, size=(nf, P)).astype(np.uint16)
    values = np.linspace(0.0, 1.0, 101)
    ops = _CpuOps(values)
    bounds = np.linspace(0, nf, world + 1).astype(int) + np.array([0] + [1] * (world - 1) + [0])[:world + 1]   # uneven shards
    mine = torch.from_numpy(levels[bounds[rank]:bounds[rank + 1]].view(np.int16).copy())
    shift = np.zeros(P, np.float32)
    mean, nf_total = sharded.global_mean(mine, P, 101, None, shift, ops=ops)
    sel = np.arange(3, P, 2).astype(np.int32)
    cov = sharded.covariance_sharded(mine, sel, 101, None, mean[sel], nf_total, ops=ops)
    block, row0 = sharded.frame_distances_sharded(mine, sel, 101, None, mean[sel], 7.0, ops=ops)
    np.savez(os.path.join(tmp, "s3rank%d.npz" % rank), levels=levels, mean=mean, nf_total=nf_total, cov=cov.numpy(),
             block=block.numpy(), row0=row0, sel=sel)
    dist.destroy_process_group()


@pytest.mark.parametrize("world,nf", [(2, 41), (3, 50)])
def test_sharded_stage3_collectives(tmp_path, world, nf):
    P = 37
    mp.spawn(_worker3, args=(world, _free_port(), nf, P, 5, str(tmp_path)), nprocs=world, join=True)
    ranks = [np.load(str(tmp_path / ("s3rank%d.npz" % r))) for r in range(world)]
    levels, sel = ranks[0]["levels"], ranks[0]["sel"]
    vals = np.linspace(0.0, 1.0, 101)[levels.astype(np.int64)]
    want_cov = np.cov(vals[:, sel].T)
    x = (vals[:, sel] - vals[:, sel].mean(axis=0)) / 7.0
    want_d = np.sqrt(np.maximum(((x[:, None, :] - x[None, :, :]) ** 2).sum(-1), 0.0))
    rows = 0
    for r in ranks:
        assert int(r["nf_total"]) == nf and np.allclose(r["mean"], vals.mean(axis=0), atol=1e-13)
        assert np.allclose(r["cov"], want_cov, atol=1e-12)                      # every rank holds the full covariance
        b, r0 = r["block"], int(r["row0"])
        assert r0 == rows and np.allclose(b, want_d[r0:r0 + b.shape[0]], atol=1e-6)   # row blocks tile the matrix
        rows += b.shape[0]
    assert rows == nf
