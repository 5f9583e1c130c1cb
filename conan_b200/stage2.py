"""Stage 2 driver: streaming per-pair statistics over the level stack (conan_pair_stats).

Stands where the reference runs the frame loop of read_process_frames (conan.py:782-874).
"""
import ctypes

import numpy as np
import torch

from . import _lib


class PairStats:
    """Device tensors with the per-pair and per-frame integer statistics of one call."""

    PAIR_FIELDS = (("sum_k", torch.int64), ("sum_k2", torch.int64), ("n_int", torch.int32),
                   ("n_enc", torch.int32), ("first_f", torch.int32), ("last_f", torch.int32),
                   ("status", torch.uint8), ("decided", torch.uint8))
    FRAME_FIELDS = (("native", torch.int32), ("nonnative", torch.int32), ("ssd_prev", torch.int64),
                    ("ssd_first", torch.int64))

    def __init__(self, nf, P, device):
        self.nf, self.P = nf, P
        for name, dt in self.PAIR_FIELDS:
            setattr(self, name, torch.empty(P, dtype=dt, device=device))
        for name, dt in self.FRAME_FIELDS:
            setattr(self, name, torch.empty(nf, dtype=dt, device=device))

    def host(self):
        """dict of numpy arrays (device -> host copy)."""
        out = {}
        for name, _ in self.PAIR_FIELDS + self.FRAME_FIELDS:
            t = getattr(self, name)
            if t.dtype == torch.uint16:
                out[name] = t.view(torch.int16).cpu().numpy().view(np.uint16)
            else:
                out[name] = t.cpu().numpy()
        return out


def pair_stats_workspace_bytes(nf, P):
    n = _lib.load().conan_pair_stats_workspace_bytes(nf, P)
    if n == 0:
        raise _lib.ConanError(-1, "conan_pair_stats_workspace_bytes rejected its arguments")
    return n


def pair_stats(d_levels, P, nlevels, lo_cnt, hi_first, equal_cut, frame0=0, first_row=None, prev_row=None,
               status_in=None, out=None, work=None, stream=None, sparse=None):
    """Launch stage 2 on a device level stack [nf, stride] (uint16); returns PairStats (device).
    sparse=True: the variant that skips 8-frame stages of nothing but last levels (same results; for stacks in which
    few pairs ever come within trunc); None: the library's default (environment variable CONAN_STATS_SPARSE)."""
    lib = _lib.load()
    if not d_levels.is_cuda:
        raise _lib.ConanError(-1, "stage 2 needs device tensors: there is no CPU fallback")
    nf, stride = d_levels.shape
    dev = d_levels.device
    if out is None:
        out = PairStats(nf, P, dev)
    if work is None:
        work = torch.empty(pair_stats_workspace_bytes(nf, P), dtype=torch.uint8, device=dev)
    po = _lib.PairStatsOut(*[getattr(out, n).data_ptr() for n, _ in PairStats.PAIR_FIELDS])
    fo = _lib.FrameStatsOut(*[getattr(out, n).data_ptr() for n, _ in PairStats.FRAME_FIELDS])
    s = stream if stream is not None else torch.cuda.current_stream(dev)
    ptr = lambda t: 0 if t is None else t.data_ptr()
    if sparse is None:
        _lib.check(lib.conan_pair_stats(d_levels.data_ptr(), stride, nf, P, frame0, nlevels, lo_cnt, hi_first, equal_cut,
                                        ptr(first_row), ptr(prev_row), ptr(status_in), ctypes.byref(po), ctypes.byref(fo),
                                        work.data_ptr(), work.numel(), s.cuda_stream))
    else:
        _lib.check(lib.conan_pair_stats_ex(d_levels.data_ptr(), stride, nf, P, frame0, nlevels, lo_cnt, hi_first, equal_cut,
                                           ptr(first_row), ptr(prev_row), ptr(status_in), ctypes.byref(po), ctypes.byref(fo),
                                           work.data_ptr(), work.numel(), 1 if sparse else 0, s.cuda_stream))
    out.work = work
    out._keepalive = (first_row, prev_row, status_in)
    return out


def shard_summary(d_levels, P, nlevels, lo_cnt, hi_first, equal_cut, stream=None):
    """(decided, status) uint8 device tensors of a frame shard (conan_shard_summary)."""
    lib = _lib.load()
    if not d_levels.is_cuda:
        raise _lib.ConanError(-1, "stage 2 needs device tensors: there is no CPU fallback")
    nf, stride = d_levels.shape
    dev = d_levels.device
    decided = torch.empty(P, dtype=torch.uint8, device=dev)
    status = torch.empty(P, dtype=torch.uint8, device=dev)
    s = stream if stream is not None else torch.cuda.current_stream(dev)
    _lib.check(lib.conan_shard_summary(d_levels.data_ptr(), stride, nf, P, nlevels, lo_cnt, hi_first, equal_cut,
                                       decided.data_ptr(), status.data_ptr(), s.cuda_stream))
    return decided, status
