"""Frame shards across GPUs: what is exchanged so that stage 2 equals one long single-GPU run.

Frames are independent for stage 1, so every rank computes the level stack of its own contiguous
frame range.  Stage 2 needs three small exchanges:
  1. the level row of GLOBAL frame 0 (defines "native" contacts and the rmsd-to-first reference)
     and the last row of the preceding shard (rmsd-to-previous of the shard's first frame);
  2. the interaction status entering each shard: every rank publishes, per pair, whether its shard
     has a frame outside the hysteresis buffer and the status it then leaves with
     (conan_shard_summary, a short backward scan from the shard's end); an exclusive scan over
     ranks gives the status entering every shard before the streaming pass runs;
  3. the per-pair sums: all-reduce SUM of the integer moments and counts, MIN / MAX of the
     first / last frame and level extremes.  Integer sums make the merge exact and
     order-independent.
The collectives are torch.distributed (NCCL over NVLink on the GPUs, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist

_INT_MAX = 2 ** 31 - 1


def _raw(t):
    """Byte view for pure data movement (the collectives have no 16-bit unsigned type)."""
    return t.contiguous().view(torch.uint8)


def exchange_rows(first_row_local, last_row_local, group=None):
    """-> (row of global frame 0, last row of the preceding rank or None for rank 0)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    first = _raw(first_row_local).clone()
    dist.broadcast(first, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    mine = _raw(last_row_local)
    lasts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(lasts, mine, group=group)
    prev = lasts[rank - 1] if rank > 0 else None
    back = lambda t: t.view(first_row_local.dtype)
    return back(first), (None if prev is None else back(prev))


def entering_status(decided, status_out, group=None):
    """Exclusive scan of the shards' (decided, status) summaries -> status entering this rank."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dec = [torch.empty_like(decided) for _ in range(world)]
    sta = [torch.empty_like(status_out) for _ in range(world)]
    dist.all_gather(dec, decided.contiguous(), group=group)
    dist.all_gather(sta, status_out.contiguous(), group=group)
    s = torch.zeros_like(status_out)
    for q in range(rank):
        s = torch.where(dec[q] != 0, sta[q], s)
    return s


def merge_pair_stats(st, group=None):
    """All-reduce the per-pair fields of `st` (an object with tensor attributes) in place."""
    world = dist.get_world_size(group)
    for name in ("sum_k", "sum_k2", "n_int", "n_enc"):
        dist.all_reduce(getattr(st, name), op=dist.ReduceOp.SUM, group=group)
    first = st.first_f
    first.copy_(torch.where(first < 0, torch.full_like(first, _INT_MAX), first))
    dist.all_reduce(first, op=dist.ReduceOp.MIN, group=group)
    first.copy_(torch.where(first == _INT_MAX, torch.full_like(first, -1), first))
    dist.all_reduce(st.last_f, op=dist.ReduceOp.MAX, group=group)
    dist.all_reduce(st.decided, op=dist.ReduceOp.MAX, group=group)
    # the status after the last frame is the last shard's
    src = dist.get_global_rank(group, world - 1) if group is not None else world - 1
    dist.broadcast(st.status, src=src, group=group)
    return st


def pair_stats_sharded(levels, P, nlevels, lo_cnt, hi_first, equal_cut, frame0, out=None, work=None, group=None,
                       local_pass=None, local_summary=None):
    """Stage 2 over this rank's frame shard, merged so that per-pair results equal a single run.

    Per-frame results stay local to the rank (frames are disjoint).  `local_pass` / `local_summary`
    exist for the CPU tests: callables with the signatures of stage2.pair_stats / shard_summary."""
    if local_pass is None:
        from . import stage2
        local_pass, local_summary = stage2.pair_stats, stage2.shard_summary
    first_row, prev_row = exchange_rows(levels[0], levels[-1], group)
    decided, status = local_summary(levels, P, nlevels, lo_cnt, hi_first, equal_cut)
    s_in = entering_status(decided, status, group)
    st = local_pass(levels, P, nlevels, lo_cnt, hi_first, equal_cut, frame0=frame0, first_row=first_row,
                    prev_row=prev_row, status_in=s_in, out=out, work=work)
    return merge_pair_stats(st, group)


# ------------------------------------------------------------------------------------------------
# stage 3 on frame shards ("GEMMs")
# ------------------------------------------------------------------------------------------------
def global_mean(levels, P, nlevels, d_lut, shift, group=None, ops=None):
    """Mean float16 distance of every pair over ALL ranks' frames (float64 [P] on the host): the local sums of
    (value - shift) are added across ranks.  `shift` must be the same on every rank (e.g. from the merged sum_k)."""
    import numpy as np
    if ops is None:
        from . import stage3 as ops          # (the CPU tests pass stand-ins for the kernels)
    nf = levels.shape[0]
    local = ops.linregress_all(levels, P, nlevels, d_lut, np.zeros((1, nf)), shift)["mean"]     # local mean, float64
    shift_h = shift.cpu().numpy() if torch.is_tensor(shift) else np.asarray(shift)
    t = torch.from_numpy((local - shift_h.astype(np.float64)) * nf).to(levels.device)
    n = torch.tensor([float(nf)], dtype=torch.float64, device=levels.device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(n, op=dist.ReduceOp.SUM, group=group)
    return shift_h.astype(np.float64) + t.cpu().numpy() / float(n.item()), int(round(float(n.item())))


def covariance_sharded(levels, sel, nlevels, d_lut, mean, nf_total, group=None, ops=None):
    """np.cov(vectors) over all ranks' frames: the frame axis is the contraction axis, so every rank contracts its
    own frames with the global means (3xTF32 tensor-core GEMM) and the [nsel, nsel] partial products are summed
    with one all-reduce.  Every rank returns the full matrix."""
    import numpy as np
    if ops is None:
        from . import stage3 as ops
    dev = levels.device
    d_sel = torch.from_numpy(np.ascontiguousarray(sel, np.int32)).to(dev)
    d_mean = torch.from_numpy(np.ascontiguousarray(mean, np.float32)).to(dev)
    hi, lo, _, s = ops.build_operand(levels, d_sel, nlevels, d_lut, d_mean, rows_are_frames=False)
    cov = ops.gemm_3xtf32(hi, lo, hi, lo, levels.shape[0], epilogue=0, alpha=1.0 / ((nf_total - 1) * s * s), symmetric=True)
    dist.all_reduce(cov, op=dist.ReduceOp.SUM, group=group)
    return cov


def frame_distances_sharded(levels, sel, nlevels, d_lut, mean, norm, group=None, ops=None):
    """Rows [this rank's frames] x [all frames] of pdist(vectors.T / norm): the output is partitioned by frame
    rows, the centred operands of the other ranks' frames arrive by broadcast (NVLink), and each rank runs the
    distance-epilogue GEMM for its row block.  Returns (block [nf_local, nf_total] float32, frame offset)."""
    import numpy as np
    if ops is None:
        from . import stage3 as ops
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = levels.device
    d_sel = torch.from_numpy(np.ascontiguousarray(sel, np.int32)).to(dev)
    d_mean = torch.from_numpy(np.ascontiguousarray(mean, np.float32)).to(dev)
    hi, lo, n2, s = ops.build_operand(levels, d_sel, nlevels, d_lut, d_mean, rows_are_frames=True, scale=1.0 / norm)
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    counts[rank] = levels.shape[0]
    dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    counts = [int(c) for c in counts.cpu()]
    total, ld = sum(counts), hi.shape[1]
    all_hi = torch.empty((total, ld), dtype=hi.dtype, device=dev)
    all_lo = torch.empty((total, ld), dtype=lo.dtype, device=dev)
    all_n2 = torch.empty(total, dtype=torch.float64, device=dev)
    o = 0
    for r in range(world):
        src = dist.get_global_rank(group, r) if group is not None else r
        for mine, everyone in ((hi, all_hi), (lo, all_lo), (n2, all_n2)):
            part = everyone[o:o + counts[r]]
            if r == rank:
                part.copy_(mine)
            dist.broadcast(part, src=src, group=group)
        o += counts[r]
    block = ops.gemm_3xtf32(hi, lo, all_hi, all_lo, len(sel), epilogue=1, alpha=1.0 / s, row_norm2=n2, col_norm2=all_n2,
                               symmetric=False)
    return block, sum(counts[:rank])
