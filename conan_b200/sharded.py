"""Frame shards across GPUs: what is exchanged so that stage 2 equals one long single-GPU run.

Frames are independent for stage 1, so every rank computes the level stack of its own contiguous
frame range.  Stage 2 needs three small exchanges:
  1. the level row of GLOBAL frame 0 (defines "native" contacts and the rmsd-to-first reference)
     and the last row of the preceding shard (rmsd-to-previous of the shard's first frame);
  2. the interaction status entering each shard: every rank publishes, per pair, whether its shard
     has a frame outside the hysteresis buffer and the status it then leaves with
     (conan_shard_summary, a short backward scan from the shard's end); an exclusive scan over
     ranks gives the status entering every shard before the streaming pass runs;
  3. the per-pair sums: all-reduce SUM of the integer moments and counts, MAX of the last frame,
     the decided flag and the negated first frame.  Integer sums make the merge exact and
     order-independent.
Exchanges 1 and 2 travel in one all-gather of a packed record per rank, exchange 3 in two all-reduces
over packed buffers.
The collectives are torch.distributed (NCCL over NVLink on the GPUs, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist

_INT_MIN = -(2 ** 31)


def exchange_summaries(first_row_local, last_row_local, decided, status_out, group=None):
    """ONE all-gather of every rank's packed record (first row, last row, decided, status: 6 bytes per pair)
    -> (row of global frame 0, last row of the preceding rank or None, status entering this rank, status after
    the last rank's last frame).

    The two status results are the exclusive / inclusive scan of the shards' (decided, status) summaries: a
    shard that has a frame outside the hysteresis buffer leaves with its own status whatever it entered with."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    P = decided.numel()
    u8 = lambda t: t.contiguous().view(torch.uint8).reshape(-1)         # the collectives have no 16-bit unsigned type
    rec = torch.cat([u8(first_row_local[:P]), u8(last_row_local[:P]), u8(decided), u8(status_out)])
    if dist.get_backend(group) == "nccl":
        recs = torch.empty((world, rec.numel()), dtype=torch.uint8, device=rec.device)
        dist.all_gather_into_tensor(recs.view(-1), rec, group=group)
    else:       # (gloo: the CPU tests, several ranks on one GPU)
        parts = [torch.empty_like(rec) for _ in range(world)]
        dist.all_gather(parts, rec, group=group)
        recs = torch.stack(parts)
    row = lambda r, k: recs[r, 2 * P * k:2 * P * (k + 1)].view(first_row_local.dtype)
    dec, sta = recs[:, 4 * P:5 * P], recs[:, 5 * P:6 * P]

    def last_decided(n, start):
        """status left behind by the last of the shards 0..n-1 that has a frame outside the buffer, else `start`."""
        if n == 0:
            return start
        order = torch.arange(1, n + 1, dtype=torch.int16, device=recs.device)[:, None]
        which = torch.where(dec[:n] != 0, order, torch.zeros_like(order)).amax(0)          # 0: none of them
        picked = sta[:n].gather(0, (which.long() - 1).clamp_(min=0)[None])[0]
        return torch.where(which > 0, picked, start)

    zero = torch.zeros_like(status_out)
    s_in = last_decided(rank, zero)
    s_end = last_decided(world, zero)
    return row(0, 0), (row(rank - 1, 1) if rank > 0 else None), s_in, s_end


def merge_pair_stats(st, final_status=None, group=None):
    """Two all-reduces over packed buffers bring the per-pair fields of `st` (an object with tensor attributes)
    to the values of a single run: SUM over (sum_k, sum_k2, n_int, n_enc), MAX over (last_f, decided, -first_f)
    with "never" mapped to the smallest integer.  The status after the last frame is the last shard's; it comes
    out of the summary scan (final_status) without another collective."""
    sums = torch.stack([st.sum_k, st.sum_k2, st.n_int.to(torch.int64), st.n_enc.to(torch.int64)])
    dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    st.sum_k.copy_(sums[0]); st.sum_k2.copy_(sums[1]); st.n_int.copy_(sums[2]); st.n_enc.copy_(sums[3])
    first = st.first_f
    neg_first = torch.where(first < 0, torch.full_like(first, _INT_MIN), -first)
    maxs = torch.stack([st.last_f, st.decided.to(torch.int32), neg_first])
    dist.all_reduce(maxs, op=dist.ReduceOp.MAX, group=group)
    st.last_f.copy_(maxs[0])
    st.decided.copy_(maxs[1])
    first.copy_(torch.where(maxs[2] == _INT_MIN, torch.full_like(first, -1), -maxs[2]))
    if final_status is not None:
        st.status.copy_(final_status)
    else:
        world = dist.get_world_size(group)
        src = dist.get_global_rank(group, world - 1) if group is not None else world - 1
        dist.broadcast(st.status, src=src, group=group)
    return st


def pair_stats_sharded(levels, P, nlevels, lo_cnt, hi_first, equal_cut, frame0, out=None, work=None, group=None,
                       local_pass=None, local_summary=None, merge_stream=None):
    """Stage 2 over this rank's frame shard, merged so that per-pair results equal a single run: one all-gather
    before the streaming pass, two all-reduces after it.

    Per-frame results stay local to the rank (frames are disjoint).  `local_pass` / `local_summary`
    exist for the CPU tests: callables with the signatures of stage2.pair_stats / shard_summary.
    merge_stream: a CUDA stream on which the merge after the pass runs (its packing kernels and the two
    all-reduces are small and latency-bound; a caller that goes straight on to the next shard's distances hides
    them behind that kernel).  The result then carries an event `merged`: wait for it before reading the per-pair
    fields or reusing `out`."""
    if local_pass is None:
        from . import stage2
        local_pass, local_summary = stage2.pair_stats, stage2.shard_summary
    decided, status = local_summary(levels, P, nlevels, lo_cnt, hi_first, equal_cut)
    first_row, prev_row, s_in, s_end = exchange_summaries(levels[0], levels[-1], decided, status, group)
    # rows travel without their padding; the streaming pass wants 16-byte aligned rows of the stack's pitch
    pad = lambda r: None if r is None else torch.nn.functional.pad(r.view(torch.int16), (0, levels.shape[1] - P)).view(levels.dtype)
    st = local_pass(levels, P, nlevels, lo_cnt, hi_first, equal_cut, frame0=frame0, first_row=pad(first_row),
                    prev_row=pad(prev_row), status_in=s_in, out=out, work=work)
    if merge_stream is None:
        return merge_pair_stats(st, s_end, group)
    passed = torch.cuda.Event()
    passed.record(torch.cuda.current_stream(levels.device))
    with torch.cuda.stream(merge_stream):
        merge_stream.wait_event(passed)
        merge_pair_stats(st, s_end, group)
        st.merged = torch.cuda.Event()
        st.merged.record(merge_stream)
    for t in (s_end, st.decided):
        t.record_stream(merge_stream)
    return st


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
