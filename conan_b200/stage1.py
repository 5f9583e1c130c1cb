"""Stage 1 driver: coordinates -> uint16 level stack on the GPU (conan_mdmat_frames).

Stands where the reference runs `gmx mdmat -frames` (conan.py:1437-1453, :1783-1784).  torch is
used for device memory and streams only; the arithmetic is in conan_b200/csrc/mdmat.cu.
"""
import numpy as np
import torch

from . import _lib
from .levels import MAX_NLEVELS, n_pairs, padded_stride


class Geometry:
    """Which residue pairs are computed: all i<j (symmetric) or an X x Y block (conan.py:541-551)."""

    def __init__(self, res_start, x0=None, xres=None, y0=None, yres=None):
        self.res_start = np.ascontiguousarray(res_start, dtype=np.int32)
        self.nres = len(self.res_start) - 1
        self.natoms = int(self.res_start[-1])
        self.max_res_atoms = int(np.diff(self.res_start).max())
        self.symmetric = xres is None
        if self.symmetric:
            self.x0, self.xres, self.y0, self.yres = 0, self.nres, 0, self.nres
            self.P = n_pairs(self.nres)
        else:
            self.x0, self.xres, self.y0, self.yres = int(x0), int(xres), int(y0), int(yres)
            self.P = n_pairs(xres=self.xres, yres=self.yres)
        self.stride = padded_stride(self.P)


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def mdmat_workspace_bytes(nf, geom):
    n = _lib.load().conan_mdmat_workspace_bytes(nf, geom.natoms, geom.nres, geom.max_res_atoms)
    if n == 0:
        raise _lib.ConanError(-1, "conan_mdmat_workspace_bytes rejected its arguments")
    return n


def mdmat_frames(d_xyz, d_box, geom, trunc, nlevels, pbc, d_levels, d_work, d_counters=None, stream=None,
                 phases=3):
    """Launch stage 1 for the frames in d_xyz ([nf, natoms, 3] float32, device) into d_levels.

    d_levels is a uint16 device tensor [nf, geom.stride]; d_counters an int64[3] device tensor
    (near-boundary results, un-pruned residue pairs, atom pairs evaluated) that is added to."""
    lib = _lib.load()
    if not d_xyz.is_cuda:
        raise _lib.ConanError(-1, "stage 1 needs device tensors: there is no CPU fallback")
    if nlevels > MAX_NLEVELS:
        raise _lib.ConanError(-4, "NLEVEL %d exceeds %d" % (nlevels, MAX_NLEVELS))
    nf = d_xyz.shape[0]
    assert d_xyz.dtype == torch.float32 and d_xyz.is_contiguous() and d_xyz.shape[1] == geom.natoms
    assert d_levels.is_contiguous() and d_levels.shape[0] >= nf and d_levels.shape[1] == geom.stride
    s = stream if stream is not None else torch.cuda.current_stream(d_xyz.device)
    _lib.check(lib.conan_mdmat_frames_ex(
        d_xyz.data_ptr(), _ptr(d_box), nf, geom.natoms, geom.res_start.ctypes.data, geom.nres,
        1 if geom.symmetric else 0, geom.x0, geom.xres, geom.y0, geom.yres, float(trunc), int(nlevels),
        1 if pbc else 0, d_levels.data_ptr(), geom.stride, _ptr(d_counters), d_work.data_ptr(),
        d_work.numel() * d_work.element_size(), s.cuda_stream, phases))
    return d_levels


STATIC_SCRATCH_BYTES = 32 * 1024 * 1024   # CONAN_MDMAT_STATIC_SCRATCH_BYTES


def mdmat_frames_static(d_xyz, d_box, geom, trunc, nlevels, pbc, d_levels, d_counters=None, stream=None):
    """Stage 1 through conan_mdmat_frames_static: no workspace tensor, nothing allocated.

    For hosts that must not allocate device memory while they run (an analysis hook inside a
    simulation's step loop); limited to the frames whose scratch fits STATIC_SCRATCH_BYTES."""
    lib = _lib.load()
    if not d_xyz.is_cuda:
        raise _lib.ConanError(-1, "stage 1 needs device tensors: there is no CPU fallback")
    nf = d_xyz.shape[0]
    s = stream if stream is not None else torch.cuda.current_stream(d_xyz.device)
    _lib.check(lib.conan_mdmat_frames_static(
        d_xyz.data_ptr(), _ptr(d_box), nf, geom.natoms, geom.res_start.ctypes.data, geom.nres,
        1 if geom.symmetric else 0, geom.x0, geom.xres, geom.y0, geom.yres, float(trunc), int(nlevels),
        1 if pbc else 0, d_levels.data_ptr(), geom.stride, _ptr(d_counters), s.cuda_stream))
    return d_levels


def compute_levels(xyz, boxes, geom, trunc, nlevels, pbc=False, device="cuda", batch=4096, out=None):
    """Host coordinates [nf, natoms, 3] -> device level stack [nf, stride] uint16, batched H2D."""
    xyz = np.ascontiguousarray(xyz, dtype=np.float32)
    nf = xyz.shape[0]
    dev = torch.device(device)
    levels = out if out is not None else torch.zeros((nf, geom.stride), dtype=torch.uint16, device=dev)
    nb = torch.zeros(3, dtype=torch.int64, device=dev)
    batch = max(1, min(batch, nf))
    work = torch.empty(mdmat_workspace_bytes(batch, geom), dtype=torch.uint8, device=dev)
    for b0 in range(0, nf, batch):
        b1 = min(nf, b0 + batch)
        d_xyz = torch.from_numpy(xyz[b0:b1]).to(dev, non_blocking=False)
        d_box = None
        if pbc:
            d_box = torch.from_numpy(np.ascontiguousarray(boxes[b0:b1], dtype=np.float32).reshape(-1, 9)).to(dev)
        mdmat_frames(d_xyz, d_box, geom, trunc, nlevels, pbc, levels[b0:b1], work, nb)
    torch.cuda.synchronize(dev)
    return levels, int(nb[0].item())
