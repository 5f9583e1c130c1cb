"""Host-side .xtc access through the C-ABI (conan_xtc_index / conan_xtc_decode / conan_xtc_encode).

Takes the place of the trajectory reading `gmx mdmat -f TRAJ` does for the reference
(conan.py:1437-1453).  The whole file is mapped once; frames are indexed by one header scan and
decoded on host threads, writing only the index group's atoms.
"""
import ctypes
import mmap
import os

import numpy as np

from . import _lib


class XtcFile:
    """An indexed .xtc: `times` (ps), `steps`, `boxes` [nf,3,3] and threaded frame decode."""

    def __init__(self, path_or_bytes):
        if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
            self._buf = np.frombuffer(bytes(path_or_bytes), dtype=np.uint8)
            self._mm = None
        else:
            f = open(path_or_bytes, "rb")
            size = os.fstat(f.fileno()).st_size
            if size == 0:
                raise _lib.ConanError(-3, "empty trajectory file %s" % path_or_bytes)
            self._mm = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
            f.close()
            self._buf = np.frombuffer(self._mm, dtype=np.uint8)
        lib = _lib.load()
        nf = ctypes.c_int64()
        na = ctypes.c_int32()
        ptr = self._buf.ctypes.data
        _lib.check(lib.conan_xtc_index(ptr, self._buf.size, 0, None, None, None, None,
                                       ctypes.byref(nf), ctypes.byref(na)))
        self.nframes = int(nf.value)
        self.natoms = int(na.value)
        self.offsets = np.zeros(self.nframes, np.int64)
        self.steps = np.zeros(self.nframes, np.int32)
        self.times = np.zeros(self.nframes, np.float32)
        self.boxes = np.zeros((self.nframes, 3, 3), np.float32)
        if self.nframes:
            _lib.check(lib.conan_xtc_index(ptr, self._buf.size, self.nframes, self.offsets.ctypes.data,
                                           self.steps.ctypes.data, self.times.ctypes.data,
                                           self.boxes.ctypes.data, ctypes.byref(nf), ctypes.byref(na)))

    def decode(self, frame_ids=None, atom_sel=None, out=None, nthreads=None):
        """Decode frames into float32 [n, nsel, 3] (nm); `out` may be a pinned buffer's numpy view."""
        lib = _lib.load()
        if frame_ids is None:
            frame_ids = np.arange(self.nframes, dtype=np.int64)
        frame_ids = np.ascontiguousarray(frame_ids, dtype=np.int64)
        if frame_ids.size and (frame_ids.min() < 0 or frame_ids.max() >= self.nframes):
            raise _lib.ConanError(-1, "frame index out of range")
        sel_ptr, nsel = None, 0
        nout = self.natoms
        if atom_sel is not None:
            atom_sel = np.ascontiguousarray(atom_sel, dtype=np.int32)
            sel_ptr, nsel, nout = atom_sel.ctypes.data, atom_sel.size, atom_sel.size
        if out is None:
            out = np.empty((frame_ids.size, nout, 3), np.float32)
        assert out.dtype == np.float32 and out.flags.c_contiguous and out.size == frame_ids.size * nout * 3
        if nthreads is None:
            nthreads = min(32, os.cpu_count() or 1)
        if frame_ids.size:
            _lib.check(lib.conan_xtc_decode(self._buf.ctypes.data, self._buf.size, self.offsets.ctypes.data,
                                            frame_ids.ctypes.data, frame_ids.size, self.natoms, sel_ptr, nsel,
                                            out.ctypes.data, nthreads))
        return out

    def decode_device(self, frame_ids=None, atom_sel=None, device="cuda", out=None, stream=None, one_pass=False,
                      work_bytes=None):
        """Decode frames on the GPU: the compressed bytes of the frame range are copied to the device and decoded
        there, in two passes with scratch memory (conan_xtc_decode_device_ws; work_bytes limits the scratch, the
        frames are then taken in rounds) or, one_pass=True, by one warp per frame (conan_xtc_decode_device).
        Returns a float32 device tensor [n, nsel, 3]; raises ConanError if any frame is malformed (this call
        synchronises to check)."""
        import torch
        lib = _lib.load()
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.ConanError(-1, "decode_device needs a CUDA device")
        if frame_ids is None:
            frame_ids = np.arange(self.nframes, dtype=np.int64)
        frame_ids = np.ascontiguousarray(frame_ids, dtype=np.int64)
        if frame_ids.size and (frame_ids.min() < 0 or frame_ids.max() >= self.nframes):
            raise _lib.ConanError(-1, "frame index out of range")
        nout = self.natoms if atom_sel is None else len(atom_sel)
        if out is None:
            out = torch.empty((frame_ids.size, nout, 3), dtype=torch.float32, device=dev)
        assert out.is_cuda and out.dtype == torch.float32 and out.is_contiguous() and out.numel() == frame_ids.size * nout * 3
        if frame_ids.size == 0:
            return out
        # the byte range that holds the requested frames (frames are 4-byte multiples, so it stays aligned)
        lo, last = int(frame_ids.min()), int(frame_ids.max())
        b0 = int(self.offsets[lo])
        b1 = int(self.offsets[last + 1]) if last + 1 < self.nframes else int(self._buf.size)
        d_data = torch.from_numpy(np.array(self._buf[b0:b1])).to(dev)      # (the mapping itself is read-only)
        d_off = torch.from_numpy(self.offsets[frame_ids] - b0).to(dev)
        d_slot = None
        if atom_sel is not None:
            slot = np.full(self.natoms, -1, np.int32)
            slot[np.asarray(atom_sel, dtype=np.int64)] = np.arange(nout, dtype=np.int32)
            d_slot = torch.from_numpy(slot).to(dev)
        d_status = torch.full((frame_ids.size,), -1, dtype=torch.int32, device=dev)
        st = stream if stream is not None else torch.cuda.current_stream(dev).cuda_stream
        with torch.cuda.device(dev):
            if one_pass:
                _lib.check(lib.conan_xtc_decode_device(d_data.data_ptr(), b1 - b0, d_off.data_ptr(), frame_ids.size,
                                                       self.natoms, None if d_slot is None else d_slot.data_ptr(), nout,
                                                       out.data_ptr(), d_status.data_ptr(), st))
            else:
                need = int(lib.conan_xtc_decode_workspace_bytes(frame_ids.size, self.natoms))
                if work_bytes is not None:
                    need = min(need, int(work_bytes))
                d_work = torch.empty(max(need, 16), dtype=torch.uint8, device=dev)
                _lib.check(lib.conan_xtc_decode_device_ws(d_data.data_ptr(), b1 - b0, d_off.data_ptr(), frame_ids.size,
                                                          self.natoms, None if d_slot is None else d_slot.data_ptr(), nout,
                                                          out.data_ptr(), d_status.data_ptr(), d_work.data_ptr(),
                                                          d_work.numel(), st))
        bad = torch.nonzero(d_status).flatten()
        if bad.numel():
            k = int(bad[0].item())
            raise _lib.ConanError(-3, "frame %d of the trajectory is malformed (decoder status %d)"
                                  % (int(frame_ids[k]), int(d_status[k].item())))
        return out


def encode_xtc(xyz, boxes=None, times=None, steps=None, precision=1000.0):
    """float32 [nf, natoms, 3] coordinates (nm) -> the bytes of an .xtc trajectory (uint8 array)."""
    lib = _lib.load()
    xyz = np.ascontiguousarray(xyz, dtype=np.float32)
    nf, na = xyz.shape[0], xyz.shape[1]
    boxes = None if boxes is None else np.ascontiguousarray(boxes, dtype=np.float32).reshape(nf, 9)
    times = None if times is None else np.ascontiguousarray(times, dtype=np.float32)
    steps = None if steps is None else np.ascontiguousarray(steps, dtype=np.int32)
    need = ctypes.c_size_t()
    args = (xyz.ctypes.data, None if boxes is None else boxes.ctypes.data,
            None if steps is None else steps.ctypes.data, None if times is None else times.ctypes.data,
            nf, na, precision)
    _lib.check(lib.conan_xtc_encode(*args, None, 0, ctypes.byref(need)))
    buf = np.empty(need.value, np.uint8)
    _lib.check(lib.conan_xtc_encode(*args, buf.ctypes.data, buf.size, ctypes.byref(need)))
    return buf


def write_xtc(path, xyz, boxes=None, times=None, steps=None, precision=1000.0):
    """Write float32 [nf, natoms, 3] coordinates (nm) as an .xtc trajectory; returns the file size."""
    buf = encode_xtc(xyz, boxes, times, steps, precision)
    with open(path, "wb") as f:
        f.write(buf.tobytes())
    return int(buf.size)
