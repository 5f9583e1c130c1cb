"""Trajectories other than .xtc for stage 1: GROMACS .trr (full precision), multi-frame .gro and multi-model .pdb.

The reference accepts whatever `gmx mdmat -f` reads (docs/index.md:143-149).  The compressed .xtc stream is
decoded on the GPU (xtc.py, conan_xtc_decode_device); these formats carry plain numbers, so they are read into
float32 arrays on the host and copied to the device from pinned memory by mdmat_run.  All readers expose what
mdmat_run needs: natoms, times (ps), boxes [nf, 3, 3] (nm) and coords(frame_ids, atom_sel).
"""
import struct

import numpy as np


class CoordTrajectory:
    """Frames held as float32 [nf, natoms, 3] in host memory."""

    def __init__(self, xyz, times, boxes):
        self.xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        self.nframes, self.natoms = self.xyz.shape[0], self.xyz.shape[1]
        self.times = np.asarray(times, dtype=np.float32)
        self.boxes = np.asarray(boxes, dtype=np.float32).reshape(self.nframes, 3, 3)

    def coords(self, frame_ids, atom_sel=None):
        x = self.xyz[np.asarray(frame_ids, dtype=np.int64)]
        return np.ascontiguousarray(x if atom_sel is None else x[:, np.asarray(atom_sel, dtype=np.int64)])


def read_trr(path):
    """GROMACS .trr: XDR records of header (magic 1993, version string, 13 sizes, natoms, step, nre, t, lambda) then
    box, virial, pressure, positions, velocities, forces, each present if its size is non-zero, in float or
    double as the sizes say.  Only box and positions are kept."""
    data = open(path, "rb").read()
    pos, xyz, times, boxes = 0, [], [], []
    natoms0 = None
    while pos + 8 <= len(data):
        magic, slen = struct.unpack_from(">ii", data, pos)
        if magic != 1993:
            raise ValueError("%s: not a .trr frame at byte %d (magic %d)" % (path, pos, magic))
        pos += 8
        nstr = struct.unpack_from(">i", data, pos)[0]          # XDR string: length, bytes padded to 4
        pos += 4 + ((nstr + 3) // 4) * 4
        sizes = struct.unpack_from(">13i", data, pos)
        pos += 52
        ir, e, box, vir, pres, top, sym, x, v, f, natoms, step, nre = sizes
        if natoms <= 0:
            raise ValueError("%s: frame with %d atoms" % (path, natoms))
        real = 8 if (box and box == 72) or (x and x == natoms * 24) or (v and v == natoms * 24) or (f and f == natoms * 24) else 4
        fmt = ">d" if real == 8 else ">f"
        t, lam = struct.unpack_from(fmt + fmt[1], data, pos)
        pos += 2 * real
        dt = np.dtype(">f8" if real == 8 else ">f4")
        b = np.zeros((3, 3), np.float32)
        if box:
            b = np.frombuffer(data, dt, 9, pos).reshape(3, 3).astype(np.float32)
            pos += box
        pos += vir + pres
        if x:
            if natoms0 is None:
                natoms0 = natoms
            if natoms != natoms0:
                raise ValueError("%s: the number of atoms changes between frames" % path)
            xyz.append(np.frombuffer(data, dt, natoms * 3, pos).reshape(natoms, 3).astype(np.float32))
            times.append(t)
            boxes.append(b)
            pos += x
        pos += v + f
    if not xyz:
        raise ValueError("%s holds no positions" % path)
    return CoordTrajectory(np.stack(xyz), times, np.stack(boxes))


def read_gro_frames(path):
    """Multi-frame .gro: title (with `t= <ps>` if written by gmx), natoms, atom lines (positions in columns 21-44
    of width 8, or free format with more decimals), box line."""
    xyz, times, boxes = [], [], []
    with open(path) as f:
        lines = f.read().splitlines()
    i = 0
    while i + 2 < len(lines) and lines[i + 1].strip():
        title = lines[i]
        n = int(lines[i + 1])
        body = lines[i + 2:i + 2 + n]
        # the width of a number: distance between the decimal points of x and y
        first = body[0][20:]
        d1 = first.find(".")
        d2 = first.find(".", d1 + 1)
        w = d2 - d1 if d1 >= 0 and d2 > d1 else 8
        xyz.append(np.array([[float(l[20 + k * w:20 + (k + 1) * w]) for k in range(3)] for l in body], dtype=np.float32))
        bl = [float(v) for v in lines[i + 2 + n].split()]
        b = np.zeros((3, 3), np.float32)
        b[0, 0], b[1, 1], b[2, 2] = bl[0], bl[1], bl[2]
        if len(bl) == 9:
            b[0, 1], b[0, 2], b[1, 0], b[1, 2], b[2, 0], b[2, 1] = bl[3:9]
        boxes.append(b)
        t = float(len(times))
        if "t=" in title:
            try:
                t = float(title.split("t=")[1].split()[0])
            except (IndexError, ValueError):
                pass
        times.append(t)
        i += n + 3
    if not xyz:
        raise ValueError("%s holds no frames" % path)
    return CoordTrajectory(np.stack(xyz), times, np.stack(boxes))


def read_pdb_frames(path):
    """Multi-model .pdb: ATOM / HETATM records between MODEL / ENDMDL (or one unnamed model), Angstrom -> nm;
    CRYST1 gives a rectangular box when its angles are 90 degrees."""
    xyz, boxes, cur = [], [], []
    box = np.zeros((3, 3), np.float32)
    with open(path) as f:
        for line in f:
            rec = line[:6]
            if rec in ("ATOM  ", "HETATM"):
                cur.append((float(line[30:38]), float(line[38:46]), float(line[46:54])))
            elif rec == "CRYST1":
                a, b, c, al, be, ga = (float(line[6:15]), float(line[15:24]), float(line[24:33]),
                                       float(line[33:40]), float(line[40:47]), float(line[47:54]))
                box = np.zeros((3, 3), np.float32)
                if abs(al - 90) < 1e-3 and abs(be - 90) < 1e-3 and abs(ga - 90) < 1e-3:
                    box[0, 0], box[1, 1], box[2, 2] = 0.1 * a, 0.1 * b, 0.1 * c
                else:       # mark as triclinic: mdmat_run refuses the minimum image for it
                    box[0, 0], box[1, 1], box[2, 2], box[1, 0] = 0.1 * a, 0.1 * b, 0.1 * c, 1.0
            elif rec.startswith("ENDMDL") or rec.startswith("END"):
                if cur:
                    xyz.append(np.array(cur, dtype=np.float32) * np.float32(0.1))
                    boxes.append(box.copy())
                    cur = []
    if cur:
        xyz.append(np.array(cur, dtype=np.float32) * np.float32(0.1))
        boxes.append(box.copy())
    if not xyz:
        raise ValueError("%s holds no atoms" % path)
    if len({len(x) for x in xyz}) != 1:
        raise ValueError("%s: the number of atoms changes between models" % path)
    return CoordTrajectory(np.stack(xyz), np.arange(len(xyz), dtype=np.float32), np.stack(boxes))


def open_trajectory(path):
    low = path.lower()
    if low.endswith(".trr"):
        return read_trr(path)
    if low.endswith(".gro"):
        return read_gro_frames(path)
    if low.endswith(".pdb") or low.endswith(".ent"):
        return read_pdb_frames(path)
    raise ValueError("unsupported trajectory %s (.xtc, .trr, .gro and .pdb are read)" % path)


def write_trr(path, xyz, boxes=None, times=None, double=False):
    """A minimal .trr writer (positions and box only), for tests and for users converting synthetic data."""
    xyz = np.asarray(xyz)
    nf, natoms = xyz.shape[0], xyz.shape[1]
    real, dt = (8, ">f8") if double else (4, ">f4")
    version = b"GMX_trn_file"
    with open(path, "wb") as f:
        for k in range(nf):
            f.write(struct.pack(">ii", 1993, len(version) + 1))
            f.write(struct.pack(">i", len(version)) + version + b"\0" * ((-len(version)) % 4))
            box = 9 * real if boxes is not None else 0
            f.write(struct.pack(">13i", 0, 0, box, 0, 0, 0, 0, natoms * 3 * real, 0, 0, natoms, k, 0))
            t = float(times[k]) if times is not None else float(k)
            f.write(struct.pack(">dd" if double else ">ff", t, 0.0))
            if boxes is not None:
                f.write(np.asarray(boxes[k], dtype=dt).tobytes())
            f.write(np.asarray(xyz[k], dtype=dt).tobytes())
