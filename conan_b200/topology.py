"""Topology reading and GROMACS-style default index groups for stage 1.

`gmx mdmat -s COORD` (conan.py:1437-1453) needs two things from the structure file: which atoms
belong to the chosen index group (the reference's tests pipe "2" = Protein-H on stdin,
test/test.py:46) and where residues start (a new residue begins whenever the residue of
consecutive group atoms changes).  This module provides both for .gro / .pdb files, for the
GROMACS .tpr files the reference ships (heuristic scan of the first molecule type), and for
.ndx index files.
"""
import struct

import numpy as np

# Residue names GROMACS' residuetypes.dat classifies as "Protein" (common force-field variants).
PROTEIN_RESIDUES = set("""
ABU ACE AIB ALA ARG ARGN ASN ASN1 ASP ASP1 ASPH ASPP ASH CT3 CYS CYS1 CYS2 CYSH DALA GLN GLU GLUH
GLUP GLH GLY HIS HIS1 HISA HISB HISH HISD HISE HISP HSD HSE HSP HYP ILE LEU LSN LYS LYSH MELEU MET
MEVAL NAC NME NHE NH2 ORN PHE PHEH PHEU PHL PRO SER THR TRP TRPH TRPU TYR TYRH TYRU VAL PGLU HID
HIE HIP LYP LYN CYN CYM CYX DAB ASF GLF PHO QLN PTR SEP TPO M3L CME MSE
""".split())
for _aa in ("ALA ARG ASN ASP CYS CYX GLN GLU GLY HID HIE HIP HIS ILE LEU LYS MET PHE PRO SER THR "
            "TRP TYR VAL").split():
    PROTEIN_RESIDUES.add("N" + _aa)
    PROTEIN_RESIDUES.add("C" + _aa)

_BACKBONE = ("N", "CA", "C")
_MAINCHAIN = ("N", "CA", "C", "O", "O1", "O2", "OC1", "OC2", "OT", "OXT")
_MAINCHAIN_CB = _MAINCHAIN + ("CB",)
_MAINCHAIN_H = _MAINCHAIN + ("H1", "H2", "H3", "H", "HN")


class Topology:
    """Per-atom names / residue names / residue ids, in file order."""

    def __init__(self, names, resnames, resids, coords=None, box=None):
        self.names = list(names)
        self.resnames = list(resnames)
        self.resids = np.asarray(resids, dtype=np.int64)   # a key that changes when the residue changes
        self.coords = coords
        self.box = box
        self.natoms = len(self.names)

    # ---- GROMACS default groups -------------------------------------------------------------
    def _is_protein(self):
        return np.array([r.upper() in PROTEIN_RESIDUES for r in self.resnames], dtype=bool)

    @staticmethod
    def _strip_digits(name):
        i = 0
        while i < len(name) and name[i].isdigit():
            i += 1
        return name[i:]

    def default_groups(self):
        """Ordered list of (name, int32 atom indices) following `gmx make_ndx` defaults 0..11."""
        prot = self._is_protein()
        base = [self._strip_digits(n).upper() for n in self.names]
        whole = [n.upper() for n in self.names]
        is_h = np.array([b.startswith("H") for b in base], dtype=bool)
        idx = np.arange(self.natoms)

        def whole_match(names):
            s = set(names)
            return np.array([w in s for w in whole], dtype=bool)

        mch = whole_match(_MAINCHAIN_H)
        groups = [
            ("System", idx),
            ("Protein", idx[prot]),
            ("Protein-H", idx[prot & ~is_h]),
            ("C-alpha", idx[prot & whole_match(("CA",))]),
            ("Backbone", idx[prot & whole_match(_BACKBONE)]),
            ("MainChain", idx[prot & whole_match(_MAINCHAIN)]),
            ("MainChain+Cb", idx[prot & whole_match(_MAINCHAIN_CB)]),
            ("MainChain+H", idx[prot & mch]),
            ("SideChain", idx[prot & ~mch]),
            ("SideChain-H", idx[prot & ~mch & ~is_h]),
            ("Prot-Masses", idx[prot]),
            ("non-Protein", idx[~prot]),
        ]
        return [(n, g.astype(np.int32)) for n, g in groups]

    def residue_runs(self, sel):
        """res_start [nres+1] for the selected atoms: a new residue whenever the residue changes."""
        sel = np.asarray(sel, dtype=np.int64)
        if sel.size == 0:
            raise ValueError("empty atom selection")
        rid = self.resids[sel]
        change = np.flatnonzero(rid[1:] != rid[:-1]) + 1
        return np.concatenate([[0], change, [sel.size]]).astype(np.int32)

    def truncated(self, natoms):
        """First `natoms` atoms (gmx uses the leading part of a topology for a shorter trajectory)."""
        return Topology(self.names[:natoms], self.resnames[:natoms], self.resids[:natoms])


def read_gro(path):
    with open(path) as f:
        lines = f.read().splitlines()
    n = int(lines[1].split()[0])
    names, resnames, resids, coords = [], [], [], []
    run, last = -1, None
    for line in lines[2:2 + n]:
        key = (line[0:5], line[5:10])
        if key != last:
            run += 1
            last = key
        resnames.append(line[5:10].strip())
        names.append(line[10:15].strip())
        resids.append(run)
        coords.append((float(line[20:28]), float(line[28:36]), float(line[36:44])))
    box = [float(v) for v in lines[2 + n].split()] if len(lines) > 2 + n else None
    return Topology(names, resnames, resids, np.array(coords, np.float32), box)


def read_pdb(path):
    names, resnames, resids, coords = [], [], [], []
    run, last = -1, None
    with open(path) as f:
        for line in f:
            if line.startswith("ENDMDL"):
                break
            if not (line.startswith("ATOM") or line.startswith("HETATM")):
                continue
            key = (line[17:21], line[21], line[22:27])
            if key != last:
                run += 1
                last = key
            names.append(line[12:16].strip())
            resnames.append(line[17:21].strip())
            resids.append(run)
            coords.append((0.1 * float(line[30:38]), 0.1 * float(line[38:46]), 0.1 * float(line[46:54])))
    if not names:
        raise ValueError("no ATOM records in %s" % path)
    return Topology(names, resnames, resids, np.array(coords, np.float32))


def read_tpr(path):
    """Heuristic reader for GROMACS .tpr run inputs (tested on tpx versions 83 and 100).

    Finds the symbol table (a count followed by that many XDR strings) and then the first block
    that parses as a molecule type's atoms: `nr, nres`, nr records of nine 4-byte fields
    (m, q, mB, qB, type, typeB, ptype, resind, atomnumber) with resind running 0..nres-1, nr atom-name
    symbol indices, 2*nr type-name indices and nres residue records.  Only that first molecule
    type is returned, which is what a protein-only trajectory needs.
    """
    with open(path, "rb") as f:
        d = f.read()
    n = len(d)

    def i32(o):
        return struct.unpack_from(">i", d, o)[0]

    # symbol table
    sym, sym_end = None, None
    for o in range(0, min(n - 8, 1 << 16), 4):
        cnt = i32(o)
        if cnt < 8 or cnt > 1000000:
            continue
        p, strs, ok = o + 4, [], True
        for _ in range(cnt):
            if p + 8 > n:
                ok = False
                break
            l1, l2 = i32(p), i32(p + 4)
            if l1 != l2 + 1 or l2 < 0 or l2 > 256 or p + 8 + l2 > n:
                ok = False
                break
            raw = d[p + 8:p + 8 + l2]
            if any(c < 32 or c > 126 for c in raw):
                ok = False
                break
            strs.append(raw.decode("ascii"))
            p += 8 + ((l2 + 3) // 4) * 4
        if ok:
            sym, sym_end = strs, p
            break
    if sym is None:
        raise ValueError("%s: could not locate the .tpr symbol table; give a .gro or .pdb as COORD" % path)
    nsym = len(sym)
    # first molecule type
    o = sym_end
    while o + 8 < n:
        nr, nres = i32(o), i32(o + 4)
        if 0 < nres <= nr < 50000000 and o + 8 + nr * 36 + nr * 12 + nres * 12 <= n:
            rec = np.frombuffer(d, dtype=">i4", count=nr * 9, offset=o + 8).reshape(nr, 9)
            resind = rec[:, 7]
            mass = np.frombuffer(d, dtype=">f4", count=nr * 9, offset=o + 8).reshape(nr, 9)[:, 0]
            step = np.diff(resind)
            if (resind[0] == 0 and resind[-1] == nres - 1 and (step >= 0).all() and (step <= 1).all()
                    and (mass >= 0).all() and (mass < 1000).all() and (rec[:, 8] >= 0).all() and (rec[:, 8] < 130).all()):
                p = o + 8 + nr * 36
                name_idx = np.frombuffer(d, dtype=">i4", count=nr, offset=p)
                p += nr * 12
                resrec = np.frombuffer(d, dtype=">i4", count=nres * 3, offset=p).reshape(nres, 3)
                if (name_idx >= 0).all() and (name_idx < nsym).all() and (resrec[:, 0] >= 0).all() and (resrec[:, 0] < nsym).all():
                    names = [sym[k] for k in name_idx]
                    resnames = [sym[resrec[r, 0]] for r in resind]
                    return Topology(names, resnames, resind.astype(np.int64))
        o += 4
    raise ValueError("%s: no molecule type found in the .tpr; give a .gro or .pdb as COORD" % path)


def read_topology(path):
    low = path.lower()
    if low.endswith(".gro"):
        return read_gro(path)
    if low.endswith(".pdb") or low.endswith(".ent"):
        return read_pdb(path)
    if low.endswith(".tpr"):
        return read_tpr(path)
    raise ValueError("unsupported COORD file %s (use .gro, .pdb or .tpr)" % path)


def read_ndx(path):
    """GROMACS index file -> ordered list of (name, 0-based int32 indices)."""
    groups, name, cur = [], None, []
    with open(path) as f:
        for line in f:
            s = line.strip()
            if s.startswith("[") and s.endswith("]"):
                if name is not None:
                    groups.append((name, np.array(cur, dtype=np.int32) - 1))
                name, cur = s[1:-1].strip(), []
            elif s:
                cur.extend(int(v) for v in s.split())
    if name is not None:
        groups.append((name, np.array(cur, dtype=np.int32) - 1))
    return groups


def write_gro(path, topo, coords, box=(0.0, 0.0, 0.0), title="conan_b200"):
    with open(path, "w") as f:
        f.write(title + "\n%5d\n" % topo.natoms)
        for a in range(topo.natoms):
            f.write("%5d%-5s%5s%5d%8.3f%8.3f%8.3f\n" % ((int(topo.resids[a]) + 1) % 100000, topo.resnames[a][:5],
                                                        topo.names[a][:5], (a + 1) % 100000,
                                                        coords[a][0], coords[a][1], coords[a][2]))
        f.write("%10.5f%10.5f%10.5f\n" % tuple(box))
