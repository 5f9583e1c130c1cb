"""Seeded synthetic trajectories of the benchmark configurations.

Residue centres: a compact self-avoiding walk (bond 0.38 nm) confined to a sphere at protein
density, so that roughly a tenth of all residue pairs come within 1 nm; eight heavy atoms per
residue scattered around the centre; dynamics = per-residue Ornstein-Uhlenbeck displacement
(sigma 0.1 nm, correlation 200 frames) plus a slow breathing mode and a small drift so that
contacts form and break.  Coordinates are rounded to 0.001 nm (xtc precision 1000).
"""
import numpy as np


def _globule(rng, nres, density=7.0, bond=0.38, min_dist=0.30):
    R = (3.0 * nres / (4.0 * np.pi * density)) ** (1.0 / 3.0)
    pos = np.zeros((nres, 3))
    cell = min_dist
    grid = {}

    def key(p):
        return tuple(np.floor(p / cell).astype(int))

    def clash(p):
        k = key(p)
        for dx in (-1, 0, 1):
            for dy in (-1, 0, 1):
                for dz in (-1, 0, 1):
                    for q in grid.get((k[0] + dx, k[1] + dy, k[2] + dz), ()):
                        if np.sum((pos[q] - p) ** 2) < min_dist * min_dist:
                            return True
        return False

    grid.setdefault(key(pos[0]), []).append(0)
    for i in range(1, nres):
        for attempt in range(2000):
            v = rng.normal(size=3)
            p = pos[i - 1] + bond * v / np.linalg.norm(v)
            if np.sum(p * p) <= R * R and (attempt > 1000 or not clash(p)):
                break
        else:
            p = pos[i - 1] * 0.9
        pos[i] = p
        grid.setdefault(key(p), []).append(i)
    return pos, R


def make_trajectory(nres, nframes, atoms_per_res=8, seed=3, two_chain=None, dtype=np.float32, closer=0.0,
                    frame_range=None, mode_period=None):
    """Returns (xyz [nframes, natoms, 3] float32 nm, res_start [nres+1] int32, box_length nm).

    two_chain=(nx, ny): two globules of nx and ny residues whose separation follows an OU process
    around 1.5 nm surface distance (the asymmetric interface configuration); closer (nm) moves the second
    globule towards the first, for small systems that should share more than a handful of contacts.
    frame_range=(f0, f1): only frames f0..f1-1 of the nframes-long trajectory are returned (a frame shard of
    one long run: the random stream is advanced through the frames before it, nothing else is kept).
    mode_period: the slow global modes (breathing, drift) repeat every mode_period frames instead of spanning the
    whole run, so that equal frame shards of a long run hold the same amount of work (weak scaling); None = nframes."""
    rng = np.random.default_rng(seed)
    if two_chain is None:
        centres, R = _globule(rng, nres)
        shift = None
    else:
        nx, ny = two_chain
        assert nx + ny == nres
        c1, R1 = _globule(rng, nx)
        c2, R2 = _globule(rng, ny)
        centres = np.concatenate([c1, c2 + np.array([R1 + R2 + 0.3 - closer, 0.0, 0.0])])
        R = R1 + R2 + 1.0
        shift = (nx, ny)
    offsets = rng.normal(0.0, 0.15, size=(nres, atoms_per_res, 3))
    f_lo, f_hi = (0, nframes) if frame_range is None else (int(frame_range[0]), int(frame_range[1]))
    assert 0 <= f_lo < f_hi <= nframes
    nkeep = f_hi - f_lo
    a = np.exp(-1.0 / 200.0)
    disp = np.empty((nkeep, nres, 3), dtype=np.float32)
    state = rng.normal(0.0, 0.1, size=(nres, 3))
    s = 0.1 * np.sqrt(1.0 - a * a)
    block = 2048
    for f0 in range(0, nframes, block):
        if f0 >= f_hi and shift is None:
            break                      # nothing after the shard draws from the stream in the one-chain system
        n = min(block, nframes - f0)
        noise = rng.normal(0.0, s, size=(n, nres, 3))
        if f0 >= f_hi:
            continue
        for k in range(n):
            state = a * state + noise[k]
            if f_lo <= f0 + k < f_hi:
                disp[f0 + k - f_lo] = state
    t = np.arange(f_lo, f_hi, dtype=np.float64)
    span = float(nframes if mode_period is None else mode_period)
    if mode_period is not None:
        t = np.mod(t, span)
    breathing = 1.0 + 0.04 * np.sin(2.0 * np.pi * t / max(span / 3.0, 50.0)) + 0.02 * t / max(span, 1.0)
    cen = centres[None, :, :] * breathing[:, None, None] + disp
    if shift is not None:
        nx, ny = shift
        sep = np.empty(nframes)
        v = 0.0
        b = np.exp(-1.0 / 100.0)
        for f in range(nframes):
            v = b * v + rng.normal(0.0, 0.35 * np.sqrt(1 - b * b))
            sep[f] = v
        cen[:, nx:, 0] += sep[f_lo:f_hi, None]
    nframes = nkeep
    L = float(4.0 * R + 4.0)
    # assembled in blocks of frames: the float64 intermediate of a 12 500-frame shard would be 2.4 GB
    xyz = np.empty((nframes, nres * atoms_per_res, 3), dtype=dtype)
    for f0 in range(0, nframes, 1024):
        blk = (cen[f0:f0 + 1024, :, None, :] + offsets[None, :, :, :]).reshape(-1, nres * atoms_per_res, 3) + L / 2.0
        xyz[f0:f0 + 1024] = np.round(blk * 1000.0) / 1000.0
    res_start = (np.arange(nres + 1) * atoms_per_res).astype(np.int32)
    return xyz, res_start, L


def make_protein_like(nres, nframes, seed=7):
    """Variable residue sizes (4..14 atoms) as in a real heavy-atom selection; small test systems."""
    rng = np.random.default_rng(seed)
    sizes = rng.integers(4, 15, size=nres)
    centres, R = _globule(rng, nres)
    res_start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    natoms = int(res_start[-1])
    offsets = rng.normal(0.0, 0.15, size=(natoms, 3))
    res_of = np.repeat(np.arange(nres), sizes)
    walk = np.cumsum(rng.normal(0.0, 0.03, size=(nframes, nres, 3)), axis=0)
    L = float(4.0 * R + 4.0)
    xyz = centres[res_of][None] + walk[:, res_of, :] + offsets[None] + rng.normal(0, 0.01, size=(nframes, natoms, 3)) + L / 2
    xyz = np.round(xyz * 1000.0) / 1000.0
    return xyz.astype(np.float32), res_start, L
