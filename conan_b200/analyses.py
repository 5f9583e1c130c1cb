"""Stage 3 of the reference's script body (conan.py:1887-2090) on the CUDA kernels.

Pearson correlation of every pair's distance series with time and with observables
(conan.py:1888-1971) comes from the streaming sums of conan_pearson_sums; the PCA covariance
(:1983) and the frame-frame contact-map distances (:2063) from the tcgen05 3xTF32 GEMM.  What the
reference leaves to scipy after those contractions -- the top-k eigenpairs (ARPACK there,
torch.linalg.eigh on the device here) and Ward linkage / fcluster (scipy in both) -- stays a
library call: small, and outside the hot path.
"""
import math
import os

import numpy as np

from . import stage3, tables


# ------------------------------------------------------------------------------------------------
# observables file (conan.py:52-79, :102-120)
# ------------------------------------------------------------------------------------------------
def read_in_observables(inputf, time_vec):
    """-> frame_indices, times, titles, obs_vec; rows whose int(time) is not a frame time are dropped."""
    lookup = {}
    for n, t in enumerate(time_vec):
        lookup.setdefault(t, n)
    frame_indices, times, titles, obs_vec = [], [], [], []
    ntitles = -1
    for line in inputf:
        f = line.split()
        if line.startswith("#") or not f:
            continue
        if ntitles == -1:
            ntitles = int(f[0])
        elif len(titles) < ntitles:
            titles.append(line[:-1])
        else:
            t = int(float(f[0]))
            if t in lookup:
                frame_indices.append(lookup[t])
                times.append(t)
                if not obs_vec:
                    obs_vec = [[] for _ in range(ntitles)]
                for n in range(ntitles):
                    obs_vec[n].append(float(f[n + 1]))
    return frame_indices, times, titles, obs_vec


def read_in_observables_blind(inputf):
    titles, obs_vec = [], []
    ntitles = -1
    for line in inputf:
        f = line.split()
        if line.startswith("#") or not f:
            continue
        if ntitles == -1:
            ntitles = int(f[0])
        elif len(titles) < ntitles:
            titles.append(line[:-1])
        else:
            if not obs_vec:
                obs_vec = [[] for _ in range(ntitles)]
            for n in range(ntitles):
                obs_vec[n].append(float(f[n + 1]))
    return titles, obs_vec


# ------------------------------------------------------------------------------------------------
# helpers over the stage-2 result
# ------------------------------------------------------------------------------------------------
class _Ctx:
    """Device stack + bookkeeping shared by the analyses."""

    def __init__(self, s2):
        self.s2 = s2
        self.d_stack = s2.d_stack
        self.layout = s2.layout
        self.P = s2.layout.P
        self.nlevels = s2.nlevels
        self.values = s2.values
        self.time_vec = np.asarray(s2.time_vec, dtype=np.float64)
        self.nf = len(self.time_vec)
        self.seen = np.asarray(s2.seen)
        self.lut = stage3.lut16(self.values, self.d_stack.device)
        self.lut_host = np.asarray(self.values).astype(np.float16)
        # per-pair shift: float16 value of the mean level keeps the moment sums free of cancellation
        mean_level = np.rint(s2.stats["sum_k"] / self.nf).astype(np.int64)
        self.shift = self.lut_host[np.clip(mean_level, 0, self.nlevels - 1)].astype(np.float32)

    def to_map(self, per_pair, diag=0.0):
        return self.layout.to_matrix(per_pair, diag=diag)

    def write_map(self, path, mat, fmt="%4d %4d %10.6f\n"):
        mat = np.asarray(mat)
        cols = [mat] if mat.ndim == 2 else [mat[:, :, c] for c in range(mat.shape[2])]
        tables.write_pair_table(path, self.layout, cols, fmt)

    def varying(self, frame_mask=None):
        """len(set(rvec)) > 1 for every pair, exactly: lowest and highest level differ as float16 values."""
        kmin, kmax = stage3.level_range(self.d_stack, self.P, frame_mask)
        ok = kmin <= kmax
        return ok & (self.lut_host[np.clip(kmin, 0, self.nlevels - 1)] != self.lut_host[np.clip(kmax, 0, self.nlevels - 1)])


def pearson_time(ctx):
    """aggregate/pearson_data.dat (conan.py:1888-1919)."""
    res = stage3.linregress_all(ctx.d_stack, ctx.P, ctx.nlevels, ctx.lut, ctx.time_vec[None], ctx.shift)
    listed = np.zeros(ctx.P, bool)
    listed[ctx.seen] = True
    r = np.where(listed & ctx.varying(), res["r"][0], 0.0)
    ctx.write_map("aggregate/pearson_data.dat", ctx.to_map(r, 0.0))
    return r


def pearson_obs(ctx, path, ignore_time=False):
    """aggregate/pearson_data_obs_N.dat: r, slope, stderr per pair and observable (conan.py:1922-1971)."""
    with open(path) as f:
        if ignore_time:
            titles, obs_vec = read_in_observables_blind(f)
            frame_indices = list(range(len(obs_vec[0]))) if obs_vec else []
        else:
            frame_indices, _, titles, obs_vec = read_in_observables(f, list(ctx.time_vec))
    if not obs_vec:
        print("No observable rows match the frame times.")
        return None
    obs = np.asarray(obs_vec, dtype=np.float64)
    if ignore_time and obs.shape[1] != ctx.nf:
        print("Pearson error! Wrong number of frames.", " len (rvec) = %i len(obs) = %i" % (ctx.nf, obs.shape[1]))
        return None
    # the reference indexes vecij[frame_indices]; duplicates are not expected in a time-ordered file
    mask = np.zeros(ctx.nf, bool)
    mask[frame_indices] = True
    full = np.zeros((obs.shape[0], ctx.nf))
    full[:, frame_indices] = obs
    res = stage3.linregress_all(ctx.d_stack, ctx.P, ctx.nlevels, ctx.lut, full, ctx.shift, frame_mask=None if mask.all() else mask)
    listed = np.zeros(ctx.P, bool)
    listed[ctx.seen] = True
    ok = listed & ctx.varying(None if mask.all() else mask)
    out = []
    for k in range(obs.shape[0]):
        trio = np.stack([np.where(ok, res[n][k], 0.0) for n in ("r", "slope", "stderr")], axis=-1)   # [P, 3]
        lay = ctx.layout
        mats = [lay.to_matrix(trio[:, c], diag=0.0) for c in range(3)]
        tables.write_pair_table("aggregate/pearson_data_obs_%d.dat" % (k + 1), lay, mats, "%4d %4d %10.6f %10.6f %10.6f\n")
        out.append(trio)
    return out


def _project_pc(ctx, sel, mean, v, change_sign):
    """project_pc (conan.py:430-452): projections of the float16 mean-free vectors on the columns of v (device,
    [nsel, n] float64), sign chosen so that a projection does not decrease with time; writes pca_projections.dat."""
    import torch
    from scipy.stats import linregress
    dev = ctx.d_stack.device
    npc = v.shape[1]
    d_sel = torch.from_numpy(sel.astype(np.int64)).to(dev)
    mean16 = torch.from_numpy(mean.astype(np.float32).astype(np.float16)).to(dev)
    lut16 = ctx.lut.to(torch.float16)
    proj = torch.empty((ctx.nf, npc), dtype=torch.float64, device=dev)
    step = max(1, int(2.0e8 // max(len(sel), 1)))
    for f0 in range(0, ctx.nf, step):
        lv = ctx.d_stack[f0:f0 + step].view(torch.int16).to(torch.int64) & 0xffff
        vm0 = (lut16[lv.index_select(1, d_sel)] - mean16[None, :]).double()      # float16 arithmetic, then widened
        proj[f0:f0 + step] = vm0 @ v
    proj = proj.cpu().numpy()
    v = v.cpu().numpy()
    print("Projecting principal components...")
    if change_sign:
        for i in range(npc):
            if linregress(proj[:, i], ctx.time_vec).rvalue < 0:
                v[:, i] = -v[:, i]
                proj[:, i] *= -1
    with open("pca/pca_projections.dat", "w") as f:
        for n in range(ctx.nf):
            f.write("%12i" % ctx.time_vec[n])
            f.write((npc * " %12.6f") % tuple(proj[n]) + "\n")
    return v, proj


def pca_read(ctx, pca_dir):
    """PCA_READ: project the trajectory on principal components written by an earlier run
    (pca/pc.N.dat of that run; conan.py:1073-1095, :2003-2010).  No sign change."""
    import torch
    os.makedirs("pca", exist_ok=True)
    sel = ctx.seen.astype(np.int32)
    print("There are a total of %i non-constant pairs." % len(sel))
    lay = ctx.layout
    pi, pj = lay.pair_labels()
    cols = []
    i = 1
    while os.path.isfile(os.path.join(pca_dir, "pc.%i.dat" % i)):
        rows = np.loadtxt(os.path.join(pca_dir, "pc.%i.dat" % i), ndmin=2)
        mat = np.zeros((lay.xres, lay.yres))
        ri, rj = rows[:, 0].astype(np.int64) - lay.xterm, rows[:, 1].astype(np.int64) - lay.yterm
        ok = (ri >= 0) & (ri < lay.xres) & (rj >= 0) & (rj < lay.yres)
        mat[ri[ok], rj[ok]] = rows[ok, 2]
        coeff = mat[pi[sel] - lay.xterm, pj[sel] - lay.yterm]
        print("Norm of PC # %i = %12.6f" % (i, math.sqrt(float((coeff ** 2).sum()))))
        cols.append(coeff)
        i += 1
    n = i - 1
    print("Reading in PC's.")
    print("I found %i principal components..." % n)
    if n == 0:
        return None
    sums = stage3.linregress_all(ctx.d_stack, ctx.P, ctx.nlevels, ctx.lut, ctx.time_vec[None], ctx.shift)
    v = torch.from_numpy(np.stack(cols, axis=1)).to(ctx.d_stack.device)
    return _project_pc(ctx, sel, sums["mean"][sel], v, change_sign=False)


def pca(ctx, npc):
    """pca/pca_output.txt, pc.N.dat, pca_projections.dat (conan.py:1973-2001, :413-452)."""
    import torch
    os.makedirs("pca", exist_ok=True)
    sel = ctx.seen.astype(np.int32)
    print("There are a total of %i non-constant pairs." % len(sel))
    print("Removing means....")
    # exact means of the float16 series (np.cov works in float64 on them)
    sums = stage3.linregress_all(ctx.d_stack, ctx.P, ctx.nlevels, ctx.lut, ctx.time_vec[None], ctx.shift)
    mean = sums["mean"][sel]
    print("Building covariance matrix....")
    cov = stage3.covariance(ctx.d_stack, sel, ctx.nlevels, ctx.lut, mean)
    print("Done.")
    sum_var = float(torch.diagonal(cov).double().sum().item())
    print("Diagonalizing....")
    w, v = stage3.top_eigenpairs(cov, npc)
    w = w.cpu().numpy()
    v, proj = _project_pc(ctx, sel, mean, v, change_sign=True)
    with open("pca/pca_output.txt", "w") as f:
        f.write("# Total variance: %f\n" % sum_var)
        f.write("# i (eigvalue i) (running sum to i) (explained variance)\n")
        partial = 0.0
        for i in range(npc):
            partial += w[i]
            f.write("%3i %12.6f %12.6f %12.6f\n" % (i + 1, w[i], partial, partial / sum_var))
            comp = np.zeros(ctx.P)
            comp[sel] = v[:, i]
            ctx.write_map("pca/pc.%i.dat" % (i + 1), ctx.to_map(comp, 0.0), fmt="%5i %5i %12.6f\n")
    return w, v, proj


def get_best_center(cluster, dist_rows):
    """Medoid: the member with the smallest summed distance to the others (conan.py:1172-1185)."""
    sub = dist_rows[np.ix_(cluster, cluster)]
    return int(cluster[int(np.argmin(sub.sum(axis=1, dtype=np.float64)))])


def cluster_trajectory(ctx, ks, trunc, trunc_inter, asymm):
    """cluster_trj/*: frame-frame distances, Ward clustering, medoids, per-cluster maps (conan.py:2034-2090, :1290-1412)."""
    import torch
    from scipy.cluster.hierarchy import fcluster, linkage
    os.makedirs("cluster_trj", exist_ok=True)
    lay = ctx.layout
    norm = np.sqrt(lay.xres * lay.yres) if asymm else lay.nres / np.sqrt(2.0)
    sel = ctx.seen.astype(np.int32)
    print("Building distance matrix...")
    sums = stage3.linregress_all(ctx.d_stack, ctx.P, ctx.nlevels, ctx.lut, ctx.time_vec[None], ctx.shift)
    D = stage3.frame_distances(ctx.d_stack, sel, ctx.nlevels, ctx.lut, sums["mean"][sel], norm)
    D.fill_diagonal_(0.0)
    print("Done.")
    nf = ctx.nf
    # The matrix stays float32 on the host (1.6 GB at 20 000 frames, not 3.2); only what scipy needs -- the condensed
    # upper triangle -- is widened, and sums over it are taken in float64.
    Dh = D.cpu().numpy()
    delta = int(nf / 1000) if nf > 1000 else 1
    if nf > 1000:
        print("There are more than 1000 frames, we will limit the interframe plots to approx. 1000x1000.")
    tn = ctx.time_vec * 0.001
    with open("cluster_trj/rmsd_interframe.dat", "w") as f:
        for i in range(0, nf, delta):
            f.write("".join("%8.3f %8.3f %12.6f\n" % (tn[i], tn[j], Dh[i, j]) for j in range(0, nf, delta)))
            f.write("\n")
    print("hierarchical...")
    from scipy.spatial.distance import squareform
    cond = squareform(Dh, checks=False).astype(np.float64)      # (the GPU result is symmetric, with a zero diagonal)
    Z = linkage(cond, 'ward')
    del cond
    with open("cluster_trj/linkage.txt", "w") as f:
        for row in Z:
            f.write("%6i %6i %12.6f %6i\n" % tuple(row))
    if ks == [0]:
        from .frontend import parse_numbers
        ks = parse_numbers(input("Number of clusters? (multiple values are allowed, or a single interval as i-j).\n"))
    listed = np.zeros(ctx.P, bool)
    listed[sel] = True
    ind_lut = stage3.lut16(ctx.values, D.device)
    # 1 where the float16 value is below trunc_inter (:1316; numpy compares the float16 with the cut-off as float16)
    ind_lut = (ind_lut < float(np.float16(trunc_inter))).float()
    zero_shift = np.zeros(ctx.P, np.float32)
    for k in ks:
        os.makedirs("cluster_trj/%i" % k, exist_ok=True)
        assign = fcluster(Z, k, criterion='maxclust')
        clusters = [np.flatnonzero(assign == c + 1) for c in range(k)]
        clusters.sort(key=len, reverse=True)
        centers = [get_best_center(c, Dh) if len(c) else 0 for c in clusters]
        cost = np.zeros(nf)
        for c, ctr in zip(clusters, centers):
            cost[c] = Dh[c, ctr]
        order = np.argsort(centers, kind="stable")
        clusters = [clusters[o] for o in order]
        centers = [centers[o] for o in order]
        # per-cluster mean distance and interaction lifetime of every pair: masked sums in one pass each
        for c0 in range(0, k, stage3.MAX_J):
            part = clusters[c0:c0 + stage3.MAX_J]
            masks = np.zeros((len(part), nf), np.float32)
            for n, c in enumerate(part):
                masks[n, c] = 1.0
            T, _ = stage3.pearson_sums(ctx.d_stack, ctx.P, ctx.nlevels, ctx.lut, masks, ctx.shift)
            L, _ = stage3.pearson_sums(ctx.d_stack, ctx.P, ctx.nlevels, ind_lut, masks, zero_shift)
            T, L = T.cpu().numpy(), L.cpu().numpy()
            for n, c in enumerate(part):
                cdir = "cluster_trj/%i/cluster%i" % (k, c0 + n)
                os.makedirs(cdir, exist_ok=True)
                cnt = max(len(c), 1)
                # np.mean of a float16 row returns float16 (:1319): round the mean accordingly
                mean_d = (ctx.shift.astype(np.float64) + T[n] / cnt).astype(np.float32).astype(np.float16).astype(np.float64)
                mean_d = np.where(listed, mean_d, trunc)
                life = np.where(listed, 100.0 * L[n] / cnt, 0.0)
                ctx.write_map(cdir + "/average_map.dat", ctx.to_map(mean_d, 0.0))
                ctx.write_map(cdir + "/lifetime.dat", ctx.to_map(life, 100.0))
        with open("cluster_trj/%i/assignment.dat" % k, "w") as f:
            f.write("# element_ind cluster_ind center? dist\n")
            lab = np.zeros(nf, int)
            for n, c in enumerate(clusters):
                lab[c] = n
            cset = set(centers)
            for i in range(nf):
                f.write("%5i %i %i %9.3f\n" % (i, lab[i], 1 if i in cset else 0, cost[i]))
    return D


def zoom_on_pairs(ctx, zoom_list, trunc, inter_low, inter_high):
    """zoom/dist_i_j.dat and zoom/2d_correlate_i_j.dat for every listed pair (conan.py:170-215, :1850-1859):
    the pair's distance trace with its hysteresis state, and its Pearson correlation with every other pair."""
    import torch
    os.makedirs("zoom", exist_ok=True)
    lay = ctx.layout
    pi, pj = lay.pair_labels()
    index = {(int(a), int(b)): n for n, (a, b) in enumerate(zip(pi, pj))}
    listed = np.zeros(ctx.P, bool)
    listed[ctx.seen] = True
    varying = ctx.varying()
    for pair in zoom_list:
        p = index.get(pair, index.get(pair[::-1]))
        if p is None or not listed[p]:
            print("Warning, pair (%i, %i) in zoom list never formed an interaction!" % pair)
            rvec = np.full(ctx.nf, trunc)
        else:
            col = ctx.d_stack[:, p].view(torch.int16).cpu().numpy().view(np.uint16)
            rvec = ctx.lut_host[col].astype(np.float64)
        inter = 0
        # the reference compares float16 distances with the cut-offs, which numpy rounds to float16 first
        lo16, hi16 = float(np.float16(inter_low)), float(np.float16(inter_high))
        with open("zoom/dist_%i_%i.dat" % pair, "w") as f:
            f.write("# t (ps) r (nm) interact?\n")
            for t, r in zip(ctx.time_vec, rvec):
                if r < lo16:
                    inter = 1
                elif r > hi16:
                    inter = 0
                f.write("%9d %12.6f %3i\n" % (t, r, inter))
        r_all = np.zeros(ctx.P)
        if len(set(rvec)) > 1:
            res = stage3.linregress_all(ctx.d_stack, ctx.P, ctx.nlevels, ctx.lut, rvec[None], ctx.shift)
            r_all = np.where(listed & varying, res["r"][0], 0.0)
        ctx.write_map("zoom/2d_correlate_%i_%i.dat" % pair, ctx.to_map(r_all, 0.0), fmt="%5i %5i %12.6f\n")


def run_all(opts, ret, s2, trunc, trunc_inter, nres, nterm, asymm_data):
    ctx = None

    def context():
        nonlocal ctx
        if ctx is None:
            ctx = _Ctx(s2)
        return ctx

    if opts.get('zoom'):
        print("Zooming in on interesting residue pairs...")
        zoom_on_pairs(context(), opts['zoom_list'], trunc, trunc_inter, opts.get('_trunc_inter_high', trunc_inter))
    if opts.get('pearson_time'):
        pearson_time(context())
        print("")
    if opts.get('pearson_obs_str'):
        pearson_obs(context(), opts['pearson_obs_str'], ignore_time=opts.get('ignore_obs_time', False))
    if opts.get('pca_make'):
        print("Starting principal component analysis...")
        pca(context(), opts['pca_make'])
    if opts.get('k_traj_clusters'):
        print("Starting contact map-based cluster analysis...")
        print("")
        cluster_trajectory(context(), opts['ks_traj'], trunc, trunc_inter, bool(asymm_data))
    if opts.get('pca_dir') and not opts.get('pca_make'):
        pca_read(context(), opts['pca_dir'])
    for k in ('blocking', 'k_res_clusters'):
        if opts.get(k):
            print("The %s analysis is outside the accelerated path and not available in this build." % k.upper())
