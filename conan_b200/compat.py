"""The reference's module-level helpers, for `import conan` users who call the pieces of the pipeline on their own
Python objects (HITS-MBM/conan README.md:53).

The command line and `read_process_frames` run on the GPU stages; the functions here take and return what the
reference's functions of the same name take and return -- sparse frame dictionaries {(i, j): r}, the
`aggregates` rows, float16 `vectors`, condensed distance vectors, open files -- so they work on host objects by
construction.  Signatures, file formats and error behaviour follow conan.py (line numbers in each docstring);
the bodies are array code, not the reference's loops.

Several of the reference's functions read script globals instead of taking arguments (`nres` in rmsd_frame,
`pairs_legend` / `asymm` in plot_pc and print_hierarchical_clusters, `gnus_path` / `domains` / `trunc_dr` in the
plotting calls).  They are looked up the same way here: in the `conan` module (or
`__main__` when conan.py is the script), where `main()` publishes the parsed options (conan.py:1748) and where an
importing program sets them by hand.
"""
import math
import os
import sys

import numpy as np


def _global(name, default=None):
    """A script global of the reference: conan.<name>, else __main__.<name>, else the default."""
    for mod in ("conan", "__main__", "conan_b200.frontend"):
        m = sys.modules.get(mod)
        if m is not None and hasattr(m, name):
            return getattr(m, name)
    if default is None:
        raise NameError("conan.%s is not set (the reference reads it as a module global)" % name)
    return default


def publish_globals(opts):
    """globals().update(opts) of conan.py:1748, for every module a caller may be looking at."""
    for mod in ("conan", "__main__", "conan_b200.frontend"):
        m = sys.modules.get(mod)
        if m is not None and (mod != "__main__" or os.path.basename(getattr(m, "__file__", "") or "") == "conan.py"):
            for k, v in opts.items():
                if not k.startswith("_"):
                    setattr(m, k, v)


# ------------------------------------------------------------------------------------------------
# dmf.xpm, one frame at a time (conan.py:486-567)
# ------------------------------------------------------------------------------------------------
def read_next_time(inputf):
    """Advance `inputf` to the next frame title and return its time in ps; -inf at the end of the file
    (conan.py:486-501: one line is consumed first, then lines are scanned for `/* title: "t=..." */`)."""
    if not inputf.readline():
        return float('-Inf')
    for line in iter(inputf.readline, ''):
        t = line.split()
        if len(t) > 1 and t[1] == 'title:':
            title = line[line.find('"') + 1:line.rfind('"')]
            return float(title[2:].split()[0])
    return float('-Inf')


def read_frame(inputf, nres, nlevels, nchar, dict_legend, nterm, trunc, asymm_data, dimer):
    """The matrix that follows the current position of `inputf` as {(i, j): r} holding only r < trunc
    (conan.py:503-567).  Symmetric runs keep the pairs i < j; asymmetric runs map the xpm indices through
    xdict / ydict; DIMER transposes the frame when the weighted contacts below the diagonal outweigh those
    above it (:553-565)."""
    line = inputf.readline()
    while line.split()[0] != "static":
        line = inputf.readline()
    for _ in range(nlevels + 1):
        line = inputf.readline()
    line = inputf.readline()
    while not line.split()[0].startswith('"'):
        line = inputf.readline()
    rows = []
    for i in range(nres):
        rows.append(line[1:1 + nres * nchar])
        line = inputf.readline()
    # rows are written top down from the last residue; column m of a row is residue m
    codes = np.array([[row[k:k + nchar] for k in range(0, nres * nchar, nchar)] for row in rows])
    uniq, inv = np.unique(codes, return_inverse=True)
    vals = np.array([dict_legend[c] for c in uniq], dtype=np.float64)[inv].reshape(nres, nres)[::-1]   # vals[n, m]
    frame = {}
    if not asymm_data:
        nn, mm = np.nonzero(np.tril(vals < trunc, -1))                 # the row of residue n holds the pairs (m, n), m < n
        for n, m in zip(nn, mm):
            frame[(int(m) + nterm, int(n) + nterm)] = float(vals[n, m])
    else:
        xdict, ydict, xres, yres, xterm, yterm = asymm_data
        for n in range(nres):
            if n + 1 not in ydict:
                continue
            for m in np.flatnonzero(vals[n] < trunc):
                if int(m) + 1 in xdict:
                    frame[(xdict[int(m) + 1], ydict[n + 1])] = float(vals[n, m])
    if dimer:
        above = sum((trunc - r) / trunc for (i, j), r in frame.items() if i < j)
        below = sum((trunc - r) / trunc for (i, j), r in frame.items() if i > j)
        if above < below:
            return {(j, i): r for (i, j), r in frame.items()}
    return frame


# ------------------------------------------------------------------------------------------------
# per-pair bookkeeping (conan.py:24-45, :599-629, :1414-1432)
# ------------------------------------------------------------------------------------------------
def update_aggregate(r, time, trunc, trunc_inter, trunc_inter_high, agg, inter_mode=0):
    """One frame of one pair: agg = [first, last, n_int, n_enc, status, sum r, sum r^2] -> the updated row
    (conan.py:599-611).  The GPU path keeps the same seven numbers per pair as exact integers
    (conan_pair_stats); this is the scalar form for callers that hold their own `aggregates` array."""
    from .frontend import update_interaction
    first, last, n_int, n_enc, status, sumr, sumr2 = tuple(agg)
    new_status = update_interaction(r, trunc_inter, trunc_inter_high, status, inter_mode)
    if new_status:
        if not status:
            n_enc += 1
            if first == -1:
                first = time
        n_int += 1
        last = time
    return np.array([first, last, n_int, n_enc, new_status, sumr + r, sumr2 + r * r])


def prepare_dimer_transpose(pairs_list, pairs_legend):
    """(npad, map): row of the transposed pair for every row of `vectors`, new rows appended for pairs whose
    transpose was never seen (conan.py:613-629)."""
    n_len = len(pairs_list)
    known = set(pairs_list)
    mapping, pad = [], []
    for i, pair in enumerate(pairs_list):
        t = pair[::-1]
        if t in known:
            mapping.append(pairs_legend[t])
        else:
            mapping.append(n_len + len(pad))
            pad.append(i)
    return len(pad), mapping + pad


def _ssd(dict_a, dict_b, trunc):
    keys = set(dict_a) | set(dict_b)
    if not keys:
        return 0.0
    a = np.fromiter((dict_a.get(k, trunc) for k in keys), dtype=np.float64, count=len(keys))
    b = np.fromiter((dict_b.get(k, trunc) for k in keys), dtype=np.float64, count=len(keys))
    return float(((a - b) ** 2).sum())


def rmsd_frame(dict_a, dict_b, trunc, xres, yres, asymm, dimer):
    """Distance between two frames: sqrt(2 sum (ra - rb)^2) / nres, or sqrt(sum / (xres yres)) for asymmetric
    runs, pairs missing from a frame standing at trunc; DIMER takes the smaller of b and b transposed
    (conan.py:1414-1432; `nres` is the module global the reference reads)."""
    s = _ssd(dict_a, dict_b, trunc)
    if dimer:
        s = min(s, _ssd(dict_a, {(j, i): r for (i, j), r in dict_b.items()}, trunc))
    if asymm:
        return math.sqrt(s / (xres * yres))
    return math.sqrt(2 * s) / _global("nres")


def plot_frames(dict_frame, first_frame, prev_frame, pairs_list, dr_mode, xterm, xres, yterm, yres, trunc, matrices,
                clean_matrices, n, time, rmsd_perframe_file, asymm, dimer):
    """matrices/NNNNN.dat, the differential maps asked for by dr_mode and one line of time_mdmat_rmsd.dat for
    frame n (conan.py:631-667).  The gnuplot calls are issued only when GNUS_PATH is set, as in the reference."""
    from .frontend import write_matrix
    gnus_path = _global("gnus_path", "")
    domains = _global("domains", 0)

    def gnuplot(cmd):
        if gnus_path:
            os.system(cmd)

    if matrices:
        with open("matrices/%05d.dat" % n, "w") as f:
            write_matrix(f, time, xterm, xres, yterm, yres, trunc, dict_frame, False, asymm)
        if clean_matrices:
            gnuplot("""gnuplot -e "domains=%d;maxz=%f;label_str='Residue index';inputfile='matrices/%05d.dat';outputfile='frames/%05d.png';"""
                    % (domains, trunc, n, n) + """title_str='t = %7.3f (ns)';cblabel_str = 'Distance (nm)" %s/script_single.gnu"""
                    % (time * 0.001, gnus_path))
            os.remove("matrices/%05d.dat" % n)
    if first_frame:
        diffs = {}
        for tag, other in (("init", first_frame), ("prev", prev_frame)):
            diffs[tag] = {p: dict_frame.get(p, trunc) - other.get(p, trunc) for p in set(dict_frame) | set(other)}
        r0 = rmsd_frame(dict_frame, first_frame, trunc, xres, yres, asymm, dimer)
        rp = rmsd_frame(dict_frame, prev_frame, trunc, xres, yres, asymm, dimer)
        rmsd_perframe_file.write("%9.4f %9.4f %9.4f\n" % (0.001 * time, rp, r0))
        if matrices:
            for tag, modes, label in (("init", (1, 3), "Change up to"), ("prev", (2, 3), "Change at")):
                if dr_mode not in modes:
                    continue
                name = "matrices/%05d_dr.%s.dat" % (n, tag)
                with open(name, "w") as f:
                    write_matrix(f, time, xterm, xres, yterm, yres, trunc, diffs[tag], True, asymm)
                if clean_matrices:
                    trunc_dr = _global("trunc_dr", trunc)
                    for dom in ((0, 1) if domains else (0,)):
                        gnuplot('gnuplot -e "domains=%d;maxz=%f;inputfile=\'%s\';outputfile=\'frames/%05d_dr.%s%s.png\';'
                                'title_time=\'%s t = %7.3f (ns)\'" %s/script_dr.gnu'
                                % (dom, trunc_dr, name, n, tag, ".domains" if dom else "", label, time, gnus_path))
                    os.remove(name)
    print("done with frame corresponding to ", time, "ps", end="\r")


# ------------------------------------------------------------------------------------------------
# inter-residue correlation (conan.py:669-695)
# ------------------------------------------------------------------------------------------------
def get_inter(residue_vectors, nterm, nres, pearson_2d_density_file):
    """Pearson correlation between the weighted-contact series of every two residues, written as an
    nres x nres table; returns {(i, j): 1 - r^2} for i != j (0-based).  A constant series correlates with
    nothing (r = 0, and 0 on its own diagonal entry)."""
    x = np.asarray(residue_vectors, dtype=np.float64)[:nres]
    varying = x.max(axis=1) != x.min(axis=1)
    with np.errstate(invalid="ignore", divide="ignore"):
        c = np.corrcoef(x) if nres > 1 else np.ones((1, 1))
    c = np.where(varying[:, None] & varying[None, :], np.atleast_2d(c), 0.0)
    np.fill_diagonal(c, np.where(varying, 1.0, 0.0))
    for i in range(nres):
        pearson_2d_density_file.write("".join("%5i %5i %15.6f\n" % (nterm + i, nterm + j, c[i, j]) for j in range(nres)))
        pearson_2d_density_file.write("\n")
    return {(i, j): 1 - c[i, j] ** 2 for i in range(nres) for j in range(nres) if i != j}


# ------------------------------------------------------------------------------------------------
# principal components (conan.py:413-452, :1073-1095)
# ------------------------------------------------------------------------------------------------
def plot_pc(xterm, xres, yterm, yres, ii, vec, pairs, pairs_dict, gnus_path, domains):
    """pca/pc.<ii+1>.dat: the component's coefficient for every residue pair, 0 where the pair has no row
    (conan.py:413-428; `pairs_legend` and `asymm` are the module globals the reference reads)."""
    legend = _global("pairs_legend", pairs_dict)
    asymm = _global("asymm", False)
    grid = np.zeros((xres, yres))
    for pair in pairs:
        v = vec[legend[pair]]
        i, j = pair
        if xterm <= i < xterm + xres and yterm <= j < yterm + yres:
            grid[i - xterm, j - yterm] = v
        if not asymm and xterm <= j < xterm + xres and yterm <= i < yterm + yres:
            grid[j - xterm, i - yterm] = v
    with open("pca/pc.%i.dat" % (ii + 1), "w+") as f:
        for i in range(xres):
            f.write("".join("%5i %5i %12.6f\n" % (xterm + i, yterm + j, grid[i, j]) for j in range(yres)))
            f.write("\n")
    if gnus_path:
        maxz = max(-min(vec), max(vec))
        os.system("""gnuplot -e "domains=%i;i=%i;maxz=%f;input_file='pca/pc.%i.dat'" %s/script_pca.gnu """
                  % (domains, ii + 1, maxz, ii + 1, gnus_path))


def project_pc(v, time_vec, vectors, pairs_list, change_sign=True):
    """pca/pca_projections.dat: every frame projected on the columns of v; with change_sign a component whose
    projection falls with time is flipped.  Returns the (possibly flipped) components (conan.py:430-452)."""
    from scipy.stats import linregress
    x = np.asarray(vectors, dtype=np.float64)              # [pairs, frames]
    ordered = np.array(v, dtype=np.float64)
    proj = x.T @ ordered                                    # [frames, components]
    print("Projecting principal components...")
    if change_sign:
        for k in range(ordered.shape[1]):
            if linregress(proj[:, k], time_vec)[2] < 0:
                ordered[:, k] *= -1
                proj[:, k] *= -1
    with open("pca/pca_projections.dat", "w+") as f:
        for t, row in zip(time_vec, proj):
            f.write("%12i" % t + (len(row) * " %12.6f") % tuple(row) + "\n")
    return ordered


def read_pca(pca_dir, xterm, xres, yterm, yres, time_vec, asymm, vectors_mean0, pairs_list):
    """Read pc.1.dat, pc.2.dat, ... of an earlier run and project this run's mean-free vectors on them
    (conan.py:1073-1095)."""
    from .frontend import read_matrix
    comps = []
    while os.path.isfile("%s/pc.%i.dat" % (pca_dir, len(comps) + 1)):
        with open("%s/pc.%i.dat" % (pca_dir, len(comps) + 1)) as f:
            pc = read_matrix(f, xterm, xres, yterm, yres, asymm, 1.0)
        vec = np.array([pc[pair] for pair in pairs_list])
        print("Norm of PC # %i = %12.6f" % (len(comps) + 1, math.sqrt(float((vec ** 2).sum()))))
        comps.append(vec)
    print("Reading in PC's.")
    print("I found %i principal components..." % len(comps))
    project_pc(np.array(comps).T, time_vec, vectors_mean0, pairs_list, change_sign=False)


# ------------------------------------------------------------------------------------------------
# hierarchical clustering (conan.py:1172-1185, :1290-1412)
# ------------------------------------------------------------------------------------------------
def _square(metric_cond, n):
    from scipy.spatial.distance import squareform
    return squareform(np.asarray(metric_cond, dtype=np.float64), checks=False) if n > 1 else np.zeros((1, 1))


def get_best_center(cluster, metric_cond, n):
    """The member of `cluster` with the smallest summed distance to the other members; metric_cond is the
    condensed distance vector over n elements (conan.py:1172-1185; the first such member on ties)."""
    idx = np.asarray(cluster, dtype=int)
    if len(idx) == 0:
        return 0
    sub = _square(metric_cond, n)[np.ix_(idx, idx)]
    return int(idx[int(np.argmin(sub.sum(axis=1)))])


def print_hierarchical_clusters(dir, vectors, trunc, trunc_inter, nframes, clusters, centers, time_vec, cost_list, xterm,
                                xres, yterm, yres, asymm, type, matrices):
    """assignment.dat and, for trajectory clusters, clusterN/average_map.dat and lifetime.dat
    (conan.py:1290-1345; `pairs_legend` is the module global the reference reads)."""
    centers_set = set(centers)
    if type == 'trj':
        legend = _global("pairs_legend")
        x = np.asarray(vectors)
        for c, members in enumerate(clusters):
            cdir = "%s/cluster%i" % (dir, c)
            os.makedirs(cdir, exist_ok=True)
            if matrices:
                for k, j in enumerate(members):
                    os.system("ln -s ../../../frames/%05i.png %s/%05i.png" % (j, cdir, k))
                    if j in centers_set:
                        os.system("ln -s ../../../frames/%05i.png %s/center.png" % (j, cdir))
            sub = x[:, np.asarray(members, dtype=int)] if len(members) else x[:, :0]
            mean_rows = sub.mean(axis=1) if sub.shape[1] else np.full(x.shape[0], np.nan)   # float16 in, float16 out (:1319)
            life_rows = 100.0 * (sub < trunc_inter).sum(axis=1) / sub.shape[1] if sub.shape[1] else np.zeros(x.shape[0])
            with open(cdir + "/average_map.dat", "w") as fa, open(cdir + "/lifetime.dat", "w") as fl:
                for i in range(xterm, xterm + xres):
                    for j in range(yterm, yterm + yres):
                        if i == j and not asymm:
                            d, life = 0, 100.0
                        elif (i, j) in legend:
                            row = legend[(i, j)]
                            d, life = mean_rows[row], life_rows[row]
                        else:
                            d, life = trunc, 0.0
                        fa.write("%4d %4d %10.6f\n" % (i, j, d))
                        fl.write("%4d %4d %10.6f\n" % (i, j, life))
                    fa.write("\n")
                    fl.write("\n")
    else:
        raise NotImplementedError("residue clusters (K_RES_CLUSTERS) are outside the trajectory contact path")
    order = sorted((j, c) for c, members in enumerate(clusters) for j in members)
    with open("%s/assignment.dat" % dir, "w") as f:
        f.write("# element_ind cluster_ind center? dist\n")
        for ind, el in order:
            f.write("%5i %i %i %9.3f\n" % (ind, el, 1 if ind in centers_set else 0, cost_list[ind]))


def call_hierarchical(dir, vectors, trunc, trunc_inter, condensed_dist, n_elements, ksi, title, time_vec, type, xterm,
                      xres, yterm, yres, asymm, matrices):
    """Ward linkage of a condensed distance vector, linkage.txt, and for every requested number of clusters
    the assignment with medoids and distances to them (conan.py:1364-1412).  No dendrogram is drawn
    (matplotlib is not a dependency); ksi == [0] asks for the numbers on standard input as the reference does
    after showing its plot."""
    from scipy.cluster.hierarchy import fcluster, linkage
    from .frontend import parse_numbers
    cond = np.asarray(condensed_dist, dtype=np.float64)
    Z = linkage(cond, 'ward')
    with open("%s/linkage.txt" % dir, "w+") as f:
        for row in Z:
            f.write("%6i %6i %12.6f %6i\n" % tuple(row))
    ks = ksi
    if ksi == [0]:
        ks = parse_numbers(input("Number of clusters? (multiple values are allowed, or a single interval as i-j).\n"))
    sq = _square(cond, n_elements)
    for k in ks:
        os.makedirs("%s/%i" % (dir, k), exist_ok=True)
        assign = fcluster(Z, k, criterion='maxclust')
        clusters = [list(np.flatnonzero(assign == c + 1)) for c in range(k)]
        clusters.sort(key=len, reverse=True)
        centers = [get_best_center(c, cond, n_elements) for c in clusters]
        cost = np.zeros(n_elements)
        for members, ctr in zip(clusters, centers):
            cost[members] = sq[members, ctr]
        order = np.argsort(centers, kind="stable")
        clusters = [clusters[o] for o in order]
        centers = [centers[o] for o in order]
        print_hierarchical_clusters("%s/%i/" % (dir, k), vectors, trunc, trunc_inter, n_elements, clusters, centers, time_vec,
                                    list(cost), xterm, xres, yterm, yres, asymm, type, matrices)
        gnus_path = _global("gnus_path", "")
        if gnus_path:
            os.system("""gnuplot -e "inputdir='%s/%i'" %s/cluster_assignment.gnu""" % (dir, k, gnus_path))
