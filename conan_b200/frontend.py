"""The reference's surface for the trajectory contact path, on top of the CUDA stages.

`./conan.py input.inp` reads the same keyword file as the reference (conan.py:1496-1713), runs
stage 1 in-process on the GPU where the reference shells out to `gmx mdmat` (conan.py:1437-1453),
and `read_process_frames` returns the same tuple (conan.py:1021/1023) and writes the same files,
with the per-frame work done by conan_pair_stats instead of Python loops.

Differences a user can observe (documented in DESIGN.md): pairs_list is ordered by residue pair
instead of by the hash order of first appearance (every consumer goes through pairs_legend);
dmf.xpm is still written after stage 1 unless WRITE_XPM no is given or the stack is very large.
"""
import os
import shutil
import sys

import numpy as np

from . import levels as lv
from . import _lib, tables, topology, xpm
from ._lib import ConanError

# ------------------------------------------------------------------------------------------------
# keyword file
# ------------------------------------------------------------------------------------------------
DEFAULT_OPTS = {
    'trunc': None, 'mean': True, 'run_mdmat': True, 'nterm': 1, 'matrices': True, 'dr_mode': 0,
    'trunc_lifetime': 0.5, 'asymm': False, 'clean_matrices': False, 'dimer': False, 'make_movie': True,
    'trunc_dr': None, 'trunc_inter': None, 'trunc_inter_high': None, 'domains': 0, 'begin': None, 'dt': None,
    'end': None, 'patch_time': False, 'economy': False, 'pearson_inter': False, 'reread': False, 'indexf': None,
    'zoom': False, 'nlevels': None,
}

_YES = lambda v: v == 'yes'
_NOT_NO = lambda v: v != 'no'
# keyword -> (option key, converter); the irregular ones are handled in read_options
_SIMPLE = {
    'TRAJ': ('traj', str), 'COORD': ('coord', str), 'NLEVEL': ('nlevels', int), 'PATCH_TIME': ('patch_time', _YES),
    'MEAN': ('mean', _YES), 'INDEX': ('indexf', str), 'BEGIN': ('begin', int), 'END': ('end', int), 'DT': ('dt', int),
    'TRUNC_INTER': ('trunc_inter', float), 'TRUNC_INTER_HIGH': ('trunc_inter_high', float),
    'TRUNC_LIFETIME': ('trunc_lifetime', float), 'TRUNC_DR': ('trunc_dr', float), 'GNUS_PATH': ('gnus_path', str),
    'RUN_MDMAT': ('run_mdmat', _YES), 'MATRICES': ('matrices', _YES), 'CLEAN_MATRICES': ('clean_matrices', _YES),
    'PEARSON_TIME': ('pearson_time', _YES), 'PEARSON_INTER': ('pearson_inter', _YES), 'PCA_READ': ('pca_dir', str),
    'PEARSON_OBS': ('pearson_obs_str', str), 'REREAD': ('reread', str), 'MAKE_MOVIE': ('make_movie', _NOT_NO),
    'ECONOMY': ('economy', _NOT_NO), 'DIMER': ('dimer', _YES), 'WRITE_XPM': ('write_xpm', _NOT_NO),
    'PBC': ('pbc', _NOT_NO), 'GROUP': ('group', int),
}


def parse_numbers(vals):
    """'1 3-5 8' -> [1, 3, 4, 5, 8] (conan.py:1347-1362)."""
    out = set()
    toks = vals.replace("-", " - ").split()
    last, i = 0, 0
    while i < len(toks):
        if toks[i] == '-':
            if i + 1 < len(toks):
                out |= set(range(last + 1, int(toks[i + 1]) + 1))
                i += 1
        else:
            last = int(toks[i])
            out.add(last)
        i += 1
    return list(out)


def read_zoom_list(zoomfile):
    pairs = []
    for line in zoomfile:
        f = line.split()
        if f and not line.startswith("#"):
            pairs.append(tuple(sorted((int(f[0]), int(f[1])))))
    return pairs


def read_stages(stagef):
    stages = []
    for line in stagef:
        f = line.split()
        if not f or line.startswith("#"):
            continue
        name = line[line.find('"') + 1:line.rfind('"')] if len(f) > 2 else "Stage %d" % (len(stages) + 1)
        stages.append((int(f[0]), int(f[1]), name))
    return stages


def read_options(finput, opts):
    """Keyword file -> opts dict, keys and messages as the reference (conan.py:1496-1713)."""
    opts['trunc_given'] = False
    for line in finput:
        f = line.split()
        if line.strip().startswith('#') or len(f) < 2:
            continue
        key, val = f[0].upper(), f[1]
        if key in _SIMPLE:
            name, conv = _SIMPLE[key]
            opts[name] = conv(val)
        elif key == 'TRUNC':
            opts['trunc'] = float(val)
            opts['trunc_given'] = True
        elif key == 'DR_MODE':
            mode = {'init': 1, 'prev': 2, 'both': 3}.get(val)
            if mode:
                opts['dr_mode'] = mode
        elif key == 'NTERM':
            opts['nterm'] = opts['xterm'] = opts['yterm'] = int(val)
        elif key in ('START_X', 'START_Y'):
            ax = key[-1].lower()
            opts['start_' + ax] = int(val)
            opts.setdefault(ax + 'term', int(val))
            opts['asymm'] = 1
        elif key in ('NRES_X', 'NRES_Y'):
            opts[key[-1].lower() + 'res'] = int(val)
            opts['asymm'] = True
        elif key in ('NTERM_X', 'NTERM_Y'):
            opts[key[-1].lower() + 'term'] = int(val)
            opts['asymm'] = True
        elif key in ('LIST_X', 'LIST_Y'):
            opts['asymm'] = 1
            opts['list_' + key[-1].lower()] = str(val)
        elif key == 'SHADOW_TOL':
            opts['shadow'] = True
            opts['shadow_tol'] = float(val)
        elif key == 'PCA_MAKE':
            if not val.isdigit() or int(val) < 1:
                print("The correct call of PCA is PCA_MAKE <# PCs>, where <# PCs> is an integer defining the number of PCs to output.")
            else:
                opts['pca_make'] = int(val)
        elif key == 'ZOOM_LIST':
            opts['zoom'] = True
            with open(str(val)) as zf:
                opts['zoom_list'] = read_zoom_list(zf)
        elif key == 'BLOCKING':
            if val != 'no':
                opts['blocking'] = True
        elif key in ('COORD_PDB', 'COORD_PDB_X', 'COORD_PDB_Y'):
            opts['coordpdb'] = True
            opts[{'COORD_PDB': 'pdbf', 'COORD_PDB_X': 'pdbx', 'COORD_PDB_Y': 'pdby'}[key]] = str(val)
        elif key == 'IGNORE_OBS_TIME':
            if val != 'no':
                opts['ignore_obs_time'] = True
        elif key in ('K_RES_CLUSTERS', 'K_TRAJ_CLUSTERS'):
            ks = parse_numbers(" ".join(f[1:]))
            what = 'residues' if key == 'K_RES_CLUSTERS' else 'frames'
            if ks != [0]:
                print("Will create the following number of clusters of %s:" % what, " ".join("%i" % k for k in ks))
            else:
                print("Will choose the number of clusters of %s interactively." % what)
            if key == 'K_RES_CLUSTERS':
                opts['k_res_clusters'] = opts['ks_res'] = ks
            else:
                opts['k_traj_clusters'] = opts['ks_traj'] = ks
        elif key in ('DOMAINS', 'DOMAINS_X', 'DOMAINS_Y'):
            opts['domains'] = True
            opts[{'DOMAINS': 'domainf', 'DOMAINS_X': 'domainf_xn', 'DOMAINS_Y': 'domainf_yn'}[key]] = str(val)
        elif key == 'STAGES':
            opts['stages'] = True
            opts['stagef'] = str(val)
        else:
            print("Cannot understand the keyword %s given in the input file." % key)
    if opts['asymm']:
        if 'shadow' in opts:
            print("Removing shadows is impossible in an asymmetric run.")
            opts.pop('shadow')
        if 'k_res_clusters' in opts:
            print("Clustering residues is impossible in an asymmetric run.")
            opts.pop('k_res_clusters')
            opts.pop('ks_res')
        for ax in 'xy':
            if 'list_' + ax in opts:
                reslist = np.atleast_1d(np.loadtxt(opts['list_' + ax], dtype=int))
                opts[ax + 'term'] = opts.get(ax + 'term', reslist[0])
                opts[ax + 'dict'] = {int(b): opts[ax + 'term'] + a - 1 for a, b in enumerate(reslist)}
                opts[ax + 'res'] = len(reslist)
                if 'start_' + ax in opts:
                    print("You cannot give a range AND a list for the residue list. We will only use the list.")
                    opts.pop('start_' + ax)
            if 'start_' + ax in opts:
                opts[ax + 'dict'] = {opts['start_' + ax] + i: opts[ax + 'term'] + i for i in range(opts[ax + 'res'])}
    if 'pca_make' in opts and 'pca_dir' in opts:
        print("PCA can either be performed or PC's can be read in, but not both. We will attempt reading them in.")
        opts.pop('pca_make')
    if 'gnus_path' not in opts:
        print(" .gnu paths have not been given. NO PLOTS will be produced.")
        opts['gnus_path'] = None
    if not opts['matrices'] and opts['make_movie']:
        print("Can't make a movie without matrices! We will turn off movie making.")
        opts['make_movie'] = False
    if opts['dimer']:
        if not opts['asymm'] or not (opts.get('xterm', 0) == opts.get('yterm', 1) and opts.get('xres', 0) == opts.get('yres', 1)):
            print("'Dimer mode' requires 1. asymmetric mode and 2. the indices of the X- and Y- residues to be identical. We will turn off the option now.")
            opts['dimer'] = False
    if opts['economy']:
        for key, msg in (('zoom', "Zooming onto residues"), ('pca_make', "PCA"), ('k_traj_clusters', "Trajectory clustering"),
                         ('pearson_time', "Pearson correlation with time"),
                         ('pearson_obs_str', "Pearson correlation with an observable")):
            if opts.get(key):
                print("%s is not available in economy mode." % msg)
                opts.pop(key)
    return opts


adjust_truncs = lv.adjust_truncs


def weight_contacts(r, trunc):
    return (trunc - r) / trunc


def update_interaction(r, trunc_inter, trunc_inter_high, status, inter_mode):
    """The hysteresis rule of conan.py:24-45, kept for `import conan` users; the GPU path applies it in level space."""
    if trunc_inter == trunc_inter_high:
        return int(r < trunc_inter)
    if r < trunc_inter:
        return 1
    if r > trunc_inter_high:
        return 0
    if inter_mode == 1:
        return (r - trunc_inter) / (trunc_inter_high - trunc_inter)
    return status


def cond_index(i, j, n):
    return int(n * (n - 1) / 2 - (n - i) * (n - i - 1) / 2 + j - i - 1)


def which(program):
    found = shutil.which(program)
    return found if found else ''


def handle_backups():
    """Move output directories of an earlier run to backup.N (conan.py:1097-1115)."""
    names = ['frames', 'input', 'matrices', 'movies', 'aggregate', 'cluster_res', 'cluster_trj', 'pca', 'zoom', 'blocking']
    present = [d for d in names if os.path.isdir(d)]
    if not present:
        return 0
    n = 1
    while os.path.isdir("backup.%i" % n):
        n += 1
    target = "backup.%i" % n
    print("Moving directories containing old calculation to: %s ..." % target)
    os.mkdir(target)
    for d in present:
        shutil.move(d, os.path.join(target, d))
    return 1


# ------------------------------------------------------------------------------------------------
# stage 1
# ------------------------------------------------------------------------------------------------
class Stage1Result:
    """What `gmx mdmat` would have left in dmf.xpm, kept on the GPU: full nres x nres geometry."""

    def __init__(self, levels, geom, times, trunc, nlevels, n_boundary, shard=None):
        self.levels, self.geom, self.times = levels, geom, np.asarray(times, dtype=np.float64)
        self.trunc, self.nlevels, self.n_boundary = trunc, nlevels, n_boundary
        # (first frame, end frame) of this rank's rows of `levels` within `times` when the run is spread over
        # several GPUs (torch.distributed); None = all frames are here
        self.shard = shard

    def square_host(self):
        """[nf, nres, nres] uint16 on the host (for dmf.xpm / matrices at small sizes)."""
        import torch
        flat = self.levels.view(torch.int16).cpu().numpy().view(np.uint16)[:, :self.geom.P]
        return xpm.square_from_condensed(flat, self.geom.nres)


def distributed():
    """(rank, world) when the program was started by torch.distributed.run on several GPUs (one process per
    GPU, frames sharded over them), else (0, 1).  The process group is created on first use: NCCL, or gloo when
    CONAN_DIST_BACKEND=gloo (several ranks on one GPU, for tests)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return 0, 1
    import torch
    import torch.distributed as dist
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = os.environ.get("CONAN_DIST_BACKEND", "nccl")
        local = int(os.environ.get("LOCAL_RANK", "0")) % max(1, torch.cuda.device_count())
        torch.cuda.set_device(local)
        if backend == "nccl":
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return dist.get_rank(), dist.get_world_size()


_last_stage1 = None   # handed from mdmat_run to read_process_frames within one process
XPM_WRITE_LIMIT = 2.0e8   # level entries above which dmf.xpm is not written unless WRITE_XPM yes
SPARSE_STACK_BELOW = 0.04  # stage 2 takes its sparse variant when fewer than this share of the entries survived stage 1's pruning


def select_frames(times, begin, end, dt):
    """The -b / -e / -dt frame filter of gmx (first frame defines the dt phase)."""
    times = np.asarray(times, dtype=np.float64)
    keep = np.ones(len(times), dtype=bool)
    if begin:
        keep &= times >= begin
    if end:
        keep &= times <= end
    if dt:
        t0 = times[0] if len(times) else 0.0
        keep &= np.abs(np.mod(times - t0 + 0.5 * 1e-3, dt)) < 1e-3 * 2
    return np.flatnonzero(keep)


def choose_group(groups, stdin=None, preset=None):
    """gmx asks for the index group on stdin (the reference's tests send "2": Protein-H)."""
    if preset is not None:
        return preset
    for n, (name, idx) in enumerate(groups):
        print("Group %5d (%15s) has %5d elements" % (n, name, len(idx)))
    print("Select a group: ", end="", flush=True)
    line = (stdin or sys.stdin).readline()
    if not line.strip():
        print("")
        raise SystemExit("No index group given on standard input.")
    tok = line.split()[0]
    if tok.isdigit():
        return int(tok)
    for n, (name, _) in enumerate(groups):
        if name == tok:
            return n
    raise SystemExit("Unknown index group %s" % tok)


def mdmat_run(program, traj, coord, nlevels, trunc, mean, begin, end, dt, indexf, pbc=True, group=None,
              write_xpm=None, device="cuda", block=None):
    """Stage 1 in-process: the work of `gmx mdmat -frames` (conan.py:1437-1453).  `program` is ignored.

    block = (x0, xres, y0, yres), 0-based residues of the index group: compute only that X x Y block
    (an asymmetric run does not need the other pairs; no dmf.xpm is written then)."""
    global _last_stage1
    import torch
    from . import stage1
    from .xtc import XtcFile
    is_xtc = traj.lower().endswith(".xtc")
    nlevels = nlevels if nlevels else 40          # gmx mdmat defaults: -nlevels 40, -t 1.5
    trunc = trunc if trunc else 1.5
    top = topology.read_topology(coord)
    if is_xtc:
        x = XtcFile(traj)
    else:       # .trr / multi-frame .gro / multi-model .pdb: plain numbers, read on the host (trajio.py)
        from . import trajio
        try:
            x = trajio.open_trajectory(traj)
        except ValueError as err:
            print(err)
            sys.exit()
    if x.natoms < top.natoms:
        top = top.truncated(x.natoms)
    elif x.natoms > top.natoms:
        print("The trajectory has %d atoms but %s describes only %d." % (x.natoms, coord, top.natoms))
        sys.exit()
    groups = topology.read_ndx(indexf) if indexf else top.default_groups()
    gsel = choose_group(groups, preset=group)
    name, sel = groups[gsel]
    print("Selected %d: '%s' (%d atoms)" % (gsel, name, len(sel)))
    res_start = top.residue_runs(sel)
    frames = select_frames(x.times, begin, end, dt)
    if len(frames) == 0:
        print("No frames selected from %s." % traj)
        sys.exit()
    boxes = x.boxes[frames]
    triclinic = np.abs(boxes - boxes * np.eye(3)[None]).max() > 1e-6
    use_pbc = bool(pbc) and boxes[:, [0, 1, 2], [0, 1, 2]].min() > 0
    if use_pbc and triclinic:
        # gmx mdmat would search the triclinic shift vectors; carrying on without images would give
        # different distances for every pair split across a box face, so stop instead.
        print("The box of %s is triclinic; only rectangular boxes have a minimum-image path here." % traj)
        print("Make the molecules whole and rerun with 'PBC no', or convert the trajectory to a rectangular box.")
        sys.exit()
    nres_all = len(res_start) - 1
    if block is not None and (block[0] < 0 or block[2] < 0 or block[0] + block[1] > nres_all or block[2] + block[3] > nres_all):
        print("The X / Y residue ranges do not fit the %d residues of the index group; computing all pairs." % nres_all)
        block = None
    geom = stage1.Geometry(res_start, *block) if block is not None else stage1.Geometry(res_start)
    rank, world = distributed()
    dev = torch.device(device) if world == 1 else torch.device("cuda", torch.cuda.current_device())
    shard = None
    all_frames = frames
    if world > 1:      # contiguous frame ranges, one per GPU; nothing is exchanged in stage 1
        bounds = np.linspace(0, len(frames), world + 1).astype(int)
        shard = (int(bounds[rank]), int(bounds[rank + 1]))
        frames = frames[shard[0]:shard[1]]
        boxes = boxes[shard[0]:shard[1]]
        print("Rank %d of %d: frames %d .. %d of %d" % (rank, world, shard[0], shard[1] - 1, len(all_frames)))
    levels = torch.zeros((max(len(frames), 1), geom.stride), dtype=torch.uint16, device=dev)[:len(frames)]
    counters = torch.zeros(3, dtype=torch.int64, device=dev)
    batch = max(1, min(len(frames), 2048))
    work = torch.empty(stage1.mdmat_workspace_bytes(batch, geom), dtype=torch.uint8, device=dev)
    d_coords = torch.empty((batch, geom.natoms, 3), dtype=torch.float32, device=dev)
    d_status = torch.zeros((max(len(frames), 1),), dtype=torch.int32, device=dev)
    d_boxes = torch.from_numpy(np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 9)).to(dev) if use_pbc else None
    copy_stream = torch.cuda.Stream(dev)
    main_stream = torch.cuda.current_stream(dev)
    host_free = [None, None]      # event after the H2D copy that read each pinned buffer
    dev_free = [None, None]       # event after the kernel that read each device staging buffer
    if is_xtc:
        # The compressed frames go to the GPU as they are on disk and are decoded there (conan_xtc_decode_device_ws).
        # Two pinned staging buffers and a copy stream: the bytes of batch b+1 travel while batch b is decoded and
        # its distances are computed; nothing waits on the host between batches.
        lib = _lib.load()
        spans = []
        for b0 in range(0, len(frames), batch):
            ids = frames[b0:b0 + batch]
            lo, last = int(ids.min()), int(ids.max())
            spans.append((b0, ids, int(x.offsets[lo]), int(x.offsets[last + 1]) if last + 1 < x.nframes else int(x._buf.size)))
        max_bytes = max((sp[3] - sp[2] for sp in spans), default=0)
        h_stage = [torch.empty(max_bytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
        h_off = [torch.empty(batch, dtype=torch.int64).pin_memory() for _ in range(2)]
        d_stage = [torch.empty(max_bytes, dtype=torch.uint8, device=dev) for _ in range(2)]
        d_off = [torch.empty(batch, dtype=torch.int64, device=dev) for _ in range(2)]
        d_status.fill_(-1)
        slot = np.full(x.natoms, -1, np.int32)
        slot[np.asarray(sel, dtype=np.int64)] = np.arange(len(sel), dtype=np.int32)
        d_slot = torch.from_numpy(slot).to(dev)
        # scratch of the two-pass decoder, at most 1 GiB (with less than a batch's worth the frames go in rounds)
        d_dec_work = torch.empty(max(96 + 8 * x.natoms, min(int(lib.conan_xtc_decode_workspace_bytes(batch, x.natoms)), 1 << 30)),
                                 dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            for n, (b0, ids, lo_b, hi_b) in enumerate(spans):
                k = n % 2
                if host_free[k] is not None:
                    host_free[k].synchronize()            # the pinned buffer is rewritten by the host below
                nbytes = hi_b - lo_b
                h_stage[k][:nbytes].numpy()[:] = x._buf[lo_b:hi_b]
                h_off[k][:len(ids)].numpy()[:] = x.offsets[ids] - lo_b
                with torch.cuda.stream(copy_stream):
                    if dev_free[k] is not None:
                        copy_stream.wait_event(dev_free[k])
                    d_stage[k][:nbytes].copy_(h_stage[k][:nbytes], non_blocking=True)
                    d_off[k][:len(ids)].copy_(h_off[k][:len(ids)], non_blocking=True)
                    copied = torch.cuda.Event()
                    copied.record(copy_stream)
                host_free[k] = copied
                main_stream.wait_event(copied)
                _lib.check(lib.conan_xtc_decode_device_ws(d_stage[k].data_ptr(), nbytes, d_off[k].data_ptr(), len(ids), x.natoms,
                                                          d_slot.data_ptr(), len(sel), d_coords.data_ptr(),
                                                          d_status[b0:b0 + len(ids)].data_ptr(), d_dec_work.data_ptr(),
                                                          d_dec_work.numel(), main_stream.cuda_stream))
                decoded = torch.cuda.Event()
                decoded.record(main_stream)
                dev_free[k] = decoded
                stage1.mdmat_frames(d_coords[:len(ids)], None if d_boxes is None else d_boxes[b0:b0 + len(ids)], geom, trunc,
                                    nlevels, use_pbc, levels[b0:b0 + len(ids)], work, counters)
    else:
        # plain coordinates: the index group's atoms of a batch are gathered into a pinned buffer and copied on
        # the copy stream while the batch before is computed
        h_xyz = [torch.empty((batch, geom.natoms, 3), dtype=torch.float32).pin_memory() for _ in range(2)]
        d_xyz = [d_coords, torch.empty_like(d_coords)]
        with torch.cuda.device(dev):
            for n, b0 in enumerate(range(0, len(frames), batch)):
                ids = frames[b0:b0 + batch]
                k = n % 2
                if host_free[k] is not None:
                    host_free[k].synchronize()
                h_xyz[k][:len(ids)].numpy()[:] = x.coords(ids, sel)
                with torch.cuda.stream(copy_stream):
                    if dev_free[k] is not None:
                        copy_stream.wait_event(dev_free[k])
                    d_xyz[k][:len(ids)].copy_(h_xyz[k][:len(ids)], non_blocking=True)
                    copied = torch.cuda.Event()
                    copied.record(copy_stream)
                host_free[k] = copied
                main_stream.wait_event(copied)
                stage1.mdmat_frames(d_xyz[k][:len(ids)], None if d_boxes is None else d_boxes[b0:b0 + len(ids)], geom, trunc,
                                    nlevels, use_pbc, levels[b0:b0 + len(ids)], work, counters)
                done = torch.cuda.Event()
                done.record(main_stream)
                dev_free[k] = done
    torch.cuda.synchronize(dev)
    bad = torch.nonzero(d_status[:len(frames)]).flatten()
    if bad.numel():
        k = int(bad[0].item())
        raise ConanError(-3, "frame %d of the trajectory is malformed (decoder status %d)" % (int(frames[k]), int(d_status[k].item())))
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(counters)
    nb = int(counters[0].item())
    print("Computed %d distance matrices of %d residues (%d results within 2 ulp of a level boundary)."
          % (len(all_frames), geom.nres, nb))
    _last_stage1 = Stage1Result(levels, geom, x.times[all_frames], trunc, nlevels, nb, shard)
    # the share of (residue pair, frame) entries that passed the bounding-sphere pruning: tells stage 2 whether the
    # stack is sparse enough for its variant that skips stretches of nothing but last levels
    _last_stage1.survivor_frac = float(counters[1].item()) / max(1.0, float(geom.P) * len(all_frames))
    entries = float(len(all_frames)) * geom.nres * geom.nres
    if world > 1:
        write_xpm = bool(write_xpm) and False      # the stack is spread over the GPUs: no dmf.xpm (WRITE_XPM is ignored)
    if write_xpm is None:
        write_xpm = entries <= XPM_WRITE_LIMIT
    if block is not None:
        print("Asymmetric run: only the %d x %d block of residue pairs was computed; dmf.xpm is not written." % (geom.xres, geom.yres))
    elif write_xpm:
        import torch
        flat = levels.view(torch.int16).cpu().numpy().view(np.uint16)
        xpm.write_xpm_stack("dmf.xpm", flat, True, geom.nres, geom.nres, _last_stage1.times, trunc, nlevels)
    else:
        print("dmf.xpm not written (%.3g entries); the level stack stays on the GPU." % entries)
    return _last_stage1


# ------------------------------------------------------------------------------------------------
# xpm reading (file contract; used when RUN_MDMAT no)
# ------------------------------------------------------------------------------------------------
def read_index(inputf):
    """Header of dmf.xpm -> nres, nlevels, nchar, dict_legend, trunc (conan.py:455-482)."""
    line = inputf.readline()
    while line.split()[0] != "static":
        line = inputf.readline()
    dims = inputf.readline()[1:-3].split()
    nres, nlevels, nchar = int(dims[0]), int(dims[2]), int(dims[3])
    codes = []
    for _ in range(nlevels):
        line = inputf.readline()
        code = line[1:].split()[0]
        codes.append(code)
        if len(code) != nchar:
            print('warning! this character is of the wrong length:', code)
            print('carrying on but it is a good idea to check.')
    trunc = float(line.split()[-2][1:-1])
    dict_legend = dict(zip(codes, lv.legend_values(trunc, nlevels)))
    return nres, nlevels, nchar, dict_legend, trunc


def _levels_from_xpm(path, asymm_data):
    """Parse dmf.xpm into (times, level stack [nf, P] uint16, layout pieces)."""
    times, frames, trunc, nlevels = xpm.read_xpm(path)
    nres = frames.shape[1]
    if asymm_data:
        xdict, ydict, xres, yres, xterm, yterm = asymm_data
        xs = np.array([m for m in range(1, nres + 1) if m in xdict], dtype=int)
        ys = np.array([m for m in range(1, nres + 1) if m in ydict], dtype=int)
        # frames[f, y-1, x-1]; output block order is x-major over the residue labels
        xo = np.array([xdict[m] - xterm for m in xs])
        yo = np.array([ydict[m] - yterm for m in ys])
        block = np.full((len(times), xres, yres), nlevels - 1, dtype=np.uint16)
        block[:, xo[:, None], yo[None, :]] = frames[:, (ys - 1)[None, :], (xs - 1)[:, None]]
        return times, block.reshape(len(times), -1), trunc, nlevels, nres
    iu = np.triu_indices(nres, 1)
    return times, np.ascontiguousarray(frames[:, iu[1], iu[0]]), trunc, nlevels, nres


# ------------------------------------------------------------------------------------------------
# REREAD: matrices/NNNNN.dat of an earlier run instead of a trajectory (conan.py:745-761, :1051-1071, :1455-1494)
# ------------------------------------------------------------------------------------------------
def read_index_folder(reread):
    """index.dat of a REREAD folder -> asymm, nterm, nres, trunc, asymm_keys."""
    asymm, nterm, trunc, nres, keys = 0, 1, 1.5, None, {}
    names = {'NTERM_X': 'xterm', 'NRES_X': 'xres', 'NTERM_Y': 'yterm', 'NRES_Y': 'yres'}
    with open('%s/index.dat' % reread) as f:
        for line in f:
            t = line.split()
            if line.strip().startswith('#') or len(t) < 2:
                continue
            key, val = t[0].upper(), t[1]
            if key == 'ASYMM' and val == 'yes':
                asymm = 1
            elif key == 'NTERM':
                nterm = int(val)
            elif key == 'NRES':
                nres = int(val)
            elif key in names:
                keys[names[key]] = int(val)
            elif key == 'TRUNC':
                trunc = float(val)
    if asymm:
        missing = [k for k in names if names[k] not in keys]
        if missing:
            print("Key not found in index.dat: %s !" % missing[0])
            print("Try again, giving all of the following keys: %s" % " ".join(names))
            sys.exit()
    elif nres is None:
        print("You need to give at least NRES in the index.dat file.")
    return asymm, nterm, nres, trunc, keys


def read_matrix(inputf, xterm, xres, yterm, yres, asymm, trunc):
    """One matrices/NNNNN.dat -> sparse {(i, j): r} with r < trunc (conan.py:1051-1071)."""
    frame = {}
    for line in inputf:
        t = line.split()
        if line.strip().startswith('#') or not t:
            continue
        n1, n2, r = int(t[0]), int(t[1]), float(t[2])
        if xterm <= n1 < xterm + xres and r < trunc and (asymm or n1 < n2):
            frame[(n1, n2)] = r
    return frame


def _levels_from_reread(reread, layout, trunc, asymm):
    """All matrices of a REREAD folder as a level stack.  The distances in those files are multiples of
    the level width of the run that wrote them; the width is recovered from the data."""
    files = []
    for i in range(100000):
        name = "%s/%05i.dat" % (reread, i)
        if os.path.exists(name):
            files.append(name)
    if not files:
        print("Reread folder is empty, will quit now.")
        sys.exit()
    print("Found %i matrices to reread from." % len(files))
    lay = layout
    dense = np.full((len(files), lay.xres, lay.yres), trunc, dtype=np.float64)
    for n, name in enumerate(files):
        a = np.loadtxt(name, ndmin=2)
        if a.size == 0:
            continue
        i, j, r = a[:, 0].astype(int) - lay.xterm, a[:, 1].astype(int) - lay.yterm, a[:, 2]
        ok = (i >= 0) & (i < lay.xres) & (j >= 0) & (j < lay.yres) & (r < trunc)
        if not asymm:
            ok &= i < j
        dense[n, i[ok], j[ok]] = r[ok]
    vals = np.unique(np.round(dense, 6))
    gaps = np.diff(vals)
    gaps = gaps[gaps > 1e-6]
    dx = gaps.min() if len(gaps) else trunc
    nlev = int(round(trunc / dx)) + 1
    for cand in (nlev, 2 * (nlev - 1) + 1, 4 * (nlev - 1) + 1):     # the smallest gap may be several levels wide
        dxc = trunc / (cand - 1)
        if cand <= lv.MAX_NLEVELS and np.abs(vals / dxc - np.rint(vals / dxc)).max() * dxc <= 0.51e-4:
            nlev, dx = cand, dxc
            break
    else:
        print("The REREAD matrices are not multiples of one level width with at most %d levels." % lv.MAX_NLEVELS)
        sys.exit()
    k = np.rint(dense / dx).astype(np.uint16)
    if asymm:
        stack = k.reshape(len(files), -1)
    else:
        iu = np.triu_indices(lay.xres, 1)
        stack = np.ascontiguousarray(k[:, iu[0], iu[1]])
    return stack, nlev


# ------------------------------------------------------------------------------------------------
# stage 2
# ------------------------------------------------------------------------------------------------
class Stage2Result:
    """Everything read_process_frames derived, kept for the analyses that follow it."""
    pass


_last_stage2 = None


def write_matrix(outputf, time, xterm, xres, yterm, yres, trunc, dict_frame, dr, asymm):
    """One matrices/NNNNN.dat from a sparse frame dictionary (conan.py:570-597), for `import conan` users."""
    if not dr:
        outputf.write("# i  j   r_ij\n# r_ij is the shortest distance between any atom of residues  i, j.\n")
        default = trunc
    else:
        outputf.write("# i  j   dr_ij\n# dr_ij is the change (compared to the previous frame)\n"
                      "# in shortest distance between any atom of residues  i, j.\n")
        default = 0.0
    for i in range(xterm, xterm + xres):
        for j in range(yterm, yterm + yres):
            if not asymm:
                r = 0 if i == j else dict_frame.get((i, j) if i < j else (j, i), default)
            else:
                r = dict_frame.get((i, j), default)
            outputf.write("%4d %4d %9.4f\n" % (i, j, r))
        outputf.write("\n")


class LazyVectors:
    """The reference's `vectors` (conan.py:731, :821-826): float16 distances [pairs seen, frames].

    The data lives on the GPU as the level stack; this object gathers the seen pairs' columns, maps
    levels to float16 distances and transposes on the device, and copies the result to the host the first
    time the array is actually looked at (np.asarray(v), v[i], v.T, any ndarray attribute).  A run that
    never touches `vectors` (the command-line path works on the device stack) never pays for the copy."""

    def __init__(self, d_stack, seen, values):
        self._d_stack, self._seen, self._values = d_stack, np.asarray(seen, dtype=np.int64), np.asarray(values)
        self._host = None
        self.shape = (len(self._seen), int(d_stack.shape[0]))
        self.dtype = np.dtype(np.float16)
        self.ndim = 2

    def _materialise(self):
        if self._host is None:
            import torch
            d_stack, dev = self._d_stack, self._d_stack.device
            nsel, nf = self.shape
            out = np.empty((nsel, nf), dtype=np.float16)
            lut = torch.from_numpy(self._values.astype(np.float16)).to(dev)
            d_sel = torch.from_numpy(self._seen).to(dev)
            chunk = max(1, int(2.0e8 // max(nf, 1)))
            for c0 in range(0, nsel, chunk):
                cols = d_stack.view(torch.int16).index_select(1, d_sel[c0:c0 + chunk]).to(torch.int64) & 0xffff
                out[c0:c0 + chunk] = lut[cols].T.contiguous().cpu().numpy()
            self._host = out
            self._d_stack = None
        return self._host

    def __array__(self, dtype=None, copy=None):
        a = self._materialise()
        return a if dtype is None else a.astype(dtype)

    def __len__(self):
        return self.shape[0]

    def __getitem__(self, idx):
        return self._materialise()[idx]

    def __iter__(self):
        return iter(self._materialise())

    def __getattr__(self, name):          # everything else an ndarray offers
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self._materialise(), name)


def _write_matrix_files(levels_host, values, layout, trunc, dr_mode, nframes, symmetric):
    """matrices/%05d.dat (+ _dr.init / _dr.prev) straight from the level stack (conan_write_matrices)."""
    lay = layout
    stack = np.ascontiguousarray(levels_host[:nframes], dtype=np.uint16)
    vals = np.ascontiguousarray(values, dtype=np.float64)
    lib = _lib.load()
    _lib.check(lib.conan_write_matrices(stack.ctypes.data, stack.shape[1], stack.shape[0], 0, 1 if symmetric else 0,
                                        lay.xterm, lay.xres, lay.yterm, lay.yres, vals.ctypes.data, len(vals),
                                        int(dr_mode), None, None, b"matrices", min(32, os.cpu_count() or 1)))


def _stage2_on_shards(d_stack, P, nlevels, lo, hi, eq, frame0, nf_total, rank, world):
    """Stage 2 when the frames are spread over several GPUs: every rank streams its own
    shard, the per-pair sums are merged (conan_b200/sharded.py), the per-frame lines are concatenated, and the
    level shards are collected on rank 0, which writes the tables and runs whatever analysis follows on the
    whole stack (a 100 000-frame stack of 1000 residues is 100 GB: it fits one B200).
    Returns (whole stack on rank 0 / own shard elsewhere, merged statistics as host arrays)."""
    import torch
    import torch.distributed as dist
    from . import sharded
    dev = d_stack.device
    st = sharded.pair_stats_sharded(d_stack, P, nlevels, lo, hi, eq, frame0=frame0)
    stats = st.host()
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    counts[rank] = d_stack.shape[0]
    dist.all_reduce(counts)
    counts = [int(c) for c in counts.cpu()]
    assert sum(counts) == nf_total
    most = max(counts)
    for name in ("native", "nonnative", "ssd_prev", "ssd_first"):
        t = getattr(st, name)
        padded = torch.zeros(most, dtype=t.dtype, device=dev)
        padded[:len(t)] = t
        parts = [torch.empty_like(padded) for _ in range(world)]
        dist.all_gather(parts, padded)
        stats[name] = torch.cat([p[:n] for p, n in zip(parts, counts)]).cpu().numpy()
    whole = torch.empty((nf_total, d_stack.shape[1]), dtype=torch.uint16, device=dev) if rank == 0 else None
    o = 0
    for r in range(world):         # one shard at a time travels to every rank; only rank 0 keeps it
        part = d_stack if r == rank else torch.empty((counts[r], d_stack.shape[1]), dtype=torch.uint16, device=dev)
        dist.broadcast(part.view(torch.uint8), src=r)
        if rank == 0:
            whole[o:o + counts[r]].copy_(part)
        o += counts[r]
        del part
    return (whole if rank == 0 else d_stack), stats


def read_process_frames(inputf, nres, nlevels, nchar, dict_legend, asymm_data, trunc, trunc_inter, trunc_inter_high,
                        begin, end, dt, patch_time, gnus_path, dr_mode, domains, pearson_inter, economy, dimer, reread,
                        stages_list, interact_pair_dict, trunc_lifetime, matrices=True, nterm=1, device="cuda"):
    """Stage 2 with the reference's signature and return tuple (conan.py:697-1023).

    The level stack comes from mdmat_run in this process if it ran, else from the xpm file
    `inputf`.  All per-frame arithmetic happens in conan_pair_stats on the GPU."""
    global _last_stage2
    import torch
    from . import stage1, stage2
    dev = torch.device(device)
    asymm = bool(asymm_data)
    s1 = _last_stage1
    if asymm:
        xdict, ydict, xres, yres, xterm, yterm = asymm_data
        layout = tables.Layout(xterm=xterm, xres=xres, yterm=yterm, yres=yres)
    else:
        layout = tables.Layout(nres=nres, nterm=nterm)
        xres = yres = nres
        xterm = yterm = nterm
    if reread:
        stack_host, nlevels = _levels_from_reread(reread, layout, trunc, asymm)
        d_stack = None
        begin = begin if begin else 0            # conan.py:751-754: REREAD times are begin + n * dt
        dt = dt if dt else 1000
        times_all = begin + dt * np.arange(stack_host.shape[0], dtype=np.float64)
        end = None
    elif s1 is not None and s1.geom.nres == nres and s1.nlevels == nlevels:
        times_all = s1.times
        if asymm and not s1.geom.symmetric:
            stack_host, d_stack = None, s1.levels      # stage 1 already computed exactly this block
        elif asymm:
            xs, ys = sorted(xdict), sorted(ydict)      # xpm residue numbers (1-based) of the X and Y selections
            sq = s1.square_host()
            full = np.full((len(times_all), xres, yres), nlevels - 1, np.uint16)
            full[:, np.array([xdict[m] - xterm for m in xs])[:, None], np.array([ydict[m] - yterm for m in ys])[None, :]] = \
                sq[:, np.array(xs)[:, None] - 1, np.array(ys)[None, :] - 1]
            stack_host, d_stack = full.reshape(len(times_all), -1), None
        else:
            stack_host, d_stack = None, s1.levels
    else:
        if inputf is None:
            print("No dmf.xpm to read and no trajectory processed.")
            sys.exit()
        path = inputf.name if hasattr(inputf, "name") else "dmf.xpm"
        times_all, stack_host, _, _, _ = _levels_from_xpm(path, asymm_data)
        d_stack = None
    # frame filter of conan.py:782-783 (on top of what stage 1 already selected)
    times_all = np.asarray(times_all, dtype=np.float64)
    t_first = times_all[0]
    if not dt:
        dt = times_all[1] - times_all[0] if len(times_all) > 1 else float('-inf')
        if dt > float('-inf'):
            print("Detected dt = %i ps" % dt)
    if not end:
        end = float('Inf')
    if not begin:
        begin = t_first
    if patch_time:
        times_all = times_all[0] + dt * np.arange(len(times_all))
    if dt == float('-inf'):
        keep = np.array([0])
    else:
        keep = np.flatnonzero((times_all >= begin) & (times_all <= end) & (np.mod(times_all - begin, dt) == 0))
    if len(keep) == 0:
        print("No frames left after applying BEGIN/END/DT.")
        sys.exit()
    time_vec = [float(t) for t in times_all[keep]]
    P, stride = layout.P, lv.padded_stride(layout.P)
    rank, world = distributed()
    shard = s1.shard if (s1 is not None and d_stack is not None) else None
    frame0 = 0
    if shard is not None:
        # several GPUs: this rank holds rows shard[0] .. shard[1]-1 of the frame axis; the frame filter above was
        # taken on the whole axis, so every rank agrees on which frames count and on their global numbers
        dev = d_stack.device
        mine = keep[(keep >= shard[0]) & (keep < shard[1])]
        frame0 = int(np.searchsorted(keep, shard[0]))
        if len(mine) != d_stack.shape[0]:
            d_stack = d_stack.view(torch.int16)[torch.from_numpy(mine - shard[0]).to(dev)].contiguous().view(torch.uint16)
        if dimer:
            print("Dimer mode runs on one GPU only; it is switched off in this multi-GPU run.")
            dimer = False
    elif d_stack is not None:
        if len(keep) != d_stack.shape[0]:
            d_stack = d_stack.view(torch.int16)[torch.from_numpy(keep).to(dev)].contiguous().view(torch.uint16)
    else:
        buf = np.zeros((len(keep), stride), np.uint16)
        buf[:, :P] = stack_host[keep]
        d_stack = torch.from_numpy(buf.view(np.int16)).to(dev).view(torch.uint16)
    nf = d_stack.shape[0]

    values = lv.legend_values(trunc, nlevels)
    lo, hi, eq = lv.level_thresholds(values, trunc_inter, trunc_inter_high)
    if dimer and asymm and xres == yres:
        # DIMER (conan.py:553-565): every frame is oriented so that the weighted contacts above the diagonal of
        # its X x Y block are the heavier half; the statistics run on the oriented frames
        lib = _lib.load()
        flipped = torch.zeros(nf, dtype=torch.uint8, device=dev)
        _lib.check(lib.conan_dimer_orient(d_stack.data_ptr(), stride, nf, xres, nlevels, flipped.data_ptr(),
                                          torch.cuda.current_stream(dev).cuda_stream))
        print("Dimer mode: %d of %d frames were transposed." % (int(flipped.sum().item()), nf))
    elif dimer:
        print("'Dimer mode' requires the asymmetric mode with identical X- and Y- residue ranges; it is ignored here.")
        dimer = False
    if shard is not None:
        d_stack, stats = _stage2_on_shards(d_stack, P, nlevels, lo, hi, eq, frame0, len(keep), rank, world)
        if rank != 0:
            return None            # rank 0 holds the merged statistics and the whole stack from here on
        nf = d_stack.shape[0]
    else:
        frac = getattr(_last_stage1, "survivor_frac", None) if _last_stage1 is not None else None
        stats = stage2.pair_stats(d_stack, P, nlevels, lo, hi, eq, sparse=(frac < SPARSE_STACK_BELOW) if frac is not None else None).host()
    if dimer:
        # rmsd_frame's dimer branch (:1420-1428): the smaller of the distance to a frame and to its transpose
        tp = torch.zeros(nf, dtype=torch.int64, device=dev)
        t0 = torch.zeros(nf, dtype=torch.int64, device=dev)
        _lib.check(lib.conan_dimer_ssd(d_stack.data_ptr(), stride, nf, xres, 0, 0, tp.data_ptr(), t0.data_ptr(),
                                       torch.cuda.current_stream(dev).cuda_stream))
        tp, t0 = tp.cpu().numpy(), t0.cpu().numpy()
        tp[0] = stats["ssd_prev"][0]                       # no frame before the first one
        stats["ssd_prev"] = np.minimum(stats["ssd_prev"], tp)
        stats["ssd_first"] = np.minimum(stats["ssd_first"], t0)
    tab = tables.Stage2Tables(stats, values, time_vec, trunc, layout)
    for d in ("aggregate", "matrices"):
        os.makedirs(d, exist_ok=True)
    tab.write_aggregate("aggregate")

    levels_host = None

    def host_levels():
        nonlocal levels_host
        if levels_host is None:
            levels_host = d_stack.view(torch.int16).cpu().numpy().view(np.uint16)[:, :P]
        return levels_host

    if matrices:
        _write_matrix_files(host_levels(), values, layout, trunc, dr_mode, nf, not asymm)
    if nf == 1:
        print("Detected a single frame. Not very dynamic, but we will plot the distances and finish.")
        sys.exit()

    # pairs_list / pairs_legend / vectors (conan.py:806-826): pairs that ever came below trunc
    pi, pj = layout.pair_labels()
    seen = np.flatnonzero(tab.seen)
    pairs_list = [(int(pi[p]), int(pj[p])) for p in seen]
    pairs_legend = {}
    for row, pair in enumerate(pairs_list):
        pairs_legend[pair] = row
        if not asymm:
            pairs_legend[pair[::-1]] = row
    if economy:
        vectors = np.zeros((0, 0), dtype=np.float16)
    else:
        vectors = LazyVectors(d_stack, seen, values)

    if interact_pair_dict:
        life = layout.to_matrix(tab.lifetime, diag=-1.0)
        with open("aggregate/interaction_types.dat", "w") as f:
            for i in range(layout.xres):
                for j in range(layout.yres):
                    if i == j and not asymm:
                        continue
                    pair = (i + xterm, j + yterm)
                    ok = pair in interact_pair_dict and life[i, j] >= trunc_lifetime
                    f.write("%4d %4d %i\n" % (pair[0], pair[1], interact_pair_dict[pair] if ok else 0))
                f.write("\n")

    if stages_list:
        hl = host_levels()
        tv = np.array(time_vec)
        for n, (t1, t2, title) in enumerate(stages_list):
            a, b = np.flatnonzero(tv >= t1), np.flatnonzero(tv >= t2)
            if len(a) == 0 or len(b) == 0:
                print("Did NOT find stage:", title)
                continue
            print("Found stage:", title)
            d = layout.to_matrix(values[hl[b[0]]], 0.0) - layout.to_matrix(values[hl[a[0]]], 0.0)
            with open("matrices/stage_%04d_dr.dat" % (n + 1), "w") as f:
                for i in range(layout.xres):
                    f.write("".join("%4d %4d %9.4f\n" % (i + xterm, j + yterm, d[i, j]) for j in range(layout.yres)))
                    f.write("\n")

    density_r2_metric = {}
    if pearson_inter:
        density_r2_metric = _pearson_inter(d_stack, nlevels, layout, asymm)

    print("number of frames=", len(time_vec))
    avg_d, life_d = tab.metric_dicts()
    res = Stage2Result()
    res.tables, res.layout, res.d_stack, res.values, res.seen = tab, layout, d_stack, values, seen
    res.time_vec, res.nlevels, res.trunc, res.stats = time_vec, nlevels, trunc, stats
    _last_stage2 = res
    if not asymm:
        return (begin, dt, time_vec, pairs_list, pairs_legend, vectors, avg_d, life_d, density_r2_metric) + tab.interact_dicts()
    return (begin, dt, time_vec, pairs_list, pairs_legend, vectors, avg_d, life_d, density_r2_metric) + tab.interact_dicts()


def _pearson_inter(d_stack, nlevels, layout, asymm):
    """Inter-residue correlation of weighted contact numbers (conan.py:669-695, :843-850, :876-887).

    The per-frame, per-residue sums come from conan_residue_weights (exact integers in level units;
    the common factor 1/(nlevels-1) does not change a correlation); the nres x nres correlation of
    those short vectors is a library call on the device."""
    import torch
    from . import stage3
    w = stage3.residue_weights(d_stack, nlevels, not asymm, layout.xres, layout.yres if asymm else 0)   # [nf, n] int32
    out = {}

    def corr_table(vecs, path, term):          # vecs: [nf, n] on the device
        n = vecs.shape[1]
        v = vecs.double()
        varying = (v.max(dim=0).values != v.min(dim=0).values)
        c = torch.corrcoef(v.T) if n > 1 else torch.ones((1, 1), dtype=torch.float64, device=v.device)
        c = torch.where(varying[:, None] & varying[None, :], c, torch.zeros_like(c))
        c = c.cpu().numpy().copy()
        np.fill_diagonal(c, np.where(varying.cpu().numpy(), 1.0, 0.0))
        with open(path, "w") as f:
            for i in range(n):
                f.write("".join("%5i %5i %15.6f\n" % (term + i, term + j, c[i, j]) for j in range(n)))
                f.write("\n")
        return c

    if asymm:
        corr_table(w[:, :layout.xres], "aggregate/pearson_map_density.x.dat", layout.xterm)
        corr_table(w[:, layout.xres:], "aggregate/pearson_map_density.y.dat", layout.yterm)
        return out
    c = corr_table(w, "aggregate/pearson_map_density.dat", layout.xterm)
    n = layout.nres
    for i in range(n):
        for j in range(n):
            if i != j:
                out[(i, j)] = 1 - c[i, j] ** 2
    return out


# ------------------------------------------------------------------------------------------------
# command line
# ------------------------------------------------------------------------------------------------
def main(argv=None):
    argv = sys.argv if argv is None else argv
    print("conan_b200: B200 implementation of the CONAN contact pipeline (no gmx needed for .xtc input)")
    opts = dict(DEFAULT_OPTS)
    if len(argv) != 2:
        print('Usage: You need to provide a single text file in which specify all the options for creating a matrix!')
        sys.exit()
    with open(argv[1]) as finput:
        opts = read_options(finput, opts)
    from .compat import publish_globals
    publish_globals(opts)          # every keyword becomes a module attribute, as conan.py:1748 does
    g = opts
    asymm_data = None
    if g['asymm']:
        print("Asymmetric run found!")
        asymm_data = (g['xdict'], g['ydict'], g['xres'], g['yres'], g['xterm'], g['yterm'])
    interact_pair_dict = None
    trunc = g['trunc']
    trunc_input = trunc
    rank, world = distributed()        # several GPUs when started by torch.distributed.run: frames are sharded
    if rank == 0:
        handle_backups()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    block = None
    if g['asymm'] and 'xdict' in g and 'ydict' in g and not g.get('write_xpm'):
        xs, ys = sorted(g['xdict']), sorted(g['ydict'])
        straight = (xs == list(range(xs[0], xs[0] + len(xs))) and ys == list(range(ys[0], ys[0] + len(ys)))
                    and [g['xdict'][m] - g['xterm'] for m in xs] == list(range(len(xs)))
                    and [g['ydict'][m] - g['yterm'] for m in ys] == list(range(len(ys)))
                    and len(xs) == g['xres'] and len(ys) == g['yres'])
        if straight:
            block = (xs[0] - 1, len(xs), ys[0] - 1, len(ys))
    if g.get('run_mdmat', False) and not g['reread']:
        mdmat_run('', g['traj'], g['coord'], g['nlevels'], trunc, g['mean'], g['begin'], g['end'], g['dt'], g['indexf'],
                  pbc=g.get('pbc', True), group=g.get('group'), write_xpm=g.get('write_xpm'), block=block)
    if g['reread']:
        asymm, nterm_r, nres, trunc, keys = read_index_folder(g['reread'])
        g['nterm'] = nterm_r
        nlevels = nchar = dict_legend = inputf = None
        if asymm:
            print("asymm, xterm, xres, yterm, yres, trunc", asymm, keys['xterm'], keys['xres'], keys['yterm'], keys['yres'], trunc)
            asymm_data = (1, 1, keys['xres'], keys['yres'], keys['xterm'], keys['yterm'])
        else:
            print("asymm, nterm, nres, trunc", asymm, nterm_r, nres, trunc)
    elif _last_stage1 is not None:
        nres, nlevels, nchar = _last_stage1.geom.nres, _last_stage1.nlevels, (1 if _last_stage1.nlevels <= 88 else 2)
        dict_legend, trunc = None, _last_stage1.trunc
        inputf = None
    else:
        with open('dmf.xpm') as f:
            nres, nlevels, nchar, dict_legend, trunc = read_index(f)
        inputf = open('dmf.xpm')
    if rank == 0:
        for d in ('frames', 'matrices', 'aggregate', 'movies', 'input'):
            os.makedirs(d, exist_ok=True)
        shutil.copy(argv[1], 'input')
    nterm = g['nterm']
    if trunc != trunc_input and g['trunc_given']:
        print("Warning! Truncation threshold given is different to the original one. Will use the given one!")
        print("Given:     %9.4f" % trunc_input)
        print("Detected:  %9.4f" % trunc)
        trunc = trunc_input
    trunc_dr, trunc_inter, trunc_inter_high = adjust_truncs(trunc, g['trunc_dr'], g['trunc_inter'], g['trunc_inter_high'])
    print("Will read in the data now...")
    stages_list = None
    if 'stages' in g:
        with open(g['stagef']) as f:
            stages_list = read_stages(f)
    ret = read_process_frames(inputf, nres, nlevels, nchar, dict_legend, asymm_data, trunc, trunc_inter, trunc_inter_high,
                              g['begin'], g['end'], g['dt'], g['patch_time'], g['gnus_path'], g['dr_mode'], g['domains'],
                              g['pearson_inter'], g['economy'], g['dimer'], g['reread'], stages_list, interact_pair_dict,
                              g['trunc_lifetime'], matrices=g['matrices'], nterm=nterm)
    if inputf is not None:
        inputf.close()
    if ret is None:                # a rank other than 0 of a multi-GPU run: its part is done
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
        return None
    # the script-level names the reference's helper functions read
    publish_globals({'nres': nres, 'nlevels': nlevels, 'nchar': nchar, 'dict_legend': dict_legend, 'trunc': trunc, 'nterm': nterm,
                     'trunc_dr': trunc_dr, 'trunc_inter': trunc_inter, 'trunc_inter_high': trunc_inter_high,
                     'time_vec': ret[2], 'pairs_list': ret[3], 'pairs_legend': ret[4], 'vectors': ret[5], 'nframes': len(ret[2])})
    from . import analyses
    g['_trunc_inter_high'] = trunc_inter_high
    analyses.run_all(g, ret, _last_stage2, trunc, trunc_inter, nres, nterm, asymm_data)
    if distributed()[1] > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return ret
