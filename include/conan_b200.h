/*
 * conan_b200.h -- C ABI of the B200-native CONAN trajectory contact path.
 *
 * The reference (HITS-MBM/conan) has no FFI: its hot path is Python that shells out to
 * `gmx mdmat` and then loops over the parsed text.  Every entry point below names the
 * reference code (file:line under the reference tree) whose work it takes over; the host
 * Python in conan_b200/ calls these through ctypes and is the only intended caller.
 *
 * Conventions
 *   - plain pointers and sizes only; pointers named d_* are DEVICE pointers, h_* / unnamed
 *     `const uint8_t* data` are HOST pointers.  All buffers are caller-owned.
 *   - no entry point allocates device memory; scratch is passed in (sizes from the
 *     *_workspace_bytes helpers).
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).  Calls are
 *     asynchronous with respect to the host unless stated otherwise.
 *   - return value: 0 = CONAN_OK, negative = error; conan_last_error() gives a
 *     thread-local message.  Nothing here prints or exits (the reference's print +
 *     sys.exit() behaviour stays on the Python side).
 *   - pair index p ("condensed" order):
 *       symmetric : residues i<j of nres, p = i*nres - i*(i+1)/2 + (j-i-1)   (P = nres(nres-1)/2)
 *       block     : x in [0,xres), y in [0,yres), p = x*yres + y            (P = xres*yres)
 *   - a level stack is uint16 [nf][stride], frame-major with a row pitch stride >= P (elements):
 *     levels[f*stride + p] is the xpm level index gmx mdmat would have written for that
 *     residue pair in that frame.  Columns P..stride-1 are padding and never read as data.
 *   - nlevels is limited to 7744 = 88^2, the two-character xpm limit of gmx (docs/index.md:161).
 */
#ifndef CONAN_B200_H
#define CONAN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CONAN_OK 0
#define CONAN_ERR_ARG (-1)    /* bad argument (null pointer, negative size, ...) */
#define CONAN_ERR_CUDA (-2)   /* a CUDA runtime / driver call failed */
#define CONAN_ERR_FORMAT (-3) /* malformed trajectory data */
#define CONAN_ERR_SIZE (-4)   /* problem exceeds a documented limit / workspace too small */
#define CONAN_ERR_IO (-5)     /* a file could not be opened or written */

/* Library version (major*10000 + minor*100 + patch). */
int conan_version(void);
/* Thread-local text of the last error returned on this thread ("" if none). */
const char* conan_last_error(void);

/* ------------------------------------------------------------------------------------ */
/* Host side: .xtc trajectory reading.  Replaces the trajectory reading `gmx mdmat -f`    */
/* does for conan.py:1437-1453 (mdmat_run) including its -b/-e/-dt frame filter, which    */
/* the host applies to the times returned here.                                          */
/* ------------------------------------------------------------------------------------ */

/* Scan an in-memory .xtc.  On return *nframes / *natoms are set.  If `offsets` is non-NULL
 * it receives min(cap, nframes) entries of: byte offset, step, time (ps), box (9 floats,
 * row-major box vectors, nm).  Call once with cap = 0 to size the arrays. */
int conan_xtc_index(const uint8_t* data, size_t nbytes, int64_t cap, int64_t* offsets,
                    int32_t* steps, float* times, float* boxes, int64_t* nframes,
                    int32_t* natoms);

/* Decode `nsel_frames` frames (frame_ids index into `offsets`) on `nthreads` host threads.
 * atom_sel (nsel entries, may be NULL = all atoms) picks the index group while decoding, so
 * xyz_out is [nsel_frames][nsel][3] float32 nm -- the array that is copied to the GPU.
 * xyz_out may be pinned host memory. */
int conan_xtc_decode(const uint8_t* data, size_t nbytes, const int64_t* offsets,
                     const int64_t* frame_ids, int64_t nsel_frames, int32_t natoms,
                     const int32_t* atom_sel, int32_t nsel, float* xyz_out, int32_t nthreads);

/* Write frames as an .xtc (xdr3dfcoord compression with runs of small differences) into `out`
 * (capacity cap bytes); *nbytes gets
 * the size needed/used.  Used to materialise synthetic trajectories as real trajectory files. */
int conan_xtc_encode(const float* xyz, const float* boxes, const int32_t* steps,
                     const float* times, int64_t nframes, int32_t natoms, float precision,
                     uint8_t* out, size_t cap, size_t* nbytes);

/* The same decode on the GPU: the compressed file bytes are copied to the device
 * as they are and every frame is decoded by one warp (the run flags are walked sequentially, the
 * groups of atoms between them in parallel).  Coordinates are bit-identical to conan_xtc_decode's.
 * d_data     the file bytes (or a 4-byte aligned slice of them) on the device, nbytes long.
 * d_offsets  [nframes] byte offset inside d_data of every frame to decode (conan_xtc_index gives
 *            them for the whole file; subtract the slice start).
 * d_slot     [natoms] output row of each atom, -1 = not in the index group; NULL = all atoms
 *            (then nsel == natoms).  d_xyz [nframes][nsel][3] float32 nm.
 * d_status   [nframes] int32 out: 0 = decoded, 1 = bad header, 2 = bad code-size index, 3 = run
 *            past the last atom, 4 = bit stream longer than the frame.  The call only enqueues the
 *            kernel; the caller checks d_status after synchronising. */
int conan_xtc_decode_device(const uint8_t* d_data, int64_t nbytes, const int64_t* d_offsets, int64_t nframes,
                            int32_t natoms, const int32_t* d_slot, int32_t nsel, float* d_xyz,
                            int32_t* d_status, void* stream);

/* The same decode in two passes, for callers that can lend scratch memory; about three times faster on
 * streams whose run flags change often.  Pass 1 walks the run flags of every frame with one thread per
 * frame (the walk is a short serial chain per group of atoms: 32 frames share an issue slot instead of
 * one frame holding a warp) and writes an 8-byte descriptor per group; pass 2 decodes the groups with
 * one thread each.  Same arguments, results and status codes as conan_xtc_decode_device.
 * d_work / work_bytes  16-byte aligned scratch; conan_xtc_decode_workspace_bytes(nframes, natoms)
 *            = nframes * (96 + 8 * natoms) holds all frames at once; with less the frames are taken in
 *            as many rounds as needed (each round costs the latency of one frame's walk, ~0.3 ms per
 *            8000 atoms), with less than one frame's worth the call is refused.
 * Systems of more than 2^21 atoms take the one-pass kernel (the descriptors keep 21 bits of atom index). */
size_t conan_xtc_decode_workspace_bytes(int64_t nframes, int32_t natoms);
int conan_xtc_decode_device_ws(const uint8_t* d_data, int64_t nbytes, const int64_t* d_offsets, int64_t nframes,
                               int32_t natoms, const int32_t* d_slot, int32_t nsel, float* d_xyz,
                               int32_t* d_status, void* d_work, size_t work_bytes, void* stream);

/* The two passes as separate calls, for a pipeline that runs the walk of a batch early -- it needs only the bytes,
 * and it is a latency-bound kernel of few warps that fits beside other kernels (bench.py runs it beside stage 2 of
 * the shard before) -- and the decode of the groups when the coordinate buffer is free.  Arguments as above; the
 * workspace must hold all nframes (conan_xtc_decode_workspace_bytes) and stay untouched between the two calls;
 * d_status is written by the walk.  Systems of more than 2^21 atoms are refused here (CONAN_ERR_SIZE). */
int conan_xtc_walk_device(const uint8_t* d_data, int64_t nbytes, const int64_t* d_offsets, int64_t nframes,
                          int32_t natoms, const int32_t* d_slot, int32_t nsel, float* d_xyz, int32_t* d_status,
                          void* d_work, size_t work_bytes, void* stream);
int conan_xtc_groups_device(const uint8_t* d_data, int64_t nbytes, int64_t nframes, int32_t natoms,
                            const int32_t* d_slot, int32_t nsel, float* d_xyz, const void* d_work,
                            size_t work_bytes, void* stream);

/* ------------------------------------------------------------------------------------ */
/* Stage 1: per-frame residue-residue minimum-distance matrix, quantised to xpm levels.  */
/* Replaces `gmx mdmat -frames` as invoked by conan.py:1437-1453 (:1783-1784) and the    */
/* text parse of its output, conan.py:455-567 (read_index / read_next_time / read_frame). */
/* ------------------------------------------------------------------------------------ */

/* Scratch bytes conan_mdmat_frames needs for nf frames (0 on bad arguments).  max_res_atoms
 * = the largest number of atoms in one residue. */
size_t conan_mdmat_workspace_bytes(int64_t nf, int32_t natoms, int32_t nres, int32_t max_res_atoms);

/*
 * d_xyz       [nf][natoms][3] float32 nm: the index group's atoms only, in trajectory order.
 * d_box       [nf][9] float32 or NULL.  Used only when pbc = 1 (rectangular minimum image with
 *             the box diagonal, as gmx pbc_dx does for rectangular boxes).
 * h_res_start [nres+1] HOST array: atoms res_start[r] .. res_start[r+1]-1 form residue r (gmx
 *             mdmat starts a new residue whenever the atom's residue index changes).
 * symmetric=1: pairs i<j over all nres residues (x0,xres,y0,yres ignored).
 * symmetric=0: block of residues [x0,x0+xres) x [y0,y0+yres) (the START_X/NRES_X/START_Y/NRES_Y
 *             selection of conan.py:541-551).
 * trunc, nlevels: gmx mdmat -t / -nlevels.  level = rint(r*(nlevels-1)/trunc) clamped to
 *             [0, nlevels-1], r = float32 min over atom pairs, truncated at trunc.
 * d_levels    [nf][levels_stride] uint16 out (columns >= P untouched).
 * d_counters  (may be NULL) three device counters that are incremented, never reset:
 *             [0] (frame, pair) results within 2 ulp of a level boundary (those may differ from
 *                 gmx by one level), [1] residue pairs that survived bounding-sphere pruning,
 *             [2] atom-pair distances evaluated for them.
 * d_work      scratch of at least conan_mdmat_workspace_bytes(...) bytes.
 */
int conan_mdmat_frames(const float* d_xyz, const float* d_box, int64_t nf, int32_t natoms,
                       const int32_t* h_res_start, int32_t nres, int32_t symmetric, int32_t x0,
                       int32_t xres, int32_t y0, int32_t yres, float trunc, int32_t nlevels,
                       int32_t pbc, uint16_t* d_levels, int64_t levels_stride,
                       unsigned long long* d_counters, void* d_work, size_t work_bytes,
                       void* stream);

/* Same call split for measurement: phases bit 0 = the coordinate re-layout / bounding-sphere pass,
 * bit 1 = the tile kernel (which needs the workspace left by a bit-0 call on the same frames).
 * conan_mdmat_frames is phases = 3.  bench.py times the two kernels separately with this. */
int conan_mdmat_frames_ex(const float* d_xyz, const float* d_box, int64_t nf, int32_t natoms,
                          const int32_t* h_res_start, int32_t nres, int32_t symmetric, int32_t x0,
                          int32_t xres, int32_t y0, int32_t yres, float trunc, int32_t nlevels,
                          int32_t pbc, uint16_t* d_levels, int64_t levels_stride,
                          unsigned long long* d_counters, void* d_work, size_t work_bytes,
                          void* stream, int32_t phases);

/*
 * conan_mdmat_frames without a workspace argument.  This entry point has no counterpart in the
 * reference; it exists for callers that cannot allocate device memory at all -- for example an
 * on-the-fly analysis hook called from inside a running simulation's step loop, where
 * cudaMalloc / cudaFree would synchronise the device and stall the host program's streams, or
 * where that program's memory pools have already claimed the GPU.  The scratch is a fixed
 * CONAN_MDMAT_STATIC_SCRATCH_BYTES array inside the library (resident on every GPU that loads
 * the library, until the process exits), so a call is limited to as many frames as
 *     conan_mdmat_workspace_bytes(nf, natoms, nres, max_res_atoms) <= CONAN_MDMAT_STATIC_SCRATCH_BYTES
 * allows (about 225 frames of a 1000-residue, 8-atoms-per-residue system); larger calls return
 * CONAN_ERR_SIZE and must be split.  It calls no cudaMalloc, cudaFree or other allocating routine.
 * The first call on each device does the set-up: it looks up the array's device address, raises the
 * shared-memory limit of the tile kernels and makes the runtime load every kernel the call can reach
 * (the periodic and the non-periodic variant, all residue-size variants), so that later calls load no
 * code either; make that first call before the step loop and before starting any stream capture.
 * h_res_start is copied to the device with cudaMemcpyAsync on every call: pass pinned host memory if the
 * call must not wait for the stream, and keep it alive until the stream has consumed it.
 * All calls on a device share the one array, which keeps the
 * previous call's contents (nothing in it needs to start at zero: each call rewrites what it
 * reads); two calls whose GPU work overlaps -- different streams, or different host threads --
 * race on it, so issue them on one stream or otherwise order them.
 * Arguments and results are otherwise exactly those of conan_mdmat_frames.
 */
#define CONAN_MDMAT_STATIC_SCRATCH_BYTES (32u * 1024u * 1024u)
int conan_mdmat_frames_static(const float* d_xyz, const float* d_box, int64_t nf, int32_t natoms,
                              const int32_t* h_res_start, int32_t nres, int32_t symmetric, int32_t x0,
                              int32_t xres, int32_t y0, int32_t yres, float trunc, int32_t nlevels,
                              int32_t pbc, uint16_t* d_levels, int64_t levels_stride,
                              unsigned long long* d_counters, void* stream);

/* ------------------------------------------------------------------------------------ */
/* Stage 2: streaming per-pair statistics over a level stack.  Replaces the per-frame     */
/* loop of read_process_frames, conan.py:782-874, with update_aggregate :599-611,         */
/* update_interaction :24-45, the native / non-native counts :841-854 and rmsd_frame      */
/* :1414-1432 (called from plot_frames :643-644).                                         */
/* ------------------------------------------------------------------------------------ */

/* Per-pair results; every array has P entries and is written, not accumulated.  Two facts the
 * reference tracks separately follow from the exact sums: a pair is in pairs_list (ever below
 * trunc, :528-539) iff sum_k < nf*(nlevels-1), and its series is constant (the len(set(rvec)) > 1
 * test, :1902) iff nf*sum_k2 == sum_k^2. */
typedef struct conan_pair_stats_out {
    int64_t* d_sum_k;    /* sum over frames of the level index            (-> sum r,  :601) */
    int64_t* d_sum_k2;   /* sum of squared level indices                  (-> sum r^2, :602) */
    int32_t* d_n_int;    /* frames with the interaction on                (:605) */
    int32_t* d_n_enc;    /* off -> on transitions                         (:607-610) */
    int32_t* d_first_f;  /* first frame with the interaction on, -1 never (:608-609) */
    int32_t* d_last_f;   /* last frame with the interaction on, -1 never  (:606) */
    uint8_t* d_status;   /* status after the last frame                   (:611) */
    uint8_t* d_decided;  /* optional: 1 if some frame of this call was outside the hysteresis buffer */
} conan_pair_stats_out;

/* Per-frame results; nf entries each, frame numbering local to the call. */
typedef struct conan_frame_stats_out {
    int32_t* d_native;     /* pairs on in this frame that were on in global frame 0 (:852) */
    int32_t* d_nonnative;  /* pairs on in this frame that were not              (:853) */
    int64_t* d_ssd_prev;   /* sum over pairs (k_f - k_{f-1})^2 (rmsd_frame vs previous, :644) */
    int64_t* d_ssd_first;  /* sum over pairs (k_f - k_0)^2     (rmsd_frame vs first,    :643) */
} conan_frame_stats_out;

/* Scratch bytes for conan_pair_stats (0 on bad arguments). */
size_t conan_pair_stats_workspace_bytes(int64_t nf, int64_t P);

/*
 * d_levels  [nf][levels_stride] uint16, 16-byte aligned, levels_stride a multiple of 8 (rows are
 *           streamed with 16-byte bulk copies).
 *           frame0 = global index of the first frame of this call (frame numbers written to
 *           first_f/last_f are global).
 * lo_cnt / hi_first: TRUNC_INTER / TRUNC_INTER_HIGH in level space: r < trunc_inter <=> k < lo_cnt,
 *           r > trunc_inter_high <=> k >= hi_first.  equal_cut = 1 selects the
 *           `trunc_inter == trunc_inter_high` branch of update_interaction (:29-30).
 * d_first_row [P] levels of GLOBAL frame 0 (defines "native" and the rmsd-to-first reference);
 *           NULL = row 0 of d_levels.
 * d_prev_row  [P] levels of the frame preceding this call's first frame; NULL = none (the
 *           first frame then gets ssd_prev 0, as the reference's first frame writes no line).
 * d_status_in [P] interaction status entering this call (frame shards after the first);
 *           NULL = all off.
 * Work is split into frame chunks x pair tiles inside; the status entering each chunk is found
 * by a short backward scan from the chunk boundary, chunk records live in d_work.
 */
int conan_pair_stats(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                     int64_t frame0, int32_t nlevels, int32_t lo_cnt, int32_t hi_first,
                     int32_t equal_cut,
                     const uint16_t* d_first_row, const uint16_t* d_prev_row,
                     const uint8_t* d_status_in, const conan_pair_stats_out* out,
                     const conan_frame_stats_out* fout, void* d_work, size_t work_bytes,
                     void* stream);

/* The same with flags.  CONAN_STATS_SPARSE: stages of eight frames in which all 128 pairs of a warp sit at the last
 * level (beyond trunc), as they did in the frame before, are only counted instead of evaluated.  Same results;
 * faster on sparse, block-structured stacks (an X x Y interface of 200 x 400 residues: 1.28 -> 0.79 ms per 20 000 frames), a few per cent slower on
 * stacks where every stretch of 128 pairs holds something (the symmetric benchmark system), hence opt-in.  Ignored
 * unless the last level lies at or above both cut-offs.  conan_pair_stats takes the flag from the environment
 * variable CONAN_STATS_SPARSE (0 / 1, default 0). */
#define CONAN_STATS_SPARSE 1
int conan_pair_stats_ex(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                     int64_t frame0, int32_t nlevels, int32_t lo_cnt, int32_t hi_first,
                     int32_t equal_cut,
                     const uint16_t* d_first_row, const uint16_t* d_prev_row,
                     const uint8_t* d_status_in, const conan_pair_stats_out* out,
                     const conan_frame_stats_out* fout, void* d_work, size_t work_bytes, int32_t flags,
                     void* stream);

/* Summary of a frame shard for the exchange between GPUs (no counterpart in the reference, which
 * is single-process): d_decided[p] = 1 if some frame of the shard lies outside the hysteresis
 * buffer, and then d_status[p] is the interaction status after the shard's last frame whatever
 * the status entering it.  An exclusive scan of these summaries over the shards gives every
 * shard's d_status_in for conan_pair_stats.  Reads only the few trailing rows it needs. */
int conan_shard_summary(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                        int32_t nlevels, int32_t lo_cnt, int32_t hi_first, int32_t equal_cut,
                        uint8_t* d_decided, uint8_t* d_status, void* stream);

/*
 * DIMER mode (conan.py:553-565 in read_frame, :1420-1428 in rmsd_frame): the stack holds n x n blocks of a
 * homodimer run (X and Y residues carry the same labels).
 * conan_dimer_orient transposes, in place, every frame whose weighted contacts below the diagonal outweigh
 * those above it (weights in exact level units, maxlevel - k); d_flipped [nf] (optional) tells which.  Run it
 * before conan_pair_stats: the reference's statistics are taken on the oriented frames.
 * conan_dimer_ssd gives the transposed twins of conan_pair_stats' per-frame sums: sum over the block of
 * (k_f[x][y] - k_ref[y][x])^2 against the previous frame (d_prev_row for the first frame of the call, NULL =
 * none, result 0) and against global frame 0 (d_first_row, NULL = row 0); the reference reports the smaller
 * of the plain and the transposed distance.
 */
int conan_dimer_orient(uint16_t* d_levels, int64_t levels_stride, int64_t nf, int32_t n, int32_t nlevels,
                       uint8_t* d_flipped, void* stream);
int conan_dimer_ssd(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int32_t n,
                    const uint16_t* d_first_row, const uint16_t* d_prev_row, int64_t* d_ssd_prev_t,
                    int64_t* d_ssd_first_t, void* stream);

/* ------------------------------------------------------------------------------------ */
/* Stage 3: the dense analyses over the level stack.  The float16 `vectors` array of the   */
/* reference (conan.py:731, :821-826) is never materialised: values come from a table of  */
/* the float16 legend values, d_lut [nlevels] (widened to float32).                       */
/* ------------------------------------------------------------------------------------ */

/* Streaming sums behind scipy.stats.linregress of every pair's series against time
 * (conan.py:1888-1919) and against observables (conan.py:1922-1971):
 *     T[j][p] = sum_f O[j][f] * (y_pf - shift[p]),   j = 0..J-1        (float64 out, [J][P])
 *     Q[p]    = sum_f O[0][f] * (y_pf - shift[p])^2                    (float64 out, [P])
 * with y_pf = d_lut[level]; O[0] is the 0/1 frame mask, O[j>0] the centred observables times the
 * mask (all float32 [J][nf]); shift[p] ~ the pair's mean keeps the moments free of cancellation.
 * J <= 8 per call.  r, slope and stderr follow on the host from T, Q and sum_f O[j]^2. */
int conan_pearson_sums(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                       const float* d_lut, int32_t nlevels, const float* d_obs, int32_t J,
                       const float* d_shift, double* d_T, double* d_Q, void* stream);

/* Smallest and largest level of every pair over the frames with d_frame_mask[f] != 0 (NULL = all
 * frames); uint32 [P] each.  A series is constant -- the len(set(rvec)) > 1 tests of conan.py:1902 and
 * :1949 fail -- iff both levels map to the same float16 legend value. */
int conan_level_range(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                      const uint8_t* d_frame_mask, uint32_t* d_kmin, uint32_t* d_kmax, void* stream);

/* Weighted number of contacts of every residue in every frame, the residue_column of conan.py:843-850
 * (weight_contacts :19-21), in level units: d_out [nf][n] int32 with n = nres (symmetric) or xres + yres,
 * out[f][r] = sum over the pairs of residue r of (nlevels-1 - level).  Feeds PEARSON_INTER (:669-695). */
int conan_residue_weights(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int32_t symmetric,
                          int32_t xres, int32_t yres, int32_t nlevels, int32_t* d_out, void* stream);

/* Builds a GEMM operand from the stack, centred, scaled and split into hi / lo parts (x = hi + lo).
 * rows_are_frames = 0: rows = the nsel selected pairs d_sel (indices into the P columns), K = frames,
 *     x = (y - d_mean[row]) * scale             -- the operand of np.cov(vectors), conan.py:1983.
 * rows_are_frames = 1: rows = frames, K = selected pairs, x = (y - d_mean[col]) * scale, and
 *     d_norm2[row] = |x|^2 (float64)            -- the operand of pdist(vectors.T / norm), conan.py:2063
 *     (centring on the mean frame leaves distances unchanged and keeps the Gram form accurate).
 * scale multiplies every centred value (1/norm of conan.py:2038-2041 for frame rows, times a power of
 * two chosen by the caller so that |value * scale| stays near 2^9: both parts of the split are binary16).
 * d_hi / d_lo are [rows][ld] binary16 (hi = half(x), lo = half(x - hi)), ld a multiple of 64 >= K;
 * the padding is written as zeros.  d_norm2 is the squared norm of the scaled rows. */
int conan_build_operand(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, const int32_t* d_sel,
                        int32_t nsel, const float* d_lut, int32_t nlevels, const float* d_mean, float scale,
                        int32_t rows_are_frames, uint16_t* d_hi, uint16_t* d_lo, int64_t ld, double* d_norm2,
                        void* stream);

/* C [M][N] (float32, row pitch ldc) = epilogue(A * B^T), A = a_hi + a_lo [M][ld], B = b_hi + b_lo [N][ld]
 * (binary16 parts), as hi*hi + hi*lo + lo*hi on the tcgen05 tensor cores: the three-product split known as
 * 3xTF32 -- the parts carry the same 11-bit significands -- stored as binary16 and issued as kind::f16 MMAs,
 * float32 accumulation in TMEM.
 * epilogue 0: alpha * acc                      (covariance: alpha = 1/((nf-1) s^2), conan.py:1983; s = the
 *                                               operand's power-of-two scale)
 * epilogue 1: alpha * sqrt(max(0, row_norm2[m] + col_norm2[n] - 2 acc))   (frame-frame distance,
 *                                               conan.py:2063; alpha = 1/s)
 * symmetric != 0 (needs A == B): only the upper triangle of tiles is computed, both halves written.
 * Two kernels: CTA pairs on 256 x 256 tiles (tcgen05.mma.cta_group::2; each CTA stages its own half of both
 * operands, so the pair does twice the MMA work per staged byte) whenever the problem is wider than one
 * 128 x 128 tile, else one CTA per 128 x 128 tile; the environment variable CONAN_GEMM_CFG (0 / 1) forces the
 * single-CTA / the pair kernel.  M, N and K are 32-bit; K (padded to 64) must stay below 2^31. */
int conan_gemm_3xtf32(const uint16_t* d_a_hi, const uint16_t* d_a_lo, int32_t M, const uint16_t* d_b_hi,
                      const uint16_t* d_b_lo, int32_t N, int64_t K, int64_t ld, int32_t epilogue, float alpha,
                      const double* d_row_norm2, const double* d_col_norm2, int32_t symmetric, float* d_c,
                      int64_t ldc, void* stream);

/* The block products of the top-k eigensolver that stands where the reference calls eigsh(covmat, k)
 * (conan.py:1985-1993).  Blocks of vectors are float64 [n][ld] row-major.  K, M, N of
 * conan_gemm_3xtf32 and all sizes here must fit 31 bits.
 *   conan_sym_block_matvec   Y [n][b] = A W, A float32 [n][lda] (the covariance conan_gemm_3xtf32 wrote),
 *                            b <= 16; float64 multiply-adds.  Reads A once: 4 n^2 bytes from HBM per product.
 *   conan_block_gram         G [mu][mv] = U^T V (orthogonalisation and Rayleigh-Ritz), mu, mv <= 160
 *   conan_block_combine      X [n][c] = U [n][m] S [m][c] (rotation of a basis by a small matrix), m, c <= 160,
 *                            X must not alias U
 */
int conan_sym_block_matvec(const float* d_A, int64_t lda, int32_t n, const double* d_W, int64_t ldw, int32_t b,
                           double* d_Y, int64_t ldy, void* stream);
int conan_block_gram(const double* d_U, int64_t ldu, int32_t mu, const double* d_V, int64_t ldv, int32_t mv,
                     int32_t n, double* d_G, void* stream);
int conan_block_combine(const double* d_U, int64_t ldu, int32_t m, const double* d_S, int32_t c, int32_t n,
                        double* d_X, int64_t ldx, void* stream);

/* ------------------------------------------------------------------------------------ */
/* Text writers of the file contract (host).  HOST pointers.         */
/* ------------------------------------------------------------------------------------ */

/* matrices/NNNNN.dat exactly as write_matrix prints them (conan.py:570-597: two comment lines, then
 * "%4d %4d %9.4f" per residue pair, a blank line after each x residue), straight from the level
 * stack, frames spread over nthreads host threads.
 * h_levels    [nf][levels_stride] uint16, pair order as conan_mdmat_frames writes it (symmetric:
 *             condensed i<j, the diagonal prints 0; else the xres x yres block, x-major).
 * file_index0 number of the first file (frame f goes to <directory>/<file_index0+f, %05d>.dat).
 * values      [nlevels] legend values (float64); values[level] is the printed distance.
 * dr_mode     bit 0: also NNNNN_dr.init.dat (change since frame 0), bit 1: NNNNN_dr.prev.dat (change
 *             since the previous frame), conan.py:829-838; frame 0 of the run gets neither.
 * h_first_row / h_prev_row: rows of the run's frame 0 and of the frame before this batch, needed
 *             only when dr_mode != 0 and file_index0 > 0. */
int conan_write_matrices(const uint16_t* h_levels, int64_t levels_stride, int64_t nf, int64_t file_index0,
                         int32_t symmetric, int32_t xterm, int32_t xres, int32_t yterm, int32_t yres,
                         const double* values, int32_t nlevels, int32_t dr_mode, const uint16_t* h_first_row,
                         const uint16_t* h_prev_row, const char* directory, int32_t nthreads);

/* The multi-frame dmf.xpm `gmx mdmat -frames` leaves behind and read_index / read_next_time /
 * read_frame parse (conan.py:455-567): per frame the comment header with `title: "t=<ps> ps"`, the
 * legend block, the axis comments and yres rows (last y residue first) of xres codes.
 * legend_block, axes_block: the text between the header and the rows (identical in every frame).
 * codes       [nlevels][nchar] characters, nchar = 1 or 2.
 * append != 0 adds the frames to an existing file (batches of one run). */
int conan_write_xpm(const uint16_t* h_levels, int64_t levels_stride, int64_t nf, int32_t symmetric, int32_t xres,
                    int32_t yres, const double* times, const char* legend_block, const char* axes_block,
                    const char* codes, int32_t nchar, int32_t nlevels, const char* path, int32_t append,
                    int32_t nthreads);

/* A residue-pair table: `header`, then one row per pair (x residue i outer, y residue j inner) and a
 * blank line after each i -- the layout of aggregate/timeline.dat and mdmat_average_rmsf.dat
 * (conan.py:917, :926), pearson_data.dat (:1916), pca/pc.N.dat (:428), the cluster maps (:1316-1326).
 * row_format is the reference's printf row format: two integer conversions for the residue numbers
 * (xterm + i, yterm + j), then one conversion per column (%d / %i columns print the value truncated to
 * an integer, as Python's % does with floats), ending in "\n".  columns[c] is [xres * yres] float64,
 * x-major.  Rows are formatted on nthreads host threads. */
int conan_write_pair_table(const char* path, const char* header, int32_t xterm, int32_t xres, int32_t yterm,
                           int32_t yres, const double* const* columns, int32_t ncol, const char* row_format,
                           int32_t nthreads);

/* ------------------------------------------------------------------------------------ */
/* Measurement helpers (bench.py only).                                                  */
/* ------------------------------------------------------------------------------------ */

/* Runs a dependent-FMA kernel on every SM and returns the achieved FP32 TFLOP/s
 * (2 flops per FMA) in *tflops; used as the roofline denominator of stage 1. */
int conan_measure_fp32_peak(double* tflops, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CONAN_B200_H */
