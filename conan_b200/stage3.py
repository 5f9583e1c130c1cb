"""Stage 3 drivers: Pearson sums, covariance and frame-frame distances over the level stack.

Stand where the reference calls scipy.stats.linregress per pair (conan.py:1888-1971), np.cov
(:1983) and scipy pdist (:2063) on its float16 `vectors`.  The float16 values are looked up from
the stack through a table (lut16), the contractions run in conan_pearson_sums (streaming) and
conan_gemm_3xtf32 (tcgen05 tensor cores).
"""
import math

import numpy as np
import torch

from . import _lib

MAX_J = 8


def lut16(values, device):
    """float32 device table of the legend values rounded to float16, as `vectors` stores them (:825)."""
    return torch.from_numpy(np.asarray(values, np.float64).astype(np.float16).astype(np.float32)).to(device)


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def pearson_sums(d_stack, P, nlevels, d_lut, obs32, shift):
    """T [J, P], Q [P] (float64, device) for float32 frame weights / centred observables obs32 [J, nf]."""
    lib = _lib.load()
    if not d_stack.is_cuda:
        raise _lib.ConanError(-1, "stage 3 needs device tensors: there is no CPU fallback")
    dev = d_stack.device
    nf, stride = d_stack.shape
    J = obs32.shape[0]
    d_obs = torch.from_numpy(np.ascontiguousarray(obs32, np.float32)).to(dev)
    d_shift = shift if torch.is_tensor(shift) else torch.from_numpy(np.ascontiguousarray(shift, np.float32)).to(dev)
    T = torch.empty((J, P), dtype=torch.float64, device=dev)
    Q = torch.empty(P, dtype=torch.float64, device=dev)
    _lib.check(lib.conan_pearson_sums(d_stack.data_ptr(), stride, nf, P, d_lut.data_ptr(), nlevels, d_obs.data_ptr(), J,
                                      d_shift.data_ptr(), T.data_ptr(), Q.data_ptr(), _stream(dev)))
    return T, Q


def level_range(d_stack, P, frame_mask=None):
    """(kmin, kmax) int64 host arrays over the (masked) frames (conan_level_range)."""
    lib = _lib.load()
    dev = d_stack.device
    nf, stride = d_stack.shape
    kmin = torch.empty(P, dtype=torch.int32, device=dev)
    kmax = torch.empty(P, dtype=torch.int32, device=dev)
    d_mask = None
    if frame_mask is not None:
        d_mask = torch.from_numpy(np.ascontiguousarray(frame_mask, np.uint8)).to(dev)
    _lib.check(lib.conan_level_range(d_stack.data_ptr(), stride, nf, P, 0 if d_mask is None else d_mask.data_ptr(),
                                     kmin.data_ptr(), kmax.data_ptr(), _stream(dev)))
    return kmin.cpu().numpy().astype(np.int64), kmax.cpu().numpy().astype(np.int64)


def residue_weights(d_stack, nlevels, symmetric, xres, yres=0):
    """int32 device tensor [nf, nres] (or [nf, xres + yres]): weighted contact numbers in level units."""
    lib = _lib.load()
    dev = d_stack.device
    nf, stride = d_stack.shape
    n = xres if symmetric else xres + yres
    out = torch.empty((nf, n), dtype=torch.int32, device=dev)
    _lib.check(lib.conan_residue_weights(d_stack.data_ptr(), stride, nf, 1 if symmetric else 0, xres, yres, nlevels,
                                         out.data_ptr(), _stream(dev)))
    return out


def linregress_all(d_stack, P, nlevels, d_lut, xs, shift, frame_mask=None):
    """scipy.stats.linregress(x_k, series_p) for every pair p and every row x_k of xs [K, nf].

    Returns dict of float64 arrays: r, slope, stderr [K, P], mean and var (population) of every
    series [P].  frame_mask (bool [nf]) restricts all sums to those frames (observables that cover
    only part of the trajectory, conan.py:70-78)."""
    xs = np.atleast_2d(np.asarray(xs, np.float64))
    K, nf = xs.shape
    w = np.ones(nf) if frame_mask is None else np.asarray(frame_mask, np.float64)
    n = w.sum()
    xm = (xs * w).sum(axis=1) / n
    out = {k: np.zeros((K, P)) for k in ("r", "slope", "stderr")}
    for k0 in range(0, K, MAX_J - 1):
        blk = xs[k0:k0 + MAX_J - 1]
        o32 = np.concatenate([w[None], (blk - xm[k0:k0 + len(blk), None]) * w[None]]).astype(np.float32)
        T, Q = pearson_sums(d_stack, P, nlevels, d_lut, o32, shift)
        T, Q = T.cpu().numpy(), Q.cpu().numpy()
        o64 = o32.astype(np.float64)
        S = o64.sum(axis=1)                       # what the rounded centred observables really sum to
        ssy = Q - T[0] ** 2 / n                   # sum (y - ybar)^2
        dev_mean = T[0] / n                       # ybar - shift
        for j in range(len(blk)):
            sxy = T[1 + j] - S[1 + j] * dev_mean
            sxx = (o64[1 + j] ** 2).sum() - S[1 + j] ** 2 / n
            with np.errstate(invalid="ignore", divide="ignore"):
                r = sxy / np.sqrt(sxx * ssy)
                slope = sxy / sxx
                r = np.clip(np.where(np.isfinite(r), r, 0.0), -1.0, 1.0)
                se = np.sqrt(np.maximum((1.0 - r * r) * ssy / sxx / (n - 2), 0.0)) if n > 2 else np.zeros(P)
            out["r"][k0 + j], out["slope"][k0 + j], out["stderr"][k0 + j] = r, np.where(np.isfinite(slope), slope, 0.0), se
        out["var"] = np.maximum(ssy / n, 0.0)
        shift_h = shift.cpu().numpy() if torch.is_tensor(shift) else np.asarray(shift)
        out["mean"] = shift_h.astype(np.float64) + dev_mean
    return out


def operand_scale(d_lut, scale=1.0):
    """Power of two that brings |centred value * scale| near 2^9 (centred values lie within +- the largest
    distance of the legend): the two binary16 parts of the split then sit well inside the half range."""
    bound = float(d_lut.float().max().item()) * float(scale)
    return 2.0 ** math.floor(math.log2(512.0 / max(bound, 1.0e-30)))


def build_operand(d_stack, d_sel, nlevels, d_lut, d_mean, rows_are_frames, scale=1.0):
    """(hi, lo, norm2, s) for conan_gemm_3xtf32: the centred operand times scale * s, split into two binary16
    parts (rows = pairs or frames); s is the power of two from operand_scale, norm2 the squared norms of the
    scaled rows (frame rows only)."""
    lib = _lib.load()
    dev = d_stack.device
    nf, stride = d_stack.shape
    nsel = d_sel.numel()
    rows, k = (nf, nsel) if rows_are_frames else (nsel, nf)
    ld = (k + 63) // 64 * 64
    s = operand_scale(d_lut, scale)
    hi = torch.empty((rows, ld), dtype=torch.float16, device=dev)
    lo = torch.empty((rows, ld), dtype=torch.float16, device=dev)
    norm2 = torch.empty(rows, dtype=torch.float64, device=dev) if rows_are_frames else None
    _lib.check(lib.conan_build_operand(d_stack.data_ptr(), stride, nf, d_sel.data_ptr(), nsel, d_lut.data_ptr(), nlevels,
                                       d_mean.data_ptr(), float(scale) * s, 1 if rows_are_frames else 0, hi.data_ptr(),
                                       lo.data_ptr(), ld, 0 if norm2 is None else norm2.data_ptr(), _stream(dev)))
    return hi, lo, norm2, s


def gemm_3xtf32(a_hi, a_lo, b_hi, b_lo, K, epilogue=0, alpha=1.0, row_norm2=None, col_norm2=None, symmetric=False, out=None):
    """C = epilogue(A B^T) on the tensor cores; operands [rows, ld] binary16 hi / lo parts."""
    lib = _lib.load()
    dev = a_hi.device
    assert a_hi.dtype == torch.float16 and b_hi.dtype == torch.float16
    M, ld = a_hi.shape
    N = b_hi.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=dev)
    p = lambda t: 0 if t is None else t.data_ptr()
    _lib.check(lib.conan_gemm_3xtf32(a_hi.data_ptr(), a_lo.data_ptr(), M, b_hi.data_ptr(), b_lo.data_ptr(), N, K, ld,
                                     epilogue, float(alpha), p(row_norm2), p(col_norm2), 1 if symmetric else 0,
                                     out.data_ptr(), out.stride(0), _stream(dev)))
    return out


def covariance(d_stack, sel, nlevels, d_lut, mean):
    """np.cov(vectors) for the selected pairs (rows), ddof = 1 (conan.py:1983): [nsel, nsel] float32 on the device."""
    dev = d_stack.device
    nf = d_stack.shape[0]
    d_sel = torch.from_numpy(np.ascontiguousarray(sel, np.int32)).to(dev)
    d_mean = torch.from_numpy(np.ascontiguousarray(mean, np.float32)).to(dev)
    hi, lo, _, s = build_operand(d_stack, d_sel, nlevels, d_lut, d_mean, rows_are_frames=False)
    return gemm_3xtf32(hi, lo, hi, lo, nf, epilogue=0, alpha=1.0 / ((nf - 1) * s * s), symmetric=True)


def frame_distances(d_stack, sel, nlevels, d_lut, mean, norm):
    """Full [nf, nf] matrix of pdist(vectors.T / norm) (conan.py:2063), float32 on the device."""
    dev = d_stack.device
    d_sel = torch.from_numpy(np.ascontiguousarray(sel, np.int32)).to(dev)
    d_mean = torch.from_numpy(np.ascontiguousarray(mean, np.float32)).to(dev)
    hi, lo, n2, s = build_operand(d_stack, d_sel, nlevels, d_lut, d_mean, rows_are_frames=True, scale=1.0 / norm)
    return gemm_3xtf32(hi, lo, hi, lo, len(sel), epilogue=1, alpha=1.0 / s, row_norm2=n2, col_norm2=n2, symmetric=True)


def _gram(u, v):
    """u^T v for two [n, m] float64 device blocks (conan_block_gram)."""
    lib = _lib.load()
    g = torch.empty((u.shape[1], v.shape[1]), dtype=torch.float64, device=u.device)
    _lib.check(lib.conan_block_gram(u.data_ptr(), u.stride(0), u.shape[1], v.data_ptr(), v.stride(0), v.shape[1], u.shape[0],
                                    g.data_ptr(), _stream(u.device)))
    return g


def _combine(u, s):
    """u [n, m] times a small s [m, c] (conan_block_combine)."""
    lib = _lib.load()
    s = s.contiguous()
    x = torch.empty((u.shape[0], s.shape[1]), dtype=torch.float64, device=u.device)
    _lib.check(lib.conan_block_combine(u.data_ptr(), u.stride(0), u.shape[1], s.data_ptr(), s.shape[1], u.shape[0],
                                       x.data_ptr(), x.stride(0), _stream(u.device)))
    return x


def _matvec(a32, w):
    """a32 [n, n] float32 times w [n, b] float64, b <= 16 (conan_sym_block_matvec)."""
    lib = _lib.load()
    y = torch.empty_like(w)
    _lib.check(lib.conan_sym_block_matvec(a32.data_ptr(), a32.stride(0), a32.shape[0], w.data_ptr(), w.stride(0), w.shape[1],
                                          y.data_ptr(), y.stride(0), _stream(w.device)))
    return y


def _orthonormalise(w, against=None):
    """Columns of w made orthonormal (and orthogonal to the orthonormal block `against`), through the small Gram
    matrix: w <- w V L^-1/2 with G = w^T w = V L V^T; directions with an eigenvalue below 1e-24 of the largest
    are dropped (the Krylov blocks lose rank as pairs converge).  Two passes, as Gram-based methods need."""
    for _ in range(2):
        if against is not None:
            w = w - _combine(against, _gram(against, w))
        lam, vec = np.linalg.eigh(_gram(w, w).cpu().numpy())   # at most 16 x 16: solved on the host
        keep = lam > 1.0e-24 * max(lam[-1], 1.0e-300)
        if not keep.any():
            return w[:, :0]
        w = _combine(w, torch.from_numpy(np.ascontiguousarray(vec[:, keep] / np.sqrt(lam[keep])[None, :])).to(w.device))
    return w


def top_eigenpairs(cov, k, tol=1.0e-11, block_extra=6, depth=8, max_cycles=40, dense_below=2048):
    """The k largest eigenpairs of a symmetric positive semi-definite float32 device matrix: eigsh(cov, k) of
    conan.py:1985-1993 without factorising the whole matrix.

    Restarted block Krylov iteration with Rayleigh-Ritz in float64.  Per cycle the block of Ritz vectors is
    multiplied `depth` times by the matrix, every new block orthonormalised against all earlier ones; the
    matrix is projected onto the basis and the small problem solved densely.  The products with the matrix
    and with the tall blocks are this library's kernels (conan_sym_block_matvec / conan_block_gram /
    conan_block_combine); the dense problems of at most (depth + 1)(k + block_extra) <= 160 unknowns in between
    are solved on the host (LAPACK through numpy, microseconds).  Stops when the residuals |A v - w v| of the k wanted pairs are below tol * w_max.
    Returns (w [k] descending, v [n, k]), float64 on the device.  Small matrices are solved densely."""
    n = cov.shape[0]
    k = int(min(k, n))
    b = min(k + block_extra, 16)
    if n <= dense_below or b >= n // 4 or (depth + 1) * b > 160 or k > b:
        w, v = torch.linalg.eigh(cov.double())
        return w.flip(0)[:k].contiguous(), v.flip(1)[:, :k].contiguous()
    a = cov if cov.dtype == torch.float32 and cov.is_contiguous() else cov.float().contiguous()
    gen = torch.Generator(device=cov.device)
    gen.manual_seed(20240229)
    x = _orthonormalise(torch.randn((n, b), dtype=torch.float64, device=cov.device, generator=gen))
    theta = res = None
    for _ in range(max_cycles):
        blocks, prods = [x], []
        basis = x
        for _ in range(depth):
            y = _matvec(a, blocks[-1])
            prods.append(y)
            w_blk = _orthonormalise(y, against=basis)
            if w_blk.shape[1] == 0:
                break
            blocks.append(w_blk)
            basis = torch.cat(blocks, dim=1)
        if len(prods) < len(blocks):
            prods.append(_matvec(a, blocks[-1]))
        aq = torch.cat(prods, dim=1)                              # A times the basis, block by block
        t = _gram(basis, aq).cpu().numpy()
        theta_all, s = np.linalg.eigh(0.5 * (t + t.T))             # at most 160 x 160: solved on the host
        nb = min(b, s.shape[1])
        s = torch.from_numpy(np.ascontiguousarray(s[:, ::-1][:, :nb])).to(cov.device)
        theta = torch.from_numpy(np.ascontiguousarray(theta_all[::-1][:nb])).to(cov.device)
        x = _combine(basis, s)
        r = _combine(aq, s) - x * theta[None, :]
        res = torch.sqrt(torch.diagonal(_gram(r, r)))
        if float(res[:k].max().item()) <= tol * float(theta[0].abs().item()):
            break
        x = _orthonormalise(x)
    else:
        raise _lib.ConanError(-4, "top_eigenpairs did not converge (residual %.3g)" % float(res[:k].max().item()))
    return theta[:k].contiguous(), x[:, :k].contiguous()
