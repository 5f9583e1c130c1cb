"""Stage 3 parity on the GPU: Pearson sums, 3xTF32 covariance and frame-frame distances.

Floating point: the bar is 1e-5 relative Frobenius error for the matrices and 1e-6 for the
regression results, against float64 numpy / scipy on the same float16 `vectors` the reference
would hold (BASELINE.json north_star).
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _stack(rng, nf, P, nlevels, frac_const=0.1):
    base = rng.uniform(0.1, 1.1, size=P)
    r = np.clip(base + np.cumsum(rng.normal(0, 0.03, size=(nf, P)), axis=0) + 0.2 * np.sin(np.arange(nf) / 17.0)[:, None], 0, 1)
    r[:, rng.random(P) < frac_const] = 1.0
    return np.rint(r * (nlevels - 1)).astype(np.uint16)


def _device(levels):
    import torch
    from conan_b200.levels import padded_stride
    nf, P = levels.shape
    buf = np.zeros((nf, padded_stride(P)), np.uint16)
    buf[:, :P] = levels
    return torch.from_numpy(buf.view(np.int16)).cuda().view(torch.uint16)


def _frob(a, b):
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.parametrize("nf,P,nlevels,trunc", [(300, 700, 1001, 1.0), (97, 130, 101, 1.5), (1000, 64, 7744, 1.2)])
def test_linregress_matches_scipy(lib, nf, P, nlevels, trunc):
    from scipy.stats import linregress
    from conan_b200 import stage3
    from conan_b200.levels import legend_values
    rng = np.random.default_rng(nf)
    levels = _stack(rng, nf, P, nlevels)
    values = legend_values(trunc, nlevels)
    vec = values.astype(np.float16)[levels].T            # the reference's `vectors`
    d = _device(levels)
    lut = stage3.lut16(values, d.device)
    times = np.arange(nf) * 100.0 + 5000.0
    obs = np.stack([times, rng.normal(3.0, 1.0, nf), np.cumsum(rng.normal(0, 1, nf))])
    shift = vec[:, 0].astype(np.float32)
    res = stage3.linregress_all(d, P, nlevels, lut, obs, shift)
    for k in range(3):
        want = np.array([tuple(linregress(obs[k], v.astype(np.float64))[i] for i in (2, 0, 4)) if len(set(v)) > 1 else (0, 0, 0)
                         for v in vec])
        varying = np.array([len(set(v)) > 1 for v in vec])
        for col, name in enumerate(("r", "slope", "stderr")):
            got = np.where(varying, res[name][k], 0.0)
            assert _frob(got, want[:, col]) < 1e-6, (k, name)
    assert np.allclose(res["mean"], vec.astype(np.float64).mean(axis=1), rtol=1e-6, atol=1e-7)
    # a frame mask restricts every sum (observables given for part of the frames)
    mask = rng.random(nf) < 0.6
    res = stage3.linregress_all(d, P, nlevels, lut, obs[1:2], shift, frame_mask=mask)
    want = np.array([linregress(obs[1][mask], v[mask].astype(np.float64))[2] if len(set(v[mask])) > 1 else 0 for v in vec])
    varying = np.array([len(set(v[mask])) > 1 for v in vec])
    assert _frob(np.where(varying, res["r"][0], 0.0), want) < 1e-6


@pytest.fixture(params=["pairs", "single"])
def gemm_kernel(request, monkeypatch):
    """conan_gemm_3xtf32 under both of its kernels: CTA pairs on 256 x 256 tiles (the default for anything wider
    than one 128 x 128 tile) and the single-CTA kernel."""
    monkeypatch.setenv("CONAN_GEMM_CFG", "1" if request.param == "pairs" else "0")
    return request.param


@pytest.mark.parametrize("nf,P,nlevels", [(500, 300, 1001), (130, 77, 101), (2000, 513, 1001)])
def test_covariance_matches_numpy(lib, gemm_kernel, nf, P, nlevels):
    from conan_b200 import stage3
    from conan_b200.levels import legend_values
    rng = np.random.default_rng(P)
    levels = _stack(rng, nf, P, nlevels)
    values = legend_values(1.0, nlevels)
    vec = values.astype(np.float16)[levels].T
    sel = np.flatnonzero((levels < nlevels - 1).any(axis=0)).astype(np.int32)
    want = np.cov(vec[sel].astype(np.float64))
    d = _device(levels)
    lut = stage3.lut16(values, d.device)
    mean = vec[sel].astype(np.float64).mean(axis=1)
    got = stage3.covariance(d, sel, nlevels, lut, mean).cpu().numpy()
    assert got.shape == want.shape
    assert _frob(got, want) < 1e-5, _frob(got, want)
    assert np.allclose(got, got.T)


@pytest.mark.parametrize("nf,P,nlevels", [(400, 350, 1001), (129, 50, 101), (1500, 2000, 1001)])
def test_frame_distances_match_pdist(lib, gemm_kernel, nf, P, nlevels):
    from scipy.spatial.distance import pdist, squareform
    from conan_b200 import stage3
    from conan_b200.levels import legend_values
    rng = np.random.default_rng(nf + P)
    levels = _stack(rng, nf, P, nlevels)
    values = legend_values(1.0, nlevels)
    vec = values.astype(np.float16)[levels].T
    sel = np.flatnonzero((levels < nlevels - 1).any(axis=0)).astype(np.int32)
    norm = 30 / np.sqrt(2.0)
    want = squareform(pdist(vec[sel].T.astype(np.float64) / norm))
    d = _device(levels)
    lut = stage3.lut16(values, d.device)
    got = stage3.frame_distances(d, sel, nlevels, lut, vec[sel].astype(np.float64).mean(axis=1), norm).cpu().numpy()
    assert got.shape == want.shape
    off = ~np.eye(nf, dtype=bool)
    assert _frob(got[off], want[off]) < 1e-5, _frob(got[off], want[off])
    assert np.abs(np.diag(got)).max() < 2e-3      # the diagonal is sqrt of rounding noise, never read by the callers


@pytest.mark.parametrize("M,N,K", [(200, 333, 70), (700, 1100, 1000), (64, 900, 130), (257, 100, 64)])
def test_gemm_general_rectangular(lib, gemm_kernel, M, N, K):
    """A != B, sizes that are not multiples of the tile (several tiles either way, one side narrower than a
    tile), plain scaling epilogue."""
    import torch
    from conan_b200 import stage3
    rng = np.random.default_rng(0)
    a, b = rng.normal(size=(M, K)), rng.normal(size=(N, K))

    def split(x, scale=64.0):
        ld = (x.shape[1] + 63) // 64 * 64
        t = torch.zeros((x.shape[0], ld), dtype=torch.float32)
        t[:, :x.shape[1]] = torch.from_numpy(x.astype(np.float32)) * scale
        hi = t.to(torch.float16)
        lo = (t - hi.float()).to(torch.float16)
        return hi.cuda(), lo.cuda()

    ah, al = split(a)
    bh, bl = split(b)
    got = stage3.gemm_3xtf32(ah, al, bh, bl, K, alpha=0.5 / (64.0 * 64.0)).cpu().numpy()
    want = 0.5 * a.astype(np.float32).astype(np.float64) @ b.astype(np.float32).astype(np.float64).T
    assert _frob(got, want) < 2e-6, _frob(got, want)


def test_top_eigenpairs_matches_dense(lib):
    """Restarted block Krylov solver against torch.linalg.eigh on a covariance-like matrix with a slowly
    decaying spectrum and a near-degenerate pair."""
    import torch
    from conan_b200 import stage3
    dev = torch.device("cuda")
    torch.manual_seed(1)
    n, F = 2600, 900
    scales = torch.logspace(0, -1.5, 30, device=dev, dtype=torch.float64)
    scales[3] = scales[2] * (1 - 1e-5)            # nearly equal eigenvalues
    X = torch.randn(n, 30, device=dev, dtype=torch.float64) @ (torch.randn(30, F, device=dev, dtype=torch.float64) * scales[:, None])
    X += 0.02 * torch.randn(n, F, device=dev, dtype=torch.float64)
    X -= X.mean(1, keepdim=True)
    A = (X @ X.T / (F - 1)).float()
    k = 8
    w, v = stage3.top_eigenpairs(A, k, dense_below=0)
    wd, vd = torch.linalg.eigh(A.double())
    wd, vd = wd.flip(0)[:k], vd.flip(1)[:, :k]
    assert torch.allclose(w, wd, rtol=1e-10, atol=0)
    # eigenvectors up to sign; the near-degenerate pair only as a subspace
    for i in (0, 1, 4, 5, 6, 7):
        d = min((v[:, i] - vd[:, i]).abs().max().item(), (v[:, i] + vd[:, i]).abs().max().item())
        assert d < 1e-7, (i, d)
    P1 = v[:, 2:4] @ v[:, 2:4].T
    P2 = vd[:, 2:4] @ vd[:, 2:4].T
    assert (P1 - P2).abs().max().item() < 1e-6
    # orthonormal and small residual
    assert (v.T @ v - torch.eye(k, device=dev, dtype=torch.float64)).abs().max().item() < 1e-12
    assert torch.linalg.vector_norm(A.double() @ v - v * w[None], dim=0).max().item() < 1e-9 * w[0].item()
