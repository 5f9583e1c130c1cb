"""Stage 2 parity on the GPU: conan_pair_stats through the C-ABI vs the CPU oracle, bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

PAIR_KEYS = ("sum_k", "sum_k2", "n_int", "n_enc", "first_f", "last_f", "status")
FRAME_KEYS = ("native", "nonnative", "ssd_prev", "ssd_first")


@pytest.fixture(autouse=True, params=["dense", "sparse"])
def stats_variant(request, monkeypatch):
    """Every test runs under both variants of the streaming pass: the default one and the one that skips 8-frame
    stages of nothing but last levels (conan_pair_stats takes the flag from CONAN_STATS_SPARSE)."""
    monkeypatch.setenv("CONAN_STATS_SPARSE", "1" if request.param == "sparse" else "0")
    return request.param


def _device_stack(levels):
    import torch
    from conan_b200.levels import padded_stride
    nf, P = levels.shape
    stride = padded_stride(P)
    buf = np.full((nf, stride), 0xABCD, dtype=np.uint16)     # padding holds junk on purpose
    buf[:, :P] = levels
    return torch.from_numpy(buf.view(np.int16)).cuda().view(torch.uint16)


def _random_walk_levels(rng, nf, P, nlevels):
    base = rng.uniform(0.1, 1.2, size=P)
    r = np.clip(base + np.cumsum(rng.normal(0, 0.05, size=(nf, P)), axis=0), 0, 1)
    r[:, rng.random(P) < 0.1] = 1.0
    return np.rint(r * (nlevels - 1)).astype(np.uint16)


def _check(got, want, first_frame=0):
    for k in PAIR_KEYS + FRAME_KEYS:
        w = want[k]
        if k in ("first_f", "last_f") and first_frame:
            w = np.where(w >= 0, w + first_frame, w)
        assert np.array_equal(got[k], w), k


@pytest.mark.parametrize("nf,P,nlevels,lo,hi,equal", [
    (200, 2850, 1001, 500, 601, 0),       # ubiquitin-sized, several frame chunks
    (1000, 333, 1001, 500, 601, 0),       # few pairs, many chunks: carries dominate
    (77, 5001, 101, 45, 45, 1),           # equal cut-offs, ragged sizes
    (40, 4096 + 3, 7744, 3000, 5000, 0),  # maximum level count
    (1, 10, 11, 5, 8, 0),                 # a single frame
    (33, 1, 40, 39, 39, 1),               # a single pair, default cut-offs (0.999999*trunc)
])
def test_matches_oracle(lib, orc, nf, P, nlevels, lo, hi, equal):
    from conan_b200 import stage2
    rng = np.random.default_rng(nf * 7 + P)
    levels = _random_walk_levels(rng, nf, P, nlevels)
    want = orc.pair_stats(levels, lo, hi, equal)
    got = stage2.pair_stats(_device_stack(levels), P, nlevels, lo, hi, equal).host()
    _check(got, want)


def test_buffer_only_series_keeps_entering_status(lib, orc):
    """Pairs that sit inside the hysteresis buffer for whole chunks must inherit the carry."""
    from conan_b200 import stage2
    nf, P = 640, 64
    levels = np.full((nf, P), 55, np.uint16)          # buffer is [50, 60]
    levels[0, ::2] = 10                                # half the pairs start on, then stay buffered
    levels[400, ::4] = 90                              # a quarter of those switch off late
    levels[500, 1::4] = 20
    want = orc.pair_stats(levels, 50, 61, 0)
    got = stage2.pair_stats(_device_stack(levels), P, 101, 50, 61, 0).host()
    _check(got, want)
    assert got["n_int"][0] == 400 and got["n_enc"][0] == 1


def test_frame_shards_compose(lib, orc):
    """Two calls over consecutive frame ranges, the second fed the first's state, equal one call."""
    import torch
    from conan_b200 import stage2
    rng = np.random.default_rng(5)
    nf, P, cut = 300, 1500, 137
    levels = _random_walk_levels(rng, nf, P, 1001)
    want = orc.pair_stats(levels, 500, 601, 0)
    d = _device_stack(levels)
    a = stage2.pair_stats(d[:cut], P, 1001, 500, 601, 0)
    b = stage2.pair_stats(d[cut:], P, 1001, 500, 601, 0, frame0=cut, first_row=d[0], prev_row=d[cut - 1],
                          status_in=a.status)
    ha, hb = a.host(), b.host()
    assert np.array_equal(ha["sum_k"] + hb["sum_k"], want["sum_k"])
    assert np.array_equal(ha["sum_k2"] + hb["sum_k2"], want["sum_k2"])
    assert np.array_equal(ha["n_int"] + hb["n_int"], want["n_int"])
    assert np.array_equal(ha["n_enc"] + hb["n_enc"], want["n_enc"])
    assert np.array_equal(np.where(ha["first_f"] >= 0, ha["first_f"], hb["first_f"]), want["first_f"])
    assert np.array_equal(np.maximum(ha["last_f"], hb["last_f"]), want["last_f"])
    assert np.array_equal(hb["status"], want["status"])
    # the shard summary used between GPUs: decided + leaving status of the first shard
    dec, sta = stage2.shard_summary(d[:cut], P, 1001, 500, 601, 0)
    assert np.array_equal(dec.cpu().numpy(), ha["decided"])
    assert np.array_equal(np.where(ha["decided"] != 0, sta.cpu().numpy(), 0), np.where(ha["decided"] != 0, ha["status"], 0))
    for k in FRAME_KEYS:
        assert np.array_equal(np.concatenate([ha[k], hb[k]]), want[k]), k


def test_full_range_jumps_at_the_level_limit(lib, orc):
    """7744 levels, every pair of a warp jumping across the whole range between frames: the split per-frame
    sums (16-bit halves of up to 128 x 7743^2) must not spill into each other."""
    from conan_b200 import stage2
    nf, P, nlevels = 70, 1024 + 128, 7744
    levels = np.zeros((nf, P), np.uint16)
    levels[1::2] = nlevels - 1                       # frame 0 at level 0, then alternating ends of the range
    levels[5, ::3] = 4000
    want = orc.pair_stats(levels, 3000, 5000, 0)
    got = stage2.pair_stats(_device_stack(levels), P, nlevels, 3000, 5000, 0).host()
    _check(got, want)
    assert int(got["ssd_first"][1]) == P * (nlevels - 1) ** 2


@pytest.mark.parametrize("fill", ["last", "none"])
def test_stack_entirely_on_or_off_the_last_level(lib, orc, fill):
    """A stack that is 100 % "beyond trunc" and one that never touches the last level."""
    from conan_b200 import stage2
    rng = np.random.default_rng(11)
    nf, P, nlevels = 150, 2000, 1001
    if fill == "last":
        levels = np.full((nf, P), nlevels - 1, np.uint16)
    else:
        levels = rng.integers(0, nlevels - 1, size=(nf, P)).astype(np.uint16)
    want = orc.pair_stats(levels, 500, 601, 0)
    got = stage2.pair_stats(_device_stack(levels), P, nlevels, 500, 601, 0).host()
    _check(got, want)


def test_misaligned_stack_is_rejected(lib):
    """Rows are streamed with 16-byte bulk copies: a pitch of 4 mod 8 or an 8-byte aligned base is an error."""
    import torch
    from conan_b200 import stage2
    from conan_b200._lib import ConanError
    d = torch.zeros((4, 12), dtype=torch.uint16, device="cuda")
    with pytest.raises(ConanError):
        stage2.pair_stats(d, 10, 101, 50, 60, 0)
    e = torch.zeros(4 * 64 + 4, dtype=torch.uint16, device="cuda")[4:].view(4, 64)
    with pytest.raises(ConanError):
        stage2.pair_stats(e, 10, 101, 50, 60, 0)


def test_bad_arguments_are_errors(lib):
    import torch
    from conan_b200 import stage2
    from conan_b200._lib import ConanError
    d = torch.zeros((4, 64), dtype=torch.uint16, device="cuda")
    with pytest.raises(ConanError):
        stage2.pair_stats(d, 10, 8000, 5, 6, 0)              # too many levels
    with pytest.raises(ConanError):
        stage2.pair_stats(d, 10, 101, 60, 50, 0)             # high cut-off below low cut-off
    with pytest.raises(ConanError):
        stage2.pair_stats(d, 100, 101, 50, 60, 0)            # P larger than the row pitch
    with pytest.raises(ConanError):
        stage2.pair_stats(d.cpu(), 10, 101, 50, 60, 0)       # host tensor: no CPU path


@pytest.mark.parametrize("nlevels,lo,hi", [(1001, 500, 601), (7744, 3000, 5000), (41, 20, 24)])
def test_sparse_block_structured_stack(lib, orc, nlevels, lo, hi):
    """Mostly last levels, with contact patches that come and go: whole 128-pair stretches are quiet for many
    stages, quiet stages sit between busy ones inside a window, patches switch on right after a quiet stage and
    windows end in the middle of quiet runs.  Both variants must give the oracle's numbers."""
    from conan_b200 import stage2
    rng = np.random.default_rng(nlevels)
    nf, P = 1000, 6000
    levels = np.full((nf, P), nlevels - 1, np.uint16)
    for _ in range(60):
        p0, w = int(rng.integers(0, P - 40)), int(rng.integers(1, 40))
        f0, d = int(rng.integers(0, nf - 5)), int(rng.integers(1, 300))
        patch = rng.integers(0, nlevels, size=(min(d, nf - f0), w))
        levels[f0:f0 + d, p0:p0 + w] = patch
    levels[0, 3000:3128] = lo // 2            # on in the first frame only: the status must be dropped, not carried
    levels[8:16, 4096:4224] = nlevels - 1     # (already quiet) one stage of a stretch
    levels[16, 4100] = 0                      # a contact right after a quiet stage
    levels[nf - 1, 5000] = 0                  # and in the last frame
    want = orc.pair_stats(levels, lo, hi, 0)
    got = stage2.pair_stats(_device_stack(levels), P, nlevels, lo, hi, 0).host()
    _check(got, want)
    # the explicit flag gives the same as the environment default
    got2 = stage2.pair_stats(_device_stack(levels), P, nlevels, lo, hi, 0, sparse=True).host()
    _check(got2, want)
