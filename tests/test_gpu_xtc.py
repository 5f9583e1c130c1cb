"""The GPU decoders -- conan_xtc_decode_device_ws (two passes: one thread per frame walks the run flags, one
thread per group of atoms decodes) and conan_xtc_decode_device (one warp per frame) -- against the host decoder
and the oracle: bit-identical coordinates on the reference's own trajectories and on synthetic compressed ones.
Every test runs with the two-pass decoder holding all frames at once, with scratch for three frames only (the
frames are then taken in rounds) and with the one-pass decoder."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch


@pytest.fixture(autouse=True, params=["two_pass", "two_pass_rounds", "one_pass"])
def decoder(request, monkeypatch):
    from conan_b200.xtc import XtcFile
    plain = XtcFile.decode_device

    def decode_device(self, *a, **k):
        if request.param == "one_pass":
            k["one_pass"] = True
        elif request.param == "two_pass_rounds":
            k["work_bytes"] = 3 * (96 + 8 * self.natoms)
        return plain(self, *a, **k)
    monkeypatch.setattr(XtcFile, "decode_device", decode_device)
    return request.param


@pytest.mark.parametrize("name", ["sample.xtc", "ubi_12.xtc"])
def test_reference_trajectories(torch_cuda, name):
    from conan_b200.xtc import XtcFile
    from oracle import oracle
    path = os.path.join(HERE, "golden", name)
    x = XtcFile(path)
    host = x.decode()
    dev = x.decode_device().cpu().numpy()
    assert np.array_equal(dev, host)
    assert np.array_equal(dev, oracle.xtc_decode(open(path, "rb").read())[2])
    # a frame subset in arbitrary order and an atom selection
    ids = np.array([x.nframes - 1, 0, 3, 3, 7])
    sel = np.arange(5, x.natoms, 3, dtype=np.int32)
    assert np.array_equal(x.decode_device(frame_ids=ids, atom_sel=sel).cpu().numpy(), host[ids][:, sel])


def test_synthetic_compressed(torch_cuda, tmp_path):
    from conan_b200 import synth
    from conan_b200.xtc import XtcFile, write_xtc
    xyz, _, L = synth.make_trajectory(300, 700, seed=4)
    p = str(tmp_path / "c3.xtc")
    write_xtc(p, xyz, boxes=np.tile(np.eye(3) * L, (700, 1, 1)), times=np.arange(700) * 100.0)
    x = XtcFile(p)
    host = x.decode()
    assert np.abs(host - xyz).max() < 1.5e-6
    assert np.array_equal(x.decode_device().cpu().numpy(), host)
    sel = np.flatnonzero(np.arange(x.natoms) % 8 < 5).astype(np.int32)
    ids = np.arange(3, 700, 7)
    assert np.array_equal(x.decode_device(frame_ids=ids, atom_sel=sel).cpu().numpy(), host[ids][:, sel])


def test_small_and_wide_systems(torch_cuda, tmp_path):
    from conan_b200.xtc import XtcFile, write_xtc
    rng = np.random.default_rng(1)
    cases = [(3, 9.0, 1000.0), (9, 9.0, 1000.0), (10, 9.0, 1000.0), (257, 9.0, 1000.0),
             (40, 200.0, 100000.0), (40, 100.0, 100000.0)]
    for natoms, width, precision in cases:
        xyz = np.round(rng.uniform(-2, width, size=(5, natoms, 3)) * 1000) / 1000
        p = str(tmp_path / "t.xtc")
        write_xtc(p, xyz, precision=precision)
        x = XtcFile(p)
        assert np.array_equal(x.decode_device().cpu().numpy(), x.decode()), (natoms, width)


def test_malformed_frames_are_reported(torch_cuda, tmp_path):
    from conan_b200 import _lib, synth
    from conan_b200.xtc import XtcFile, write_xtc
    xyz, _, L = synth.make_protein_like(30, 6, seed=2)
    p = str(tmp_path / "m.xtc")
    write_xtc(p, xyz)
    data = bytearray(open(p, "rb").read())
    x = XtcFile(bytes(data))
    # break the code-size index of frame 2 (header field at +84): the index scan does not look at it
    off = int(x.offsets[2])
    data[off + 84:off + 88] = (200).to_bytes(4, "big")
    bad = XtcFile(bytes(data))
    with pytest.raises(_lib.ConanError) as e:
        bad.decode_device()
    assert "frame 2" in str(e.value)
    # the other frames still decode
    ids = np.array([0, 1, 3, 4, 5])
    assert np.array_equal(bad.decode_device(frame_ids=ids).cpu().numpy(), x.decode(frame_ids=ids))


def test_random_systems(torch_cuda, tmp_path):
    """Clustered / spread / wide random systems at several precisions: the GPU decoder equals the host decoder."""
    from conan_b200.xtc import XtcFile, write_xtc
    rng = np.random.default_rng(11)
    for case in range(24):
        natoms = int(rng.integers(10, 400))
        nframes = int(rng.integers(1, 40))
        spread = [0.02, 0.3, 3.0, 60.0][case % 4]
        cluster = [0.0, 0.05, 0.4][case % 3]
        precision = [10.0, 1000.0, 100000.0][(case // 4) % 3]
        centres = rng.uniform(-spread, spread, size=(nframes, natoms // 5 + 1, 3))
        xyz = (centres[:, np.arange(natoms) // 5] + rng.normal(0, cluster, size=(nframes, natoms, 3))).astype(np.float32)
        p = str(tmp_path / "r.xtc")
        write_xtc(p, xyz, precision=precision)
        x = XtcFile(p)
        assert np.array_equal(x.decode_device().cpu().numpy(), x.decode()), (case, natoms, nframes, spread, cluster, precision)


def test_walk_and_groups_as_separate_calls(torch_cuda, tmp_path):
    """conan_xtc_walk_device + conan_xtc_groups_device (the two passes apart, as a pipeline runs them) give what the
    one call gives; a workspace that cannot hold all frames is refused."""
    import torch
    from conan_b200 import _lib, synth
    from conan_b200.xtc import XtcFile, write_xtc
    lib = _lib.load()
    xyz, _, L = synth.make_trajectory(120, 90, seed=9)
    p = str(tmp_path / "w.xtc")
    write_xtc(p, xyz)
    x = XtcFile(p)
    want = x.decode()
    dev = torch.device("cuda")
    d_data = torch.from_numpy(np.array(x._buf)).to(dev)
    d_off = torch.from_numpy(x.offsets).to(dev)
    nf, na = x.nframes, x.natoms
    need = int(lib.conan_xtc_decode_workspace_bytes(nf, na))
    d_work = torch.empty(need, dtype=torch.uint8, device=dev)
    out = torch.zeros((nf, na, 3), dtype=torch.float32, device=dev)
    status = torch.full((nf,), -1, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(lib.conan_xtc_walk_device(d_data.data_ptr(), d_data.numel(), d_off.data_ptr(), nf, na, None, na, out.data_ptr(),
                                         status.data_ptr(), d_work.data_ptr(), d_work.numel(), st))
    torch.cuda.synchronize()
    assert int(status.abs().sum().item()) == 0 and float(out.abs().sum().item()) == 0.0      # nothing decoded yet
    _lib.check(lib.conan_xtc_groups_device(d_data.data_ptr(), d_data.numel(), nf, na, None, na, out.data_ptr(),
                                           d_work.data_ptr(), d_work.numel(), st))
    assert np.array_equal(out.cpu().numpy(), want)
    with pytest.raises(_lib.ConanError):
        _lib.check(lib.conan_xtc_walk_device(d_data.data_ptr(), d_data.numel(), d_off.data_ptr(), nf, na, None, na, out.data_ptr(),
                                             status.data_ptr(), d_work.data_ptr(), need // 2, st))
