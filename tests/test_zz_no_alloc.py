"""conan_mdmat_frames_static: stage 1 for callers that may not allocate device memory.

An analysis hook running inside another program's step loop cannot call cudaMalloc (it would
synchronise the device under the host program), so this entry point keeps its scratch in a fixed
array inside the library.  The test drives it through the C-ABI twice in a row -- a larger system
first, then a smaller one with different residue sizes, so the second call runs over the first
call's leftovers -- and requires the level stacks to equal conan_mdmat_frames' bit for bit.  It
lives in the last-collected test file so a run that stops at the first failure has already run
everything else, and it must also pass on its own in a fresh process.
"""
import gc

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both_entry_points(xyz, boxes, geom, trunc, nlevels, pbc):
    import torch
    from conan_b200 import stage1
    dev = torch.device("cuda")
    nf = xyz.shape[0]
    d_xyz = torch.from_numpy(xyz).to(dev)
    d_box = None if boxes is None else torch.from_numpy(np.ascontiguousarray(boxes, np.float32).reshape(nf, 9)).to(dev)
    ref = torch.zeros((nf, geom.stride), dtype=torch.uint16, device=dev)
    work = torch.empty(stage1.mdmat_workspace_bytes(nf, geom), dtype=torch.uint8, device=dev)
    c_ref = torch.zeros(3, dtype=torch.int64, device=dev)
    stage1.mdmat_frames(d_xyz, d_box, geom, trunc, nlevels, pbc, ref, work, c_ref)
    got = torch.zeros((nf, geom.stride), dtype=torch.uint16, device=dev)
    c_got = torch.zeros(3, dtype=torch.int64, device=dev)
    torch.cuda.synchronize(dev)
    # (tensors left over from earlier tests must not be collected in the middle of the comparison)
    gc.collect()
    gc.disable()
    try:
        before = (torch.cuda.memory_allocated(dev), torch.cuda.mem_get_info(dev)[0])
        stage1.mdmat_frames_static(d_xyz, d_box, geom, trunc, nlevels, pbc, got, c_got)
        torch.cuda.synchronize(dev)
        # neither torch's allocator nor the driver saw an allocation during the call
        assert (torch.cuda.memory_allocated(dev), torch.cuda.mem_get_info(dev)[0]) == before
    finally:
        gc.enable()
    to_np = lambda t: t.view(torch.int16).cpu().numpy().view(np.uint16)[:, :geom.P]
    return to_np(ref), to_np(got), c_ref.cpu().numpy(), c_got.cpu().numpy()


def test_static_scratch_matches_workspace_entry_point_twice(lib, orc):
    from conan_b200 import stage1, synth
    # call 1: uniform residues, symmetric, several tiles
    xyz, res_start, _ = synth.make_trajectory(200, 24, seed=21)
    geom = stage1.Geometry(res_start)
    ref, got, c_ref, c_got = _both_entry_points(xyz, None, geom, 1.0, 1001, False)
    assert np.array_equal(got, ref)
    assert np.array_equal(c_got, c_ref)
    want = orc.condensed(orc.mdmat_levels(xyz, None, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 0))
    assert np.array_equal(got, want)
    # call 2, straight after: smaller system, ragged residue sizes, block mode, periodic box
    xyz2, res_start2, L = synth.make_protein_like(61, 9, seed=22)
    boxes = np.tile(np.eye(3, dtype=np.float32) * np.float32(L), (9, 1, 1))
    geom2 = stage1.Geometry(res_start2, x0=5, xres=30, y0=20, yres=41)
    ref2, got2, c_ref2, c_got2 = _both_entry_points(xyz2, boxes, geom2, 0.9, 501, True)
    assert np.array_equal(got2, ref2)
    assert np.array_equal(c_got2, c_ref2)


def test_static_scratch_rejects_calls_that_do_not_fit(lib):
    import torch
    from conan_b200 import stage1, synth
    from conan_b200._lib import ConanError
    xyz, res_start, _ = synth.make_trajectory(1000, 1, seed=23)
    geom = stage1.Geometry(res_start)
    nf = 400                                   # needs ~98 MB of scratch, the array holds 32 MiB
    assert stage1.mdmat_workspace_bytes(nf, geom) > stage1.STATIC_SCRATCH_BYTES
    d_xyz = torch.from_numpy(np.repeat(xyz, nf, axis=0)).cuda()
    lv = torch.zeros((nf, geom.stride), dtype=torch.uint16, device="cuda")
    with pytest.raises(ConanError) as e:
        stage1.mdmat_frames_static(d_xyz, None, geom, 1.0, 1001, False, lv)
    assert e.value.code == -4 and "scratch" in str(e.value)
    # a call that fits still works afterwards
    stage1.mdmat_frames_static(d_xyz[:100], None, geom, 1.0, 1001, False, lv[:100])
    torch.cuda.synchronize()
