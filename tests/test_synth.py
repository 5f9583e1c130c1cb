"""The synthetic benchmark trajectories (conan_b200/synth.py): frame shards of one long run, CPU only."""
import numpy as np


def test_frame_shards_are_ranges_of_the_whole_run():
    from conan_b200 import synth
    whole, rs, L = synth.make_trajectory(40, 300, 8, seed=4)
    for f0, f1 in ((0, 100), (100, 200), (200, 300)):
        part, rs2, L2 = synth.make_trajectory(40, 300, 8, seed=4, frame_range=(f0, f1))
        assert np.array_equal(part, whole[f0:f1]) and np.array_equal(rs, rs2) and L == L2


def test_periodic_modes_leave_the_first_shard_alone():
    """bench.py: with the global modes repeating every shard length, the first shard of an N-GPU run is the one-GPU
    workload itself, and the later shards follow the same breathing profile with their own noise."""
    from conan_b200 import synth
    one, _, _ = synth.make_trajectory(40, 100, 8, seed=4)
    first, _, _ = synth.make_trajectory(40, 400, 8, seed=4, frame_range=(0, 100), mode_period=100)
    assert np.array_equal(one, first)
    third, _, _ = synth.make_trajectory(40, 400, 8, seed=4, frame_range=(200, 300), mode_period=100)
    assert third.shape == one.shape and not np.array_equal(third, one)
    # same radius of gyration profile up to the noise of the per-residue displacements
    rg = lambda x: np.sqrt(((x - x.mean(axis=1, keepdims=True)) ** 2).sum(-1).mean(-1))
    assert np.abs(rg(third) - rg(one)).max() < 0.05 * rg(one).mean()
