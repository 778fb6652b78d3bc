"""The torch form of the synthetic input generator (used by bench.py and the full-size GPU tests to build the 16384^2
grid on the device) is bit-identical with the numpy one the goldens were made from."""
import numpy as np
import pytest

from mazetd_b200 import synth


@pytest.mark.parametrize("W,H,k,lanes", [(300, 260, 21, None), (67, 45, 11, None), (513, 257, 3, 4096), (640, 640, 1, None)])
def test_random_grid_torch_matches_numpy(W, H, k, lanes):
    e, f = synth.random_grid(W, H, seed_k=k, wall_pct=30, elev_mod=512, lanes=lanes)
    et, ft = synth.random_grid_torch(W, H, seed_k=k, wall_pct=30, elev_mod=512, lanes=lanes, device="cpu")
    assert np.array_equal(e, et.numpy()) and np.array_equal(f, ft.numpy().view(np.uint16))
    assert np.array_equal(synth.pick_open_tiles(f, 6, 1), synth.pick_open_tiles_torch(ft, 6, 1))


def test_numa_helper_is_a_noop_without_a_gpu_topology():
    """mazetd_b200/numa.py (bench e2e leg at N > 1): without a readable PCI / NUMA topology it must leave the process alone"""
    import os

    from mazetd_b200 import numa

    before = os.sched_getaffinity(0)
    info = numa.bind_to_gpu_node(0)
    assert set(info) == {"node", "cpus", "bound"}
    if not info["bound"]:
        assert os.sched_getaffinity(0) == before
    assert numa._parse_cpulist("0-3,8,10-11") == {0, 1, 2, 3, 8, 10, 11}
    os.sched_setaffinity(0, before)
