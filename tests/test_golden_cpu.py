"""The C oracle against the committed golden fixtures (tests/golden/*.npz, produced by tests/golden/make_golden.py
from the reference's own compiled code). CPU only; these fixtures are what pins the oracle on the GPU box, where
/root/reference does not exist."""
import os

import numpy as np

from mazetd_b200 import synth
from oracle_lib import KAT_SEED

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_oracle_mapgen_matches_golden(oracle):
    gold = np.load(os.path.join(GOLD, "mapgen_k0_k7.npz"))
    for k in range(8):
        o = oracle.mapgen(KAT_SEED[0] + k, KAT_SEED[1], cap=1024)
        for key in ("elev", "flags", "wp_tiles", "agg", "full_distance", "fullpath_len", "fullpath"):
            assert np.array_equal(o[key], gold[key][k]), (k, key)
        for l in range(4):
            assert int(np.bitwise_xor.reduce(o["dist"][l])) == int(gold["dist_xor"][k, l])
            assert int(o["dist"][l].astype(np.uint64).sum() & 0xFFFFFFFF) == int(gold["dist_sum"][k, l])
            assert int(o["tiles"][l].astype(np.uint64).sum() & 0xFFFFFFFF) == int(gold["tiles_sum"][k, l])


def test_oracle_directions_match_golden(oracle):
    gold = np.load(os.path.join(GOLD, "directions.npz"))
    cases = {
        "walls": (67, 45) + synth.random_grid(67, 45, seed_k=11, wall_pct=30, elev_mod=512),
        "maze": (67, 45) + synth.maze_grid(67, 45, 3),
    }
    for name, (W, H, elev, flags) in cases.items():
        t, d = oracle.directions_update(W, H, elev, flags, int(gold[name + "_src"]))
        assert np.array_equal(d, gold[name + "_dist"]) and np.array_equal(t, gold[name + "_tiles"]), name
    W, H = 300, 260
    elev, flags = synth.random_grid(W, H, seed_k=21, wall_pct=30, elev_mod=512)
    t, d = oracle.directions_update(W, H, elev, flags, int(gold["big_src"]))
    assert int(np.bitwise_xor.reduce(d)) == int(gold["big_dist_xor"])
    assert int(d.astype(np.uint64).sum() & 0xFFFFFFFF) == int(gold["big_dist_sum"])
    assert int(t.astype(np.uint64).sum() & 0xFFFFFFFF) == int(gold["big_tiles_sum"])
    assert np.array_equal(d[100 * W : 101 * W], gold["big_dist_row100"])
    assert np.array_equal(t[100 * W : 101 * W], gold["big_tiles_row100"])
