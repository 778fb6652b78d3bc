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


def test_dots_tick_hand_checked_cases(oracle):
    """"damage monsters" (sources/game/monsters.cpp:101-183), small cases worked out by hand from the source"""
    n = 5
    dur = np.zeros((5, n), np.uint32)
    dmg = np.zeros((5, n), np.uint32)
    life = np.full(n, 1000, np.int64)
    alive = np.ones(n, np.uint8)
    resist = np.zeros(n, np.uint8)
    money = np.arange(1, n + 1, dtype=np.uint32) * 10
    # 0: plain physical dot 100 over 4 ticks -> 25 now, 75 left, duration 3
    dur[0, 0], dmg[0, 0] = 4, 100
    # 1: fire + water both active -> both cancelled, durations forced to 30 then decremented to 29, no damage
    dur[1, 1], dmg[1, 1], dur[2, 1], dmg[2, 1] = 5, 500, 2, 300
    # 2: fire dot strengthened by magic (x2) and resisted (/4): 80/2 = 40 -> 80 -> 20; magic itself 10/1... duration 1 -> dmg 10, super by magic? magic strengthens itself: 10 x2 = 20
    dur[1, 2], dmg[1, 2], dur[4, 2], dmg[4, 2] = 2, 80, 1, 10
    resist[2] = 1 << 1
    # 3: lethal: life 30, physical 64 with duration 0 (instant) -> dies, money 40
    dmg[0, 3] = 64
    life[3] = 30
    # 4: dead already: untouched
    alive[4] = 0
    dmg[0, 4] = 999
    ms, killed = oracle.monsters_dots_tick(dur, dmg, life, alive, resist, money)
    assert (dur[0, 0], dmg[0, 0], life[0]) == (3, 75, 975)
    assert (dur[1, 1], dmg[1, 1], dur[2, 1], dmg[2, 1], life[1]) == (29, 0, 29, 0, 1000)
    # monster 2: fire first (magic still holds 10 > 0 -> super): 40 -> 80 -> /4 = 20; then magic: 10, super by itself (10 - 10 = 0 left, but `super` was read before the subtraction) -> 20
    assert (dur[1, 2], dmg[1, 2], dur[4, 2], dmg[4, 2], life[2]) == (1, 40, 0, 0, 1000 - 20 - 20)
    assert (life[3], alive[3]) == (30 - 64, 0) and (ms, killed) == (40, 1)
    assert (dmg[0, 4], life[4], alive[4]) == (999, 1000, 0)


GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_oracle_reproduces_sdf_golden(oracle):
    """tests/golden/sdf_lattice.npz was written by the compiled reference pieces (make_golden.py: ref_sdf_lattice)"""
    g = np.load(os.path.join(GOLD, "sdf_lattice.npz"))
    got = oracle.sdf_lattice(KAT_SEED[0] + int(g["seed_k"]), KAT_SEED[1], 101, 101, 50, 50, g["flags"], res=tuple(int(v) for v in g["res"]))
    assert np.array_equal(got.view(np.uint32), g["sdf"].view(np.uint32))


def test_oracle_reproduces_game_events_golden(oracle):
    """the scripted place/remove workload at reduced size ends in the committed state (flags, monsters, layers)"""
    from mazetd_b200 import game_events
    from test_gpu_parity import _OracleGame

    g = np.load(os.path.join(GOLD, "game_events.npz"))
    o = oracle.mapgen(KAT_SEED[0], KAT_SEED[1])
    mon = {k[5:]: g[k].copy() for k in g.files if k.startswith("mon0_")}
    og = _OracleGame(oracle, 101, 101, o["elev"], o["flags"], o["wp_tiles"], mon)
    stats = game_events.run_events(og, 300)
    assert [stats[k] for k in ("placed", "cleared", "rejected_flags", "rejected_trial")] == list(g["stats"])
    assert np.array_equal(og.flags, g["flags"])
    for k in ("tile", "tileStart", "tileEnd", "timeStart", "timeEnd", "bits", "alive"):
        assert np.array_equal(og.a[k], g["mon_" + k]), k
    assert np.array_equal(og.w["agg"], g["agg"]) and np.array_equal(og.w["full_distance"], g["full_distance"])
