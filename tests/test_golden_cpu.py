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


def _towers(rows):
    """rows: list of dicts with the reference's component fields; defaults are DamageComponent / AttackComponent's (game.h:226-245)"""
    base = dict(tile=0, damage=0, overTime=0, firingPeriod=30, firingRange=5.0, splashRadius=0.0, manaCost=12, invalidClasses=0, element=0, bonus=0, targeting=0, firingDelay=0, mana=0)
    dt = dict(firingRange=np.float32, splashRadius=np.float32, invalidClasses=np.uint8, element=np.uint8, bonus=np.uint8, targeting=np.uint8)
    return {k: np.array([dict(base, **r)[k] for r in rows], dt.get(k, np.uint32)) for k in base}


def test_towers_attack_hand_checked_cases(oracle):
    """"attacks" (sources/game/attacks.cpp:132-216) on an 11 x 11 flat grid (offset 5: tile (x, y) sits at world (x-5, y-5)),
    five standing monsters, ten towers at the origin; every expectation below is worked out by hand from the source"""
    W = H = 11
    T = lambda x, z: (z + 5) * W + (x + 5)  # noqa: E731
    elev = np.zeros(W * H, np.int16)
    mt = np.array([T(2, 0), T(0, 1), T(1, 1), T(2, 1), T(-1, 0)], np.uint32)
    n = len(mt)
    mon = dict(tileStart=mt.copy(), tileEnd=mt.copy(), timeStart=np.zeros(n, np.uint32), timeEnd=np.zeros(n, np.uint32), alive=np.ones(n, np.uint8),
               timeToArrive=np.array([100, 90, 90, 10, 5], np.uint32))
    life = np.array([50, 70, 70, 5, 1], np.int64)
    cls = np.array([0, 0, 0, 0, 1], np.uint8)  # monster 4 is a flier
    o = T(0, 0)
    tw = _towers([
        dict(tile=o, firingRange=2, targeting=1, invalidClasses=1, damage=7, manaCost=0),                      # A Front, no fliers: tta tie 90 -> lowest index 1
        dict(tile=o, firingRange=2, targeting=3, damage=1, manaCost=0),                                        # B Closest: d2 = 1 for monsters 1 and 4 -> 1
        dict(tile=o, firingRange=2, targeting=5, element=1, damage=16, splashRadius=1, manaCost=0),            # C Strongest (life 70 tie -> 1), fire, splash 1 hits 1 and 2
        dict(tile=o, firingRange=2, firingDelay=3),                                                            # D waits
        dict(tile=o, firingRange=2, targeting=1, manaCost=12, mana=5),                                         # E no mana
        dict(tile=o, firingRange=2, targeting=6, bonus=1, damage=4, manaCost=12, mana=30),                     # F Weakest -> 4, damage x2, mana cost x2
        dict(tile=o, firingRange=2, targeting=2, damage=1, manaCost=0),                                        # G Back: largest tta -> 0
        dict(tile=T(-5, -5), firingRange=0.5, targeting=1, damage=1, manaCost=0),                              # H nobody in range
        dict(tile=o, firingRange=2, targeting=4, bonus=3, damage=1, manaCost=0),                               # I range 2 + 4: Farthest -> 3 (d2 = 5)
        dict(tile=o, firingRange=2, targeting=1, bonus=2, damage=1, manaCost=0, firingPeriod=31),              # J Front, fliers allowed -> 4; firing rate: period 31 / 2 = 15
    ])
    dur = np.zeros((5, n), np.uint32)
    dmg = np.zeros((5, n), np.uint32)
    target, hits = oracle.towers_attack(W, H, 5, 5, elev, 0, tw, mon, cls, life, dur, dmg)
    assert target.tolist() == [1, 1, 1, -1, -1, 4, 0, -1, 3, 4]
    assert hits.tolist() == [1, 1, 2, 0, 0, 1, 1, 0, 1, 1]
    assert tw["firingDelay"].tolist() == [30, 30, 30, 2, 0, 30, 30, 0, 30, 15]
    assert tw["mana"].tolist() == [0, 0, 0, 0, 5, 6, 0, 0, 0, 0]
    want = np.zeros((5, n), np.uint32)
    want[0, 1] = 7 + 1      # A, B
    want[0, 4] = 8 + 1      # F, J (Front with fliers allowed: tta 5)
    want[0, 0] = 1          # G
    want[0, 3] = 1          # I
    want[1, 1] = want[1, 2] = 48  # C: fire, damage x3
    assert np.array_equal(dmg, want)
    wd = np.zeros((5, n), np.uint32)
    wd[1, 1] = wd[1, 2] = 80  # sqrt(16) * 20
    assert np.array_equal(dur, wd)
    # dead monsters are not in the index; Random takes candidate number word % count in index order
    mon["alive"][1] = 0
    tw2 = _towers([dict(tile=o, firingRange=2, targeting=0, damage=1, manaCost=0), dict(tile=o, firingRange=2, targeting=3, damage=1, manaCost=0)])
    target, hits = oracle.towers_attack(W, H, 5, 5, elev, 0, tw2, mon, cls, life, dur, dmg, random_words=np.array([5, 0], np.uint64))
    assert target.tolist() == [4, 4]  # candidates 0, 2, 4: 5 % 3 = 2 -> monster 4; closest now monster 4 (d2 = 1)
