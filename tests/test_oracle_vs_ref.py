"""Pins the C oracle (oracle/mazetd_oracle.c) against (a) the reference's own golden vectors and
(b) the reference's own code compiled in place (oracle/_ref). CPU only."""
import ctypes as C

import numpy as np
import pytest

from mazetd_b200 import synth
from oracle_lib import KAT_SEED, M, ptr, u32p


# ---------------------------------------------------------------------------------------------------
# golden vectors held by the reference's tests
# ---------------------------------------------------------------------------------------------------
def test_prng_known_answers(oracle):
    """externals/cage/sources/test-core/random.cpp:47-78"""
    g = oracle.rng(*KAT_SEED)
    v = [oracle.L.mo_rng_range_i32(C.byref(g), 10, 1000) for _ in range(4)]
    assert v == [730, 955, 866, 863]
    g = oracle.rng(*KAT_SEED)
    f = [oracle.L.mo_rng_range_f(C.byref(g), 10.0, 1000.0) for _ in range(4)]
    # the reference test compares with tolerance 1e-4 relative ("test(v, Vec..)"); we hold float32-exact prints
    assert np.allclose(f, [10.4162464, 931.966248, 645.401428, 848.857239], rtol=1e-7, atol=0)
    # python generator used for synthetic inputs is the same stream
    x = synth.Xoroshiro(*KAT_SEED)
    assert [x.range_u32(10, 1000) for _ in range(4)] == [730, 955, 866, 863]
    s = synth.stream(KAT_SEED[0], KAT_SEED[1], 3, lanes=1)
    g = oracle.rng(*KAT_SEED)
    assert [int(v) for v in s] == [oracle.L.mo_rng_next(C.byref(g)) for _ in range(3)]


def test_numeric_cast_truncation(oracle):
    """externals/cage/sources/test-core/numericTypes.cpp:49-60: float -> int truncates toward zero"""
    assert oracle.L.mo_rtos32(5.1 / 255) in (5,)
    assert oracle.L.mo_rtos32(C.c_float(-5.9 / 255)) == -5
    assert oracle.L.mo_rtos16(C.c_float(0.35)) == 89  # slope threshold, sources/map/grid.cpp:166
    assert oracle.L.mo_rtos16(C.c_float(-7)) == -1785 and oracle.L.mo_rtos16(C.c_float(6)) == 1530


def test_grid_tester_vectors(oracle):
    """GridTester, sources/map/grid.cpp:214-251 (10x5 grid, offset (4,2))"""
    elev = np.zeros(50, np.int16)
    g = oracle.grid(10, 5, 4, 2, elev, None)
    L = oracle.L
    idx = lambda x, y: L.mo_grid_index_i(C.byref(g), x, y)
    idf = lambda x, y: L.mo_grid_index_f(C.byref(g), x, y)
    assert idx(0, 0) == 24 and idx(-4, -2) == 0 and idx(5, -2) == 9 and idx(-4, 2) == 40 and idx(5, 2) == 49
    assert idx(-10, 0) == M and idx(0, 100) == M
    assert idf(0, 0) == 24
    for x, y in [(-0.45, 0), (0.45, 0), (0, -0.45), (0, 0.45)]:
        assert idf(x, y) == 24
    for x, y in [(4.9, 1.9), (5.1, 1.9), (4.9, 2.1), (5.1, 2.1)]:
        assert idf(x, y) == 49

    def pos(i):
        o = (C.c_int32 * 2)()
        L.mo_grid_position(C.byref(g), i, o)
        return (o[0], o[1])

    def center(i):
        o = (C.c_float * 3)()
        L.mo_grid_center(C.byref(g), i, o)
        return tuple(o)

    assert pos(24) == (0, 0) and pos(0) == (-4, -2) and pos(9) == (5, -2) and pos(40) == (-4, 2) and pos(49) == (5, 2)
    assert center(0) == (-4, 0, -2) and center(24) == (0, 0, 0) and center(49) == (5, 0, 2)
    for i in (13, 42):
        assert idx(*pos(i)) == i
        c = center(i)
        assert idf(c[0], c[2]) == i


# ---------------------------------------------------------------------------------------------------
# against the reference's own compiled code
# ---------------------------------------------------------------------------------------------------
def test_ref_passes_its_own_kats(ref):
    out = np.zeros(4, np.int32)
    ref.L.ref_rng_kat_range4i(*KAT_SEED, 10, 1000, out.ctypes.data)
    assert list(out) == [730, 955, 866, 863]
    f = np.zeros(4, np.float32)
    ref.L.ref_rng_kat_range4(*KAT_SEED, 10.0, 1000.0, f.ctypes.data)
    assert np.allclose(f, [10.4162464, 931.966248, 645.401428, 848.857239], rtol=1e-7, atol=0)
    assert ref.L.ref_grid_index_i(10, 5, 4, 2, 0, 0) == 24
    assert ref.L.ref_grid_index_f(10, 5, 4, 2, 4.9, 1.9) == 49


def test_rng_stream_matches_ref(oracle, ref):
    g = oracle.rng(*KAT_SEED)
    ref.L.ref_rng_seed(*KAT_SEED)
    for _ in range(1000):
        assert oracle.L.mo_rng_next(C.byref(g)) == ref.L.ref_rng_next()
    for _ in range(1000):
        assert oracle.L.mo_rng_chance(C.byref(g)) == ref.L.ref_rng_chance()
    for hi in (1, 2, 101 * 101, 2**32 - 1):
        for _ in range(200):
            assert oracle.L.mo_rng_range_u32(C.byref(g), 0, hi) == ref.L.ref_rng_range_u32(0, hi)


def test_fixed_point_matches_ref(oracle, ref):
    rng = np.random.default_rng(1)
    for v in np.concatenate([rng.uniform(-130, 130, 2000), [0, -0.0, 0.35, -7, 1, 6, 100, -100]]).astype(np.float32):
        assert oracle.L.mo_rtos32(v) == ref.L.ref_rtos32(v)
        assert oracle.L.mo_rtos16(v) == ref.L.ref_rtos16(v)
    for v in list(range(-33000, 33000, 7)) + [2**31 - 1, -(2**31), 123456789]:
        assert oracle.L.mo_stor(v) == ref.L.ref_stor(v)


def test_edge_weights_match_ref_all_int16_pairs_sampled(oracle, ref):
    """Grid::neighborDistance (sources/map/grid.cpp:141-145) depends on the operand PAIR, not only on the
    difference: sweep a dense band of pairs plus random full-range int16 pairs."""
    rng = np.random.default_rng(7)
    W, H = 257, 129
    e = rng.integers(-32768, 32768, W * H).astype(np.int16)
    ro, do = oracle.edge_weights(W, H, e)
    rr, dr = ref.edge_weights(W, H, e)
    assert np.array_equal(ro, rr) and np.array_equal(do, dr)
    # small-range band like generated maps (|de| <= 89) and C4's (|de| <= 511)
    for amp in (90, 512):
        e = rng.integers(-amp, amp, W * H).astype(np.int16)
        ro, do = oracle.edge_weights(W, H, e)
        rr, dr = ref.edge_weights(W, H, e)
        assert np.array_equal(ro, rr) and np.array_equal(do, dr)
        assert ro[ro > 0].min() >= 255


@pytest.mark.parametrize("case", ["open", "walls30", "walls45", "maze", "tiny", "row", "col", "blocked_source"])
def test_directions_update_matches_ref(oracle, ref, case):
    if case == "tiny":
        W, H = 1, 1
    elif case == "row":
        W, H = 37, 1
    elif case == "col":
        W, H = 1, 29
    else:
        W, H = 67, 45
    if case == "maze":
        elev, flags = synth.maze_grid(W, H, 3)
    else:
        pct = {"open": 0, "walls30": 30, "walls45": 45}.get(case, 20)
        elev, flags = synth.random_grid(W, H, seed_k=11, wall_pct=pct, elev_mod=512)
    if case == "blocked_source":
        src = int(np.flatnonzero(flags & synth.FLAG_INVALID)[5])  # paths.cpp never tests the source tile itself
    else:
        src = int(synth.pick_open_tiles(flags, 1, 2)[0]) if W * H > 1 else 0
    to, do = oracle.directions_update(W, H, elev, flags, src)
    tr, dr = ref.directions_update(W, H, elev, flags, src)
    assert np.array_equal(do, dr)
    assert np.array_equal(to, tr)
    assert do[src] == 0 and to[src] == M


def test_directions_tiebreak_flat_grid(oracle, ref):
    """all weights equal (flat): maximal number of ties -> the (d[u], u) lexicographic rule is what decides"""
    W, H = 33, 31
    elev = np.zeros(W * H, np.int16)
    flags = np.full(W * H, synth.FLAG_SUN, np.uint16)
    for src in (0, W * H // 2, W * H - 1, W - 1):
        to, do = oracle.directions_update(W, H, elev, flags, src)
        tr, dr = ref.directions_update(W, H, elev, flags, src)
        assert np.array_equal(do, dr) and np.array_equal(to, tr)


@pytest.mark.parametrize("n", [1, 2, 4, 5])
def test_waypoints_update_matches_ref(oracle, ref, n):
    W, H = 41, 37
    elev, flags = synth.random_grid(W, H, seed_k=5, wall_pct=25, elev_mod=200)
    wp = synth.pick_open_tiles(flags, n, 9)
    # make sure the waypoints are mutually reachable; otherwise m-wrapping sums are exercised on purpose below
    o = oracle.waypoints_update(W, H, elev, flags, wp)
    r = ref.waypoints_update(W, H, elev, flags, wp)
    for k in ("tiles", "dist", "fullpath_len", "full_distance", "agg", "fullpath"):
        assert np.array_equal(o[k], r[k]), k


def test_waypoints_find_matches_ref_including_unreachable(oracle, ref):
    W, H = 29, 23
    elev, flags = synth.random_grid(W, H, seed_k=8, wall_pct=38, elev_mod=300)  # fragmented: some waypoints unreachable
    wp = synth.pick_open_tiles(flags, 4, 1)
    o = oracle.waypoints_update(W, H, elev, flags, wp)
    h = ref.map_create(W, H, elev, flags, wp)
    ref.L.ref_waypoints_update(h)
    wv = oracle.waypoints_view(W, H, elev, flags, wp, o["tiles"], o["dist"])
    rng = np.random.default_rng(3)
    for _ in range(600):
        pos = int(rng.integers(0, W * H))
        bits = int(rng.integers(0, 16))
        assert oracle.find(wv, pos, bits) == ref.find(h, pos, bits)
    ref.L.ref_map_destroy(h)


@pytest.mark.parametrize("k", [0, 1, 2, 3, 17])
def test_mapgen_matches_ref(oracle, ref, k):
    """newProcedural -> newGrid -> newWaypoints on seed (s1+k, s2). Noise parity vs upstream FastNoiseLite is
    UNPINNED (oracle/fnl_subset.h); this checks that everything AROUND the noise -- RNG draw order, offset,
    fractal settings via noiseFunction.cpp, rtos16/clamp, flags, components, border, mana, spawn picking and
    Waypoints::update -- matches the reference's compiled code given the same noise restatement."""
    o = oracle.mapgen(KAT_SEED[0] + k, KAT_SEED[1])
    r = ref.mapgen(KAT_SEED[0] + k, KAT_SEED[1])
    for key in ("elev", "flags", "wp_tiles", "dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
        assert np.array_equal(o[key], r[key]), key
    valid = int(((o["flags"] & synth.FLAG_INVALID) == 0).sum())
    assert 500 < valid < 7000  # designer's sizing comment: ~3000 playable tiles (monsterSpawning.cpp:261-263)


def test_grid_stages_match_ref_on_synthetic_elevations(oracle, ref):
    """flags/slope/radius/components/border/mana validated WITHOUT the noise: feed the reference's newGrid a
    Procedural-free path is impossible, so compare the oracle's integer stages against the reference on the
    reference-generated elevations AND check internal consistency on random elevations."""
    r = ref.mapgen(KAT_SEED[0] + 5, KAT_SEED[1])
    flags = oracle.grid_stages(101, 101, 50, 50, 45, r["elev"])
    for t in r["wp_tiles"]:
        flags[t] |= synth.FLAG_WAYPOINT
    assert np.array_equal(flags, r["flags"])


def test_steering_matches_ref(oracle, ref):
    r = ref.mapgen(KAT_SEED[0] + 2, KAT_SEED[1])
    W = H = 101
    h = ref.map_create(W, H, r["elev"], r["flags"], r["wp_tiles"], 50, 50)
    ref.L.ref_waypoints_update(h)
    wv = oracle.waypoints_view(W, H, r["elev"], r["flags"], r["wp_tiles"], r["tiles"], r["dist"], 50, 50)
    spawner = int(r["agg"][2])
    n = 300
    rng = np.random.default_rng(5)
    speeds = rng.uniform(0.02, 0.6, n).astype(np.float32)

    def fresh():
        a = {k: np.zeros(n, np.uint32) for k in ("tile", "tileStart", "tileEnd", "timeStart", "timeEnd", "bits", "timeToArrive", "damage")}
        a["dotflags"] = (np.arange(n) % 4).astype(np.uint8)
        a["alive"] = np.ones(n, np.uint8)
        a["damage"][:] = np.arange(n) % 7 + 1
        return a

    ao, ar = fresh(), fresh()
    for i in range(n):
        so = oracle.monster_spawn(wv, spawner, i, float(speeds[i]))
        sr = ref.monster_spawn(h, i, float(speeds[i]))
        assert np.array_equal(so, sr)
        for a in (ao, ar):
            a["tile"][i] = a["tileStart"][i] = a["tileEnd"][i] = so[0]
            a["bits"][i] = so[1]
            a["timeToArrive"][i] = so[2]
            a["timeStart"][i] = a["timeEnd"][i] = i
    dmg_o = dmg_r = 0
    for t in range(0, 6000):
        dmg_o += oracle.monsters_tick(wv, t, ao)
        dmg_r += ref.monsters_tick(h, t, ar)
        if t % 500 == 0 or t == 5999:
            for k in ao:
                assert np.array_equal(ao[k], ar[k]), (t, k)
            po = oracle.monsters_position(W, H, 50, 50, r["elev"], t, ao["tileStart"], ao["tileEnd"], ao["timeStart"], ao["timeEnd"])
            pr = ref.monsters_position(h, t, ar["tileStart"], ar["tileEnd"], ar["timeStart"], ar["timeEnd"])
            assert np.array_equal(po.view(np.uint32), pr.view(np.uint32))
    assert dmg_o == dmg_r and dmg_o > 0  # some monsters completed all waypoints
    assert (ao["alive"] == 0).any()
    ref.L.ref_map_destroy(h)


def test_position_edge_cases_match_ref(oracle, ref):
    """0/0 -> NaN -> saturate gives 0 (cage-core/math.h:1021-1052); negative and >1 factors clamp"""
    r = ref.mapgen(KAT_SEED[0], KAT_SEED[1])
    h = ref.map_create(101, 101, r["elev"], r["flags"], r["wp_tiles"], 50, 50)
    ok = np.flatnonzero((r["flags"] & synth.FLAG_INVALID) == 0).astype(np.uint32)
    ts = ok[:8].copy()
    te = ok[1:9].copy()
    t0 = np.array([5, 5, 5, 5, 10, 0, 4000000000, 7], np.uint32)
    t1 = np.array([5, 9, 3, 6, 20, 0xFFFFFFFF, 4000000010, 7], np.uint32)
    for gt in (0, 5, 6, 8, 15, 4000000005):
        po = oracle.monsters_position(101, 101, 50, 50, r["elev"], gt, ts, te, t0, t1)
        pr = ref.monsters_position(h, gt, ts, te, t0, t1)
        assert np.array_equal(po.view(np.uint32), pr.view(np.uint32))
    ref.L.ref_map_destroy(h)


def test_sdf_lattice_matches_ref(oracle, ref):
    """Maker::sdf on the marching-cubes lattice: the oracle's restatement against the compiled
    reference Procedural::sdf / newGrid / Grid::index, bit for bit, on a reduced lattice and on one slab of the game's own"""
    for k, res in ((0, (40, 12, 36)), (3, (220, 7, 220))):
        sr, flags = ref.sdf_lattice(KAT_SEED[0] + k, KAT_SEED[1], res=res)
        so = oracle.sdf_lattice(KAT_SEED[0] + k, KAT_SEED[1], 101, 101, 50, 50, flags, res=res)
        assert np.array_equal(so.view(np.uint32), sr.view(np.uint32))
        assert np.isfinite(sr).all() and (sr > 0).any() and (sr < 0).any()
