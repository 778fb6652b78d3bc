"""Drop-in test of the host adapter (mazetd_b200/host/paths_b200.cpp + grid_b200.cpp).

oracle/_ref/libmazetd_adapter.so links the adapter's replacement bodies of Directions::update, Waypoints::update,
Waypoints::find, newWaypoints, newGrid and the Grid geometry UNDER THE SAME C++ test entry layer (oracle/ref_shim) that
oracle/_ref/libmazetd_ref.so puts over the reference's own sources/map/paths.cpp + grid.cpp. Both libraries are driven
through the unchanged sources/grid.h interface by identical call sequences; every array must come out bit-identical.
Both libraries are prebuilt where /root/reference exists (oracle/Makefile) and travel to the GPU box."""
import os

import numpy as np
import pytest

import oracle_lib
from mazetd_b200 import synth
from oracle_lib import KAT_SEED

pytestmark = pytest.mark.gpu

def vp(a):
    import ctypes

    return ctypes.c_void_p(a.ctypes.data)


KEYS = ("wp_tiles", "tiles", "dist", "fullpath", "fullpath_len", "full_distance", "agg")


@pytest.fixture(scope="module")
def both():
    if os.path.isdir("/root/reference/sources/map"):
        oracle_lib.build_ref()
        oracle_lib.build_adapter()
    if not (oracle_lib.Ref.available() and os.path.exists(oracle_lib.ADAPTER_SO)):
        pytest.skip("oracle/_ref libraries not prebuilt (they need /root/reference at build time)")
    return oracle_lib.Ref(), oracle_lib.Ref(oracle_lib.ADAPTER_SO)


@pytest.mark.parametrize("k", [0, 1, 7, 23])
def test_new_procedural_grid_waypoints_sequence(both, k):
    """mapGenTaskEntry's sequence newProcedural -> newGrid -> newWaypoints (sources/map/generate.cpp:155-173) from one
    RNG seed: the adapter consumes the thread-local RNG stream exactly like the reference, so even the spawn tiles agree"""
    ref, ada = both
    r = ref.mapgen(KAT_SEED[0] + k, KAT_SEED[1])
    a = ada.mapgen(KAT_SEED[0] + k, KAT_SEED[1])
    for key in ("elev", "flags") + KEYS:
        assert np.array_equal(a[key], r[key]), f"map {k}: {key} differs between adapter and reference"


def test_waypoints_update_after_tower_placements(both):
    """placeStructure's flag flip + globalWaypoints->update() (sources/player/buildingsPlacement.cpp:120-130)"""
    ref, ada = both
    r = ref.mapgen(KAT_SEED[0] + 3, KAT_SEED[1])
    W = H = 101
    flags = r["flags"].copy()
    rng = np.random.default_rng(1)
    path = r["fullpath"][0][: r["fullpath_len"][0]]
    for step in range(6):
        cand = int(path[rng.integers(3, len(path) - 3)])
        if flags[cand] & (synth.FLAG_WAYPOINT | synth.FLAG_BUILDING):
            continue
        flags[cand] |= synth.FLAG_BUILDING
        wr = ref.waypoints_update(W, H, r["elev"], flags, r["wp_tiles"], 50, 50)
        wa = ada.waypoints_update(W, H, r["elev"], flags, r["wp_tiles"], 50, 50)
        for key in KEYS:
            assert np.array_equal(wa[key], wr[key]), f"placement {step}: {key} differs"


def test_directions_update_single_layer_and_unreachable(both):
    """Directions::update as placingBuildingBlocksMonsters uses it (sources/player/buildingsPlacement.cpp:18-29)"""
    ref, ada = both
    W, H = 61, 47
    elev, flags = synth.random_grid(W, H, seed_k=2, wall_pct=40)
    for tile in synth.pick_open_tiles(flags, 3, 5):
        tr, dr = ref.directions_update(W, H, elev, flags, int(tile))
        ta, da = ada.directions_update(W, H, elev, flags, int(tile))
        assert np.array_equal(da, dr) and np.array_equal(ta, tr)
        assert (dr == oracle_lib.M).any()  # 40 % walls leave unreachable pockets


def test_find_and_steering_through_the_adapter(both):
    """Waypoints::find for every visited-bits combination, the monster tick and position() on top of adapter-built layers"""
    ref, ada = both
    r = ref.mapgen(KAT_SEED[0] + 2, KAT_SEED[1])
    W = H = 101
    hr = ref.map_create(W, H, r["elev"], r["flags"], r["wp_tiles"], 50, 50)
    ha = ada.map_create(W, H, r["elev"], r["flags"], r["wp_tiles"], 50, 50)
    ref.L.ref_waypoints_update(hr)
    ada.L.ref_waypoints_update(ha)
    rng = np.random.default_rng(3)
    open_tiles = np.flatnonzero((r["flags"] & synth.FLAG_INVALID) == 0)
    for pos in rng.choice(open_tiles, 200):
        for bits in range(16):
            assert ada.find(ha, int(pos), bits) == ref.find(hr, int(pos), bits)
    # blocked / unreachable positions give { m, m } or wrapped sums: same on both
    for pos in rng.choice(np.flatnonzero(r["flags"] & synth.FLAG_INVALID), 50):
        for bits in (0, 5, 14):
            assert ada.find(ha, int(pos), bits) == ref.find(hr, int(pos), bits)
    n = 500
    path = r["fullpath"][int(r["agg"][2])][: r["fullpath_len"][int(r["agg"][2])]]
    tile = rng.choice(path, n).astype(np.uint32)
    mk = lambda: dict(tile=tile.copy(), tileStart=tile.copy(), tileEnd=tile.copy(), timeStart=np.zeros(n, np.uint32), timeEnd=np.zeros(n, np.uint32),
                      bits=np.full(n, 1 << int(r["agg"][2]), np.uint32), timeToArrive=rng.integers(2000, 9000, n).astype(np.uint32) * 0 + 6000,
                      dotflags=(np.arange(n) % 4).astype(np.uint8), damage=np.ones(n, np.uint32), alive=np.ones(n, np.uint8))
    A, R = mk(), mk()
    for t in range(0, 600, 7):
        assert ada.monsters_tick(ha, t, A) == ref.monsters_tick(hr, t, R)
        for key in ("tile", "tileStart", "tileEnd", "timeStart", "timeEnd", "bits", "alive"):
            assert np.array_equal(A[key], R[key]), (t, key)
    pa = ada.monsters_position(ha, 601, A["tileStart"], A["tileEnd"], A["timeStart"], A["timeEnd"])
    pr = ref.monsters_position(hr, 601, R["tileStart"], R["tileEnd"], R["timeStart"], R["timeEnd"])
    assert np.array_equal(pa.view(np.uint32), pr.view(np.uint32))  # float32 positions: bit-identical (same host arithmetic)
    ref.L.ref_map_destroy(hr)
    ada.L.ref_map_destroy(ha)


def test_grid_geometry_golden(both):
    """GridTester's assertions (sources/map/grid.cpp:214-251) hold for the adapter's geometry as for the reference's"""
    ref, ada = both
    for lib in (ref, ada):
        L = lib.L
        assert L.ref_grid_index_i(10, 5, 2, 1, -2, -1) == 0
        assert L.ref_grid_index_i(10, 5, 2, 1, 7, 3) == 49
        assert L.ref_grid_index_i(10, 5, 2, 1, 8, 3) == oracle_lib.M
    rng = np.random.default_rng(0)
    W, H = 33, 21
    elev = rng.integers(-3000, 3000, W * H).astype(np.int16)
    a = rng.integers(0, W * H, 500).astype(np.uint32)
    b = rng.integers(0, W * H, 500).astype(np.uint32)
    outs = []
    for lib in (ref, ada):
        o = np.zeros(500, np.uint32)
        lib.L.ref_neighbor_distances(W, H, 7, 3, vp(elev), vp(a), vp(b), 500, vp(o))
        c = np.zeros((500, 3), np.float32)
        lib.L.ref_grid_centers(W, H, 7, 3, vp(elev), vp(a), 500, vp(c))
        outs.append((o, c))
    assert np.array_equal(outs[0][0], outs[1][0])
    assert np.array_equal(outs[0][1].view(np.uint32), outs[1][1].view(np.uint32))
