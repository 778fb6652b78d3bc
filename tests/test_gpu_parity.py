"""GPU parity tests: every call goes through the C ABI (mazetd_b200.capi -> libmazetd_b200.so) and is compared
bit for bit with the CPU oracle (oracle/libmazetd_oracle.so) and, where the prebuilt library travelled, with the
reference's own compiled code (oracle/_ref/libmazetd_ref.so). Run with `-m gpu` on the B200 box."""
import os

import numpy as np
import pytest

from mazetd_b200 import capi, synth
from oracle_lib import KAT_SEED, M

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from mazetd_b200 import capi

    c = capi.Context(0)
    yield c
    c.close()


@pytest.fixture(params=["dataflow", "rounds"])
def tctx(ctx, request):
    """the tiled (large-grid) path under both schedulers: k_tile_async (default) and the round-based k_tile_rounds"""
    ctx.debug_set(ctx.DEBUG_BIG_ASYNC, 1 if request.param == "dataflow" else 0)
    yield ctx
    ctx.debug_set(ctx.DEBUG_BIG_ASYNC, 1)


@pytest.fixture(params=["fused", "two_kernels"])
def gctx(ctx, request):
    """mtd_mapgen_batch with waypoints under both forms: generator fused into k_waypoints_small (default: the next map's
    elevations are computed by the warps without a layer while the current map is relaxed) and k_mapgen_grid first"""
    ctx.debug_set(ctx.DEBUG_FUSED_MAPGEN, 1 if request.param == "fused" else 0)
    yield ctx
    ctx.debug_set(ctx.DEBUG_FUSED_MAPGEN, 1)


GOLDEN_BIG = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "big_configs.npz")


def _checks(a):
    return int(np.bitwise_xor.reduce(a)), int(a.astype(np.uint64).sum() & np.uint64(0xFFFFFFFFFFFFFFFF))


def diff_report(a, b, W):
    bad = np.flatnonzero(a != b)
    i = int(bad[0])
    return f"{len(bad)} mismatches; first at tile {i} (x={i % W}, y={i // W}): got {a.reshape(-1)[i]} want {b.reshape(-1)[i]}"


# ---------------------------------------------------------------------------------------------------
# G4: edge weights
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("amp", [90, 512, 32768])
def test_edge_weights_bit_exact(ctx, oracle, amp):
    rng = np.random.default_rng(amp)
    W, H = 257, 129
    elev = rng.integers(-amp, amp, W * H).astype(np.int16)
    flags = np.zeros(W * H, np.uint16)
    ctx.grid_upload(W, H, elev, flags)
    r, d = ctx.edge_weights()
    ro, do = oracle.edge_weights(W, H, elev)
    assert np.array_equal(r, ro), diff_report(r, ro, W)
    assert np.array_equal(d, do), diff_report(d, do, W)


def test_neighbor_distance_general_pairs(ctx, oracle):
    """Grid::neighborDistance also for non-adjacent tiles (the reference API allows any pair)"""
    import ctypes as C

    rng = np.random.default_rng(3)
    W, H = 101, 101
    elev = rng.integers(-2200, 1200, W * H).astype(np.int16)
    ctx.grid_upload(W, H, elev, np.zeros(W * H, np.uint16), 50, 50)
    a = rng.integers(0, W * H, 4000).astype(np.uint32)
    b = rng.integers(0, W * H, 4000).astype(np.uint32)
    got = ctx.neighbor_distances(a, b)
    g = oracle.grid(W, H, 50, 50, elev, None)
    want = np.array([oracle.L.mo_neighbor_distance(C.byref(g), int(x), int(y)) for x, y in zip(a, b)], np.uint32)
    assert np.array_equal(got, want)


# ---------------------------------------------------------------------------------------------------
# P1: Directions::update -- small (one CTA) path
# ---------------------------------------------------------------------------------------------------
SMALL_CASES = ["open", "walls30", "walls45", "maze", "tiny", "row", "col", "blocked_source", "flat", "default101"]


def make_case(case):
    if case == "tiny":
        W, H = 1, 1
    elif case == "row":
        W, H = 37, 1
    elif case == "col":
        W, H = 1, 29
    elif case == "default101":
        W, H = 101, 101
    else:
        W, H = 67, 45
    if case == "maze":
        elev, flags = synth.maze_grid(W, H, 3)
    elif case == "flat":
        elev = np.zeros(W * H, np.int16)
        flags = np.full(W * H, synth.FLAG_SUN, np.uint16)
    else:
        pct = {"open": 0, "walls30": 30, "walls45": 45, "default101": 35}.get(case, 20)
        elev, flags = synth.random_grid(W, H, seed_k=11, wall_pct=pct, elev_mod=512)
    if case == "blocked_source":
        src = int(np.flatnonzero(flags & synth.FLAG_INVALID)[5])
    else:
        src = int(synth.pick_open_tiles(flags, 1, 2)[0]) if W * H > 1 else 0
    return W, H, elev, flags, src


@pytest.mark.parametrize("case", SMALL_CASES)
def test_directions_small_bit_exact(ctx, oracle, case):
    W, H, elev, flags, src = make_case(case)
    ctx.grid_upload(W, H, elev, flags)
    tiles, dist = ctx.directions_update(src)
    to, do = oracle.directions_update(W, H, elev, flags, src)
    assert np.array_equal(dist, do), diff_report(dist, do, W)
    assert np.array_equal(tiles, to), diff_report(tiles, to, W)


def test_directions_small_vs_reference_code(ctx, ref):
    """same comparison against the reference's own compiled paths.cpp (oracle/_ref)"""
    for case in ("walls30", "maze", "flat", "blocked_source"):
        W, H, elev, flags, src = make_case(case)
        ctx.grid_upload(W, H, elev, flags)
        tiles, dist = ctx.directions_update(src)
        tr, dr = ref.directions_update(W, H, elev, flags, src)
        assert np.array_equal(dist, dr) and np.array_equal(tiles, tr), case


def test_directions_extra_blocked_is_the_trial_grid(ctx, oracle):
    """placingBuildingBlocksMonsters: a copy of the grid with Building on the cursor tile (buildingsPlacement.cpp:20-29)"""
    W, H, elev, flags, src = make_case("walls30")
    ctx.grid_upload(W, H, elev, flags)
    open_tiles = np.flatnonzero((flags & (synth.FLAG_INVALID | synth.FLAG_BUILDING)) == 0)
    for cand in open_tiles[:: max(1, len(open_tiles) // 12)]:
        if cand == src:
            continue
        tiles, dist = ctx.directions_update(src, extra_blocked=int(cand))
        f2 = flags.copy()
        f2[cand] |= synth.FLAG_BUILDING
        to, do = oracle.directions_update(W, H, elev, f2, src)
        assert np.array_equal(dist, do) and np.array_equal(tiles, to)


def test_directions_multi_source_extension(ctx, oracle):
    W, H, elev, flags, _ = make_case("walls30")
    ctx.grid_upload(W, H, elev, flags)
    srcs = synth.pick_open_tiles(flags, 7, 4)
    tiles, dist = ctx.directions_update(srcs)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(dist, do) and np.array_equal(tiles, to)


def test_every_single_building_toggle_on_one_map(ctx, oracle):
    """flip Building on every passable tile of one generated map, rebuild, compare"""
    o = oracle.mapgen(KAT_SEED[0], KAT_SEED[1])
    W = H = 101
    flags = o["flags"].copy()
    ctx.grid_upload(W, H, o["elev"], flags, 50, 50)
    src = int(o["wp_tiles"][0])
    passable = np.flatnonzero((flags & (synth.FLAG_INVALID | synth.FLAG_WAYPOINT)) == 0)
    for t in passable[::37]:
        f2 = flags.copy()
        f2[t] |= synth.FLAG_BUILDING
        ctx.grid_set_flags(int(t), int(f2[t]))
        tiles, dist = ctx.directions_update(src)
        to, do = oracle.directions_update(W, H, o["elev"], f2, src, 50, 50)
        assert np.array_equal(dist, do) and np.array_equal(tiles, to), int(t)
        ctx.grid_set_flags(int(t), int(flags[t]))


# ---------------------------------------------------------------------------------------------------
# P1: tiled path for large grids
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape,pct", [((300, 260), 30), ((513, 257), 0), ((640, 384), 40), ((1024, 1024), 30)])
def test_directions_tiled_bit_exact(tctx, oracle, shape, pct):
    ctx = tctx
    W, H = shape
    elev, flags = synth.random_grid(W, H, seed_k=21, wall_pct=pct, elev_mod=512)
    src = int(synth.pick_open_tiles(flags, 1, 5)[0])
    ctx.grid_upload(W, H, elev, flags)
    tiles, dist = ctx.directions_update(src)
    to, do = oracle.directions_update(W, H, elev, flags, src)
    assert np.array_equal(dist, do), diff_report(dist, do, W)
    assert np.array_equal(tiles, to), diff_report(tiles, to, W)


def test_directions_tiled_maze_and_multisource(tctx, oracle):
    ctx = tctx
    W, H = 290, 270
    elev, flags = synth.maze_grid(W, H, 7)
    ctx.grid_upload(W, H, elev, flags)
    tiles, dist = ctx.directions_update(0)
    to, do = oracle.directions_update(W, H, elev, flags, 0)
    assert np.array_equal(dist, do), diff_report(dist, do, W)
    assert np.array_equal(tiles, to)
    elev, flags = synth.random_grid(W, H, seed_k=2, wall_pct=30, elev_mod=512)
    ctx.grid_upload(W, H, elev, flags)
    srcs = synth.pick_open_tiles(flags, 16, 6)
    tiles, dist = ctx.directions_update(srcs)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(dist, do), diff_report(dist, do, W)
    assert np.array_equal(tiles, to)


def test_directions_tiled_blocked_source_and_border_sources(tctx, oracle):
    ctx = tctx
    W, H = 300, 300
    elev, flags = synth.random_grid(W, H, seed_k=9, wall_pct=25, elev_mod=300)
    ctx.grid_upload(W, H, elev, flags)
    blocked = np.flatnonzero(flags & synth.FLAG_INVALID)
    # sources straddling tile borders (x or y = 63, 64, 127, 128 ...) and a blocked source
    cands = [63 + 64 * W, 64 + 64 * W, 63 + 63 * W, 128 + 127 * W, int(blocked[100]), 0, W * H - 1, W - 1]
    for s in cands:
        tiles, dist = ctx.directions_update(s)
        to, do = oracle.directions_update(W, H, elev, flags, s)
        assert np.array_equal(dist, do), (s, diff_report(dist, do, W))
        assert np.array_equal(tiles, to), s


def test_dataflow_scheduler_really_runs_and_windows_agree(ctx, oracle):
    """k_tile_async is the path taken by default (statistics say so), and its priority window changes nothing but time"""
    W, H = 700, 520
    elev, flags = synth.random_grid(W, H, seed_k=5, wall_pct=30, elev_mod=512)
    srcs = synth.pick_open_tiles(flags, 3, 9)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    ctx.grid_upload(W, H, elev, flags)
    for delta in (1, 4096, 32768, 0xFFFFFFFF):
        ctx.debug_set(ctx.DEBUG_ASYNC_DELTA, delta)
        tiles, dist = ctx.directions_update(srcs)
        st = ctx.big_stats()
        assert st["dataflow"] and st["rounds"] == 0 and st["visits"] >= ((W + 31) // 32) * ((H + 31) // 32) // 2
        assert np.array_equal(dist, do), (delta, diff_report(dist, do, W))
        assert np.array_equal(tiles, to), delta
    ctx.debug_set(ctx.DEBUG_ASYNC_DELTA, 65536)
    # the scheduler's optional forms: no late look at the apron / two of them, revisits continued on the last visitor's frame,
    # fewer workers, one mailbox per worker -- all performance knobs, the fixpoint is the same
    for key, value, default in ((ctx.DEBUG_ASYNC_LATE_LOOKS, 0, 1), (ctx.DEBUG_ASYNC_LATE_LOOKS, 2, 1), (ctx.DEBUG_ASYNC_CONT, 1, 0), (ctx.DEBUG_ASYNC_WORKERS, 5, 16), (ctx.DEBUG_ASYNC_DEPTH, 1, 2)):
        try:
            ctx.debug_set(key, value)
            tiles, dist = ctx.directions_update(srcs)
        finally:
            ctx.debug_set(key, default)
        assert ctx.big_stats()["dataflow"]
        assert np.array_equal(dist, do), (key, value, diff_report(dist, do, W))
        assert np.array_equal(tiles, to), (key, value)


def test_c5_full_size_against_reference_golden_and_oracle(tctx, oracle):
    """BASELINE config 5 at its own size (4096^2, 8 layers): every layer against the committed golden made from the
    reference's compiled Directions::update (tests/golden/make_golden_big.py), and, array for array, against a CPU
    Dijkstra run here (the compiled reference when it travelled, else the oracle restatement)"""
    import oracle_lib
    import torch

    ctx = tctx
    g = np.load(GOLDEN_BIG)
    W = H = 4096
    elev, flags = synth.random_grid(W, H, seed_k=1, wall_pct=30, elev_mod=512)
    srcs = synth.pick_open_tiles(flags, 8, 1)
    assert np.array_equal(srcs, g["C5_sources"])
    ctx.grid_upload(W, H, elev, flags)
    dt = torch.empty((8, W * H), dtype=torch.int32, device="cuda")
    dd = torch.empty((8, W * H), dtype=torch.int32, device="cuda")
    ctx.layers_update_dev(srcs, dt.data_ptr(), dd.data_ptr())
    ctx.synchronize()
    cpu = oracle_lib.Ref() if oracle_lib.Ref.available() else None
    rows = [int(r) for r in g["C5_rows"]]
    for l in range(8):
        dist = dd[l].cpu().numpy().view(np.uint32)
        tiles = dt[l].cpu().numpy().view(np.uint32)
        assert _checks(dist) == (int(g["C5_dist_xor"][l]), int(g["C5_dist_sum"][l])), l
        assert _checks(tiles) == (int(g["C5_tiles_xor"][l]), int(g["C5_tiles_sum"][l])), l
        assert int((dist == M).sum()) == int(g["C5_unreachable"][l])
        for j, r in enumerate(rows):
            assert np.array_equal(dist[r * W:(r + 1) * W], g["C5_dist_rows"][l][j]) and np.array_equal(tiles[r * W:(r + 1) * W], g["C5_tiles_rows"][l][j])
        if ctx.big_stats()["dataflow"] or l in (0, 6):  # all 8 layers array for array under the default scheduler
            to, do = cpu.directions_update(W, H, elev, flags, int(srcs[l])) if cpu else oracle.directions_update(W, H, elev, flags, int(srcs[l]))
            assert np.array_equal(dist, do), diff_report(dist, do, W)
            assert np.array_equal(tiles, to), diff_report(tiles, to, W)


def test_c4_full_size_against_reference_golden(ctx):
    """BASELINE config 4 at its own size (16384^2, 16 sources, one layer) against the committed golden: distances = minimum
    of 16 runs of the reference's compiled Directions::update, next hops = closed form of the heap order on them"""
    import torch

    g = np.load(GOLDEN_BIG)
    if "C4_dist_sum" not in g:
        pytest.skip("C4 golden not generated")
    W = H = 16384
    elev, flags = synth.random_grid(W, H, seed_k=1, wall_pct=30, elev_mod=512)
    srcs = synth.pick_open_tiles(flags, 16, 1)
    assert np.array_equal(srcs, g["C4_sources"])
    ctx.grid_upload(W, H, elev, flags)
    del elev, flags
    dt = torch.empty(W * H, dtype=torch.int32, device="cuda")
    dd = torch.empty(W * H, dtype=torch.int32, device="cuda")
    ctx.directions_update_dev(srcs, dt.data_ptr(), dd.data_ptr())
    ctx.synchronize()
    assert ctx.big_stats()["dataflow"]
    for name, t in (("dist", dd), ("tiles", dt)):
        v = t.to(torch.int64) & 0xFFFFFFFF
        s = 0
        for part in torch.chunk(v, 16):  # 64-bit wrap-around sum without overflow surprises
            s = (s + int(part.sum().item())) & 0xFFFFFFFFFFFFFFFF
        x = t.view(-1)
        xr = x[0:1].clone()
        h = x
        while h.numel() > 1:  # xor-reduce by halving
            n = h.numel() // 2
            rest = h[2 * n:]
            h = torch.bitwise_xor(h[:n], h[n:2 * n])
            if rest.numel():
                h = torch.cat([h, rest])
        del xr
        assert (int(h.item()) & 0xFFFFFFFF) == int(g[f"C4_{name}_xor"]), name
        assert s == int(g[f"C4_{name}_sum"]), name
        for j, r in enumerate(int(r) for r in g["C4_rows"]):
            assert np.array_equal(t[r * W:(r + 1) * W].cpu().numpy().view(np.uint32), g[f"C4_{name}_rows"][j]), (name, r)
    assert int((dd == -1).sum().item()) == int(g["C4_unreachable"])
    ctx.grid_upload(64, 64, np.zeros(4096, np.int16), np.zeros(4096, np.uint16))  # release the 2 GiB of derived data


def test_tiled_properties_at_4096(ctx):
    """full-size property checks where the oracle would take too long: the result is a fixpoint of relaxation
    (every edge satisfied, every finite non-source distance attained by its next hop) and next hops are neighbours"""
    W = H = 4096
    elev, flags = synth.random_grid(W, H, seed_k=1, wall_pct=30, elev_mod=512)
    srcs = synth.pick_open_tiles(flags, 8, 1)
    ctx.grid_upload(W, H, elev, flags)
    tiles, dist = ctx.directions_update(srcs)
    r, d = ctx.edge_weights()
    D = dist.reshape(H, W).astype(np.int64)
    D[D == M] = 1 << 40
    blocked = (flags.reshape(H, W) & synth.FLAG_INVALID) != 0
    R = r.reshape(H, W).astype(np.int64)
    Dn = d.reshape(H, W).astype(np.int64)
    # no edge into an unblocked tile can improve it
    assert np.all((D[:, :-1] + R[:, :-1] >= D[:, 1:]) | blocked[:, 1:])
    assert np.all((D[:, 1:] + R[:, :-1] >= D[:, :-1]) | blocked[:, :-1])
    assert np.all((D[:-1, :] + Dn[:-1, :] >= D[1:, :]) | blocked[1:, :])
    assert np.all((D[1:, :] + Dn[:-1, :] >= D[:-1, :]) | blocked[:-1, :])
    # every reached non-source tile has a next hop that is a neighbour and attains its distance exactly
    t = tiles.astype(np.int64)
    reached = (dist != M) & (dist != 0)
    idx = np.flatnonzero(reached)
    nh = t[idx]
    assert np.all(nh != M)
    dx = np.abs(nh % W - idx % W) + np.abs(nh // W - idx // W)
    assert np.all(dx == 1)
    lo = np.minimum(nh, idx)
    w = np.where(np.abs(nh - idx) == 1, r[lo], d[lo]).astype(np.int64)
    assert np.all(dist[nh].astype(np.int64) + w == dist[idx].astype(np.int64))
    assert np.all(tiles[dist == M] == M) and np.all(tiles[srcs] == M) and np.all(dist[srcs] == 0)
    assert reached.sum() > 0.5 * (~blocked).sum()


# ---------------------------------------------------------------------------------------------------
# P2-P4: Waypoints::update / find
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [1, 2, 3, 4])
def test_waypoints_update_bit_exact(ctx, oracle, n):
    W, H = 41, 37
    elev, flags = synth.random_grid(W, H, seed_k=5, wall_pct=25, elev_mod=200)
    wp = synth.pick_open_tiles(flags, n, 9)
    ctx.grid_upload(W, H, elev, flags)
    g = ctx.waypoints_update(wp)
    o = oracle.waypoints_update(W, H, elev, flags, wp)
    for k in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
        assert np.array_equal(g[k], o[k]), k


def test_waypoints_update_more_than_four_layers(ctx, oracle):
    W, H = 41, 37
    elev, flags = synth.random_grid(W, H, seed_k=5, wall_pct=20, elev_mod=200)
    wp = synth.pick_open_tiles(flags, 6, 9)
    ctx.grid_upload(W, H, elev, flags)
    g = ctx.waypoints_update(wp)
    o = oracle.waypoints_update(W, H, elev, flags, wp)
    for k in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
        assert np.array_equal(g[k], o[k]), k


def test_dense_fallback_when_the_frontier_queue_overflows(ctx, oracle):
    """a tiny queue segment forces the queue-less dense rounds of the one-CTA kernel; results must not change"""
    W, H = 101, 101
    elev, flags = synth.random_grid(W, H, seed_k=11, wall_pct=10, elev_mod=300)
    wp = synth.pick_open_tiles(flags, 4, 2)
    ctx.grid_upload(W, H, elev, flags, 50, 50)
    o = oracle.waypoints_update(W, H, elev, flags, wp, 50, 50)
    try:
        for cap in (32, 40, 64):
            ctx.debug_set(ctx.DEBUG_QSEG_CAP, cap)
            g = ctx.waypoints_update(wp)
            for k in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
                assert np.array_equal(g[k], o[k]), (cap, k)
            t, d = ctx.directions_update(wp[:1])
            assert np.array_equal(d, o["dist"][0]) and np.array_equal(t, o["tiles"][0])
    finally:
        ctx.debug_set(ctx.DEBUG_QSEG_CAP, 0)


@pytest.mark.parametrize("shape,n", [((150, 140), 4), ((170, 160), 3), ((200, 200), 3), ((255, 255), 2), ((33, 7), 6)])
def test_waypoints_update_fewer_resident_layers_than_waypoints(ctx, oracle, shape, n):
    """maps whose layers do not all fit shared memory at once go through several passes and walk on the L2 copies"""
    W, H = shape
    elev, flags = synth.random_grid(W, H, seed_k=21, wall_pct=22, elev_mod=400)
    wp = synth.pick_open_tiles(flags, n, 5)
    ctx.grid_upload(W, H, elev, flags)
    g = ctx.waypoints_update(wp, cap=8192)
    o = oracle.waypoints_update(W, H, elev, flags, wp, cap=8192)
    for k in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
        assert np.array_equal(g[k], o[k]), k


def test_open_map_wide_frontier(ctx, oracle):
    """no walls: the widest frontiers the 101x101 map can produce"""
    W, H = 101, 101
    rng = np.random.default_rng(5)
    elev = rng.integers(-40, 40, W * H).astype(np.int16)
    flags = np.zeros(W * H, np.uint16)
    wp = np.array([0, W * H - 1, 50 * W + 50, 100], np.uint32)
    ctx.grid_upload(W, H, elev, flags, 50, 50)
    g = ctx.waypoints_update(wp)
    o = oracle.waypoints_update(W, H, elev, flags, wp, 50, 50)
    for k in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
        assert np.array_equal(g[k], o[k]), k


def test_waypoints_empty_is_a_noop(ctx):
    W, H, elev, flags, _ = make_case("open")
    ctx.grid_upload(W, H, elev, flags)
    g = ctx.waypoints_update(np.zeros(0, np.uint32))  # paths.cpp:113-114
    assert g["agg"].tolist() == [0, 0, 0]


def test_waypoints_on_generated_maps_vs_reference_code(ctx, ref):
    for k in (0, 3):
        r = ref.mapgen(KAT_SEED[0] + k, KAT_SEED[1])
        ctx.grid_upload(101, 101, r["elev"], r["flags"], 50, 50)
        g = ctx.waypoints_update(r["wp_tiles"])
        for key in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
            assert np.array_equal(g[key], r[key]), (k, key)


def test_find_batch_bit_exact_including_unreachable(ctx, oracle):
    W, H = 29, 23
    elev, flags = synth.random_grid(W, H, seed_k=8, wall_pct=38, elev_mod=300)
    wp = synth.pick_open_tiles(flags, 4, 1)
    ctx.grid_upload(W, H, elev, flags)
    ctx.waypoints_update(wp)
    o = oracle.waypoints_update(W, H, elev, flags, wp)
    wv = oracle.waypoints_view(W, H, elev, flags, wp, o["tiles"], o["dist"])
    rng = np.random.default_rng(3)
    pos = rng.integers(0, W * H, 3000).astype(np.uint32)
    bits = rng.integers(0, 16, 3000).astype(np.uint32)
    t, d = ctx.waypoints_find(pos, bits)
    want = np.array([oracle.find(wv, int(p), int(b)) for p, b in zip(pos, bits)], np.uint32)
    assert np.array_equal(t, want[:, 0]) and np.array_equal(d, want[:, 1])


def test_placing_blocks_matches_oracle(ctx, oracle):
    o = oracle.mapgen(KAT_SEED[0] + 1, KAT_SEED[1])
    W = H = 101
    ctx.grid_upload(W, H, o["elev"], o["flags"], 50, 50)
    ctx.waypoints_update(o["wp_tiles"])
    passable = np.flatnonzero((o["flags"] & (synth.FLAG_INVALID | synth.FLAG_WAYPOINT)) == 0).astype(np.uint32)
    rng = np.random.default_rng(2)
    mt = rng.choice(passable, 500).astype(np.uint32)
    alive = (rng.integers(0, 5, 500) > 0).astype(np.uint8)
    z = np.zeros(500, np.uint32)
    ctx.monsters_upload(dict(tile=mt, tileStart=mt, tileEnd=mt, timeStart=z, timeEnd=z, bits=z, timeToArrive=z, alive=alive))
    got_any = set()
    for cand in np.concatenate([o["fullpath"][0][: int(o["fullpath_len"][0])][1:60:3], rng.choice(passable, 25)]):
        want = oracle.placing_blocks(W, H, o["elev"], o["flags"], o["wp_tiles"], int(cand), mt, alive, 50, 50)
        got = ctx.placing_blocks(int(cand))
        assert got == bool(want), int(cand)
        got_any.add(got)
    assert got_any  # ran


# ---------------------------------------------------------------------------------------------------
# M1-M7: procedural generator
# ---------------------------------------------------------------------------------------------------
def test_elevation_samples_bit_exact(ctx, oracle):
    import ctypes as C

    from oracle_lib import MoProcedural

    rng = np.random.default_rng(4)
    xy = np.concatenate([rng.uniform(-60, 60, (6000, 2)), np.array([[0, 0], [-50, -50], [50, 50], [30, 0], [0, -30.0001]])]).astype(np.float32)
    for k in (0, 5):
        got = ctx.elevation_samples(KAT_SEED[0] + k, KAT_SEED[1], xy)
        g = oracle.rng(KAT_SEED[0] + k, KAT_SEED[1])
        p = MoProcedural()
        oracle.L.mo_procedural_init(C.byref(p), C.byref(g))
        want = np.array([oracle.L.mo_elevation(C.byref(p), float(x), float(y)) for x, y in xy], np.float32)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), f"{(got != want).sum()} of {len(xy)} samples differ"


def test_mapgen_batch_bit_exact(gctx, oracle):
    ctx = gctx
    n = 24
    g = ctx.mapgen_batch(KAT_SEED[0], KAT_SEED[1], 0, n, cap=2048)
    for k in range(n):
        o = oracle.mapgen(KAT_SEED[0] + k, KAT_SEED[1], cap=2048)
        for key in ("elev", "flags", "wp_tiles", "dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
            assert np.array_equal(g[key][k], o[key]), (k, key)


@pytest.mark.parametrize("gw", [1, 2, 4])
@pytest.mark.parametrize("walker", [1, 2])
def test_mapgen_batch_walker_warp_and_group_widths(gctx, oracle, gw, walker):
    """the batched kernel's variants -- 1/2/4 warps per layer, paths walked by the workers (1) or by the walker warps on
    finished maps' layers in global memory (2), generator fused or not -- must all produce the reference's bits"""
    ctx = gctx
    n = 10
    try:
        ctx.debug_set(ctx.DEBUG_GROUP_WARPS, gw)
        ctx.debug_set(ctx.DEBUG_WALKER, walker)
        g = ctx.mapgen_batch(KAT_SEED[0], KAT_SEED[1], 100, n, cap=2048)
    finally:
        ctx.debug_set(ctx.DEBUG_GROUP_WARPS, 2)
        ctx.debug_set(ctx.DEBUG_WALKER, 0)
    for k in range(n):
        o = oracle.mapgen(KAT_SEED[0] + 100 + k, KAT_SEED[1], cap=2048)
        for key in ("elev", "flags", "wp_tiles", "dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
            assert np.array_equal(g[key][k], o[key]), (k, key)


def test_mapgen_batch_many_maps_per_cta(gctx, oracle):
    """more maps than CTAs: every CTA loops over several maps and its walker warps run behind the workers"""
    ctx = gctx
    n = 700
    g = ctx.mapgen_batch(KAT_SEED[0], KAT_SEED[1], 1000, n, cap=1024)
    for k in list(range(0, n, 37)) + [n - 1]:
        o = oracle.mapgen(KAT_SEED[0] + 1000 + k, KAT_SEED[1], cap=1024)
        for key in ("elev", "flags", "wp_tiles", "dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
            assert np.array_equal(g[key][k], o[key]), (k, key)


@pytest.mark.parametrize("walker", [1, 2])
def test_walks_with_unreachable_waypoints_take_the_per_hop_path(ctx, oracle, walker):
    """a wall splits the map: some inter-waypoint distances are m, find()'s uint32 sums wrap, and the walk must follow the
    reference's per-hop evaluation instead of the segment shortcut"""
    W, H = 40, 30
    rng = np.random.default_rng(3)
    elev = rng.integers(-100, 100, W * H).astype(np.int16)
    flags = np.zeros(W * H, np.uint16)
    flags[np.arange(H) * W + 20] = synth.FLAG_BUILDING  # full-height wall at x = 20
    for wp in ([5 * W + 3, 9 * W + 30, 20 * W + 10, 25 * W + 35], [5 * W + 3, 20 * W + 10, 9 * W + 30], [2 * W + 2, 3 * W + 20, 4 * W + 8]):
        wp = np.array(wp, np.uint32)
        ctx.grid_upload(W, H, elev, flags)
        try:
            ctx.debug_set(ctx.DEBUG_WALKER, walker)
            g = ctx.waypoints_update(wp)
        finally:
            ctx.debug_set(ctx.DEBUG_WALKER, 0)
        o = oracle.waypoints_update(W, H, elev, flags, wp)
        for k in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
            assert np.array_equal(g[k], o[k]), (list(wp), k)


def test_mapgen_batch_vs_reference_code(gctx, ref):
    ctx = gctx
    g = ctx.mapgen_batch(KAT_SEED[0], KAT_SEED[1], 40, 3, cap=2048)
    for j in range(3):
        r = ref.mapgen(KAT_SEED[0] + 40 + j, KAT_SEED[1], cap=2048)
        for key in ("elev", "flags", "wp_tiles", "dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
            assert np.array_equal(g[key][j], r[key]), (j, key)


def test_mapgen_against_committed_golden(ctx):
    """tests/golden/mapgen_k0_k7.npz was generated from the reference's compiled code (tests/golden/make_golden.py)"""
    path = os.path.join(os.path.dirname(__file__), "golden", "mapgen_k0_k7.npz")
    gold = np.load(path)
    g = ctx.mapgen_batch(KAT_SEED[0], KAT_SEED[1], 0, 8, cap=1024)
    assert np.array_equal(g["elev"], gold["elev"])
    assert np.array_equal(g["flags"], gold["flags"])
    assert np.array_equal(g["wp_tiles"], gold["wp_tiles"])
    assert np.array_equal(g["agg"], gold["agg"])
    assert np.array_equal(g["full_distance"], gold["full_distance"])
    assert np.array_equal(g["fullpath_len"], gold["fullpath_len"])
    for k in range(8):
        for l in range(4):
            assert int(np.bitwise_xor.reduce(g["dist"][k, l])) == int(gold["dist_xor"][k, l])
            assert int(g["dist"][k, l].astype(np.uint64).sum() & 0xFFFFFFFF) == int(gold["dist_sum"][k, l])
            assert int(g["tiles"][k, l].astype(np.uint64).sum() & 0xFFFFFFFF) == int(gold["tiles_sum"][k, l])


# ---------------------------------------------------------------------------------------------------
# S1-S3: steering
# ---------------------------------------------------------------------------------------------------
def test_steering_bit_exact_over_many_ticks(ctx, oracle):
    o = oracle.mapgen(KAT_SEED[0] + 2, KAT_SEED[1])
    W = H = 101
    ctx.grid_upload(W, H, o["elev"], o["flags"], 50, 50)
    g = ctx.waypoints_update(o["wp_tiles"])
    assert np.array_equal(g["agg"], o["agg"])
    wv = oracle.waypoints_view(W, H, o["elev"], o["flags"], o["wp_tiles"], o["tiles"], o["dist"], 50, 50)
    spawner = int(o["agg"][2])
    n = 2000
    rng = np.random.default_rng(5)
    speeds = rng.uniform(0.3, 30.0, n).astype(np.float32)  # fast ones finish all waypoints well inside 3000 ticks
    a = {k: np.zeros(n, np.uint32) for k in ("tile", "tileStart", "tileEnd", "timeStart", "timeEnd", "bits", "timeToArrive", "damage")}
    a["dotflags"] = (np.arange(n) % 4).astype(np.uint8)
    a["alive"] = np.ones(n, np.uint8)
    a["damage"][:] = np.arange(n) % 7 + 1
    for i in range(n):
        so = oracle.monster_spawn(wv, spawner, i % 50, float(speeds[i]))
        if i < 40:
            assert np.array_equal(ctx.monster_spawn(i % 50, float(speeds[i])), so)
        a["tile"][i] = a["tileStart"][i] = a["tileEnd"][i] = so[0]
        a["bits"][i] = so[1]
        a["timeToArrive"][i] = so[2]
        a["timeStart"][i] = a["timeEnd"][i] = i % 50
    ctx.monsters_upload(a)
    dmg_g = dmg_o = 0
    for t in range(0, 3000):
        dg, alive_g = ctx.monsters_tick(t)
        dmg_g += dg
        dmg_o += oracle.monsters_tick(wv, t, a)
        if t % 250 == 0 or t == 2999:
            got = ctx.monsters_download()
            for k in got:
                assert np.array_equal(got[k], a[k]), (t, k)
            assert alive_g == int(a["alive"].sum())
            pg = ctx.monsters_positions(t)
            po = oracle.monsters_position(W, H, 50, 50, o["elev"], t, a["tileStart"], a["tileEnd"], a["timeStart"], a["timeEnd"])
            # north_star tolerance for float positions: abs 1e-4; the kernels are unfused so we hold them exact
            assert np.array_equal(pg.view(np.uint32), po.view(np.uint32))
    assert dmg_g == dmg_o and dmg_o > 0


def test_directions_against_committed_golden(ctx):
    """tests/golden/directions.npz: Directions::update of the reference's compiled paths.cpp (make_golden.py)"""
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "directions.npz"))
    cases = {
        "walls": (67, 45) + synth.random_grid(67, 45, seed_k=11, wall_pct=30, elev_mod=512),
        "maze": (67, 45) + synth.maze_grid(67, 45, 3),
    }
    for name, (W, H, elev, flags) in cases.items():
        ctx.grid_upload(W, H, elev, flags)
        t, d = ctx.directions_update(int(gold[name + "_src"]))
        assert np.array_equal(d, gold[name + "_dist"]) and np.array_equal(t, gold[name + "_tiles"]), name
    W, H = 300, 260
    elev, flags = synth.random_grid(W, H, seed_k=21, wall_pct=30, elev_mod=512)
    ctx.grid_upload(W, H, elev, flags)
    t, d = ctx.directions_update(int(gold["big_src"]))
    assert int(np.bitwise_xor.reduce(d)) == int(gold["big_dist_xor"])
    assert int(d.astype(np.uint64).sum() & 0xFFFFFFFF) == int(gold["big_dist_sum"])
    assert int(t.astype(np.uint64).sum() & 0xFFFFFFFF) == int(gold["big_tiles_sum"])
    assert np.array_equal(d[100 * W : 101 * W], gold["big_dist_row100"])
    assert np.array_equal(t[100 * W : 101 * W], gold["big_tiles_row100"])


# ---------------------------------------------------------------------------------------------------
# row stripes of one large grid: several stripe contexts on one GPU run the multi-GPU protocol
# ---------------------------------------------------------------------------------------------------
def _striped_on_one_gpu(W, H, cuts, elev, flags, srcs, round_budget=0):
    from mazetd_b200 import shard, striped

    stripes = []
    for y0, y1 in zip(cuts[:-1], cuts[1:]):
        ya, yb = shard.stripe_halo_rows(y0, y1, H)
        stripes.append(striped.GpuStripe(0, W, H, y0, y1, elev.reshape(H, W)[ya:yb], flags.reshape(H, W)[ya:yb], srcs, round_budget=round_budget))
    results, sweeps = striped.solve_local(stripes)
    tiles = np.concatenate([t.cpu().numpy().view(np.uint32).reshape(-1) for t, _ in results])
    d = np.concatenate([x.cpu().numpy().view(np.uint32).reshape(-1) for _, x in results])
    launches = sum(s.ctx.launches for s in stripes)
    for s in stripes:
        s.close()
    return tiles, d, sweeps, launches


@pytest.mark.parametrize("W,H,cuts,nsrc,pct", [
    (384, 320, [0, 128, 320], 1, 30),        # tile-aligned cut
    (384, 320, [0, 64, 128, 192, 320], 4, 30),
    (200, 301, [0, 97, 203, 301], 3, 35),    # ragged width, cuts inside tiles
    (256, 256, [0, 1, 255, 256], 2, 20),     # one-row stripes at both ends
    (1024, 1024, [0, 256, 512, 768, 1024], 16, 30),
])
def test_stripes_bit_exact_with_whole_grid(oracle, W, H, cuts, nsrc, pct):
    elev, flags = synth.random_grid(W, H, seed_k=W + len(cuts), wall_pct=pct)
    srcs = synth.pick_open_tiles(flags, nsrc, 2)
    tiles, d, sweeps, launches = _striped_on_one_gpu(W, H, cuts, elev, flags, srcs)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do), diff_report(d, do, W)
    assert np.array_equal(tiles, to), diff_report(tiles, to, W)
    assert sweeps >= 1 and launches > 0


_OVERLAP = []  # [True / False] once probed: do kernels of two contexts of this process run side by side on this GPU?


def _overlap_probe():
    """The smallest peer-coupled run there is (two stripes of one tile row each, 64 CTAs per stripe): it can only time out
    at a barrier if the box runs the two kernels one after the other. Probed once; after a successful probe a barrier
    timeout in a real test is a protocol failure, not a skip."""
    from mazetd_b200 import shard, striped

    if _OVERLAP:
        return _OVERLAP[0]
    # CUDA loads kernels lazily, and a load while one stripe's kernel waits at a barrier holds the other stripes back:
    # launch every kernel of the protocol once, alone
    w0, h0 = 64, 64
    e0, f0 = synth.random_grid(w0, h0, seed_k=1, wall_pct=10)
    warm = striped.GpuStripe(0, w0, h0, 0, h0, e0, f0, synth.pick_open_tiles(f0, 1, 1))
    for dataflow in (1, 0):  # both relaxation kernels of the peer-coupled mode
        warm.ctx.debug_set(warm.ctx.DEBUG_BIG_ASYNC, dataflow)
        striped.solve_peer_local([warm])
    warm.close()
    W, H = 64, 128
    e, f = synth.random_grid(W, H, seed_k=2, wall_pct=10)
    srcs = synth.pick_open_tiles(f, 1, 1)
    stripes = []
    for y0, y1 in ((0, 64), (64, 128)):
        ya, yb = shard.stripe_halo_rows(y0, y1, H)
        stripes.append(striped.GpuStripe(0, W, H, y0, y1, e.reshape(H, W)[ya:yb], f.reshape(H, W)[ya:yb], srcs))
    try:
        striped.solve_peer_local(stripes, ctas=64)
        _OVERLAP.append(True)
    except capi.MtdError as err:
        if "did not reach a barrier" not in str(err):
            raise
        _OVERLAP.append(False)
    for st in stripes:
        st.close()
    return _OVERLAP[0]


@pytest.fixture(params=["dataflow", "rounds"])
def peer_kernel(request):
    """the two relaxation kernels of the peer-coupled stripes: k_tile_async (default: border offers and termination waves go
    rank to rank, no rounds) and k_tile_rounds (device-side round barriers)"""
    return request.param


def _peer_striped_on_one_gpu(W, H, cuts, elev, flags, srcs, repeat=1, kernel="dataflow"):
    """stripes as contexts of this process whose kernels run side by side on the one GPU and write into each other's peer
    blocks. Skipped only if the overlap probe says the box serialises the kernels. Returns rounds = None for the dataflow
    kernel (after checking that it really ran)."""
    from mazetd_b200 import shard, striped

    if not _overlap_probe():
        pytest.skip("the stripes' kernels do not run side by side on this GPU (overlap probe)")
    stripes = []
    for y0, y1 in zip(cuts[:-1], cuts[1:]):
        ya, yb = shard.stripe_halo_rows(y0, y1, H)
        stripes.append(striped.GpuStripe(0, W, H, y0, y1, elev.reshape(H, W)[ya:yb], flags.reshape(H, W)[ya:yb], srcs))
        stripes[-1].ctx.debug_set(capi.Context.DEBUG_BIG_ASYNC, 1 if kernel == "dataflow" else 0)
    try:
        for _ in range(repeat):
            results, rounds = striped.solve_peer_local(stripes, ctas=64)
    except capi.MtdError:
        raise  # the probe showed that the kernels overlap: a barrier timeout here is a failure of the protocol, not a skip
    ran_dataflow = [s.ctx.big_stats()["dataflow"] for s in stripes]
    assert all(ran_dataflow) if kernel == "dataflow" else not any(ran_dataflow)
    if kernel == "dataflow":
        rounds = None
    tiles = np.concatenate([t.cpu().numpy().view(np.uint32).reshape(-1) for t, _ in results])
    d = np.concatenate([x.cpu().numpy().view(np.uint32).reshape(-1) for _, x in results])
    for s in stripes:
        s.close()
    return tiles, d, rounds


@pytest.mark.parametrize("W,H,cuts,nsrc,pct", [
    (384, 320, [0, 128, 320], 1, 30),
    (384, 320, [0, 64, 128, 192, 320], 4, 30),
    (200, 301, [0, 97, 203, 301], 3, 35),    # cuts inside tiles
    (256, 256, [0, 1, 255, 256], 2, 20),     # one-row stripes at both ends
    (1024, 1024, [0, 256, 512, 768, 1024], 16, 30),
])
def test_peer_coupled_stripes_side_by_side_on_one_gpu(oracle, peer_kernel, W, H, cuts, nsrc, pct):
    elev, flags = synth.random_grid(W, H, seed_k=W + len(cuts), wall_pct=pct)
    srcs = synth.pick_open_tiles(flags, nsrc, 2)
    tiles, d, rounds = _peer_striped_on_one_gpu(W, H, cuts, elev, flags, srcs, repeat=2, kernel=peer_kernel)  # the second run reuses the peer blocks
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do), diff_report(d, do, W)
    assert np.array_equal(tiles, to), diff_report(tiles, to, W)
    assert rounds is None or rounds >= 1


def test_peer_coupled_stripes_maze_and_sources_on_cut_rows(oracle, peer_kernel):
    W, H = 320, 256
    elev, flags = synth.maze_grid(W, H, seed_k=4)
    srcs = synth.pick_open_tiles(flags, 1, 1)
    tiles, d, _ = _peer_striped_on_one_gpu(W, H, [0, 64, 128, 192, 256], elev, flags, srcs, kernel=peer_kernel)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do), diff_report(d, do, W)
    assert np.array_equal(tiles, to), diff_report(tiles, to, W)
    W, H = 192, 192
    elev, flags = synth.random_grid(W, H, seed_k=9, wall_pct=25)
    flags = flags.copy()
    srcs = np.array([63 * W + 10, 64 * W + 100, 127 * W + 191, 128 * W + 0], np.uint32)
    flags[srcs] &= ~np.uint16(synth.FLAG_BUILDING | 1)
    flags[srcs[1]] |= synth.FLAG_BUILDING  # a blocked source still expands (paths.cpp:63-67)
    tiles, d, _ = _peer_striped_on_one_gpu(W, H, [0, 64, 128, 192], elev, flags, srcs, kernel=peer_kernel)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do), diff_report(d, do, W)
    assert np.array_equal(tiles, to), diff_report(tiles, to, W)


def test_peer_coupled_stripes_ties_across_the_cuts(oracle, peer_kernel):
    """flat terrain: every edge weighs 255, so most tiles have several optimal predecessors and the next hop is decided by
    the tie-break (smallest label, then smallest tile id: paths.cpp:16-17, 69) -- across a cut that needs the neighbour's
    boundary row to be exact even where it never improved anything on this side (found at 8 ranks, profiles/r03i_*)"""
    W, H = 256, 256
    _, flags = synth.random_grid(W, H, seed_k=31, wall_pct=15)
    elev = np.zeros(W * H, np.int16)
    srcs = synth.pick_open_tiles(flags, 3, 7)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    for cuts in ([0, 64, 128, 192, 256], [0, 100, 256]):
        tiles, d, _ = _peer_striped_on_one_gpu(W, H, cuts, elev, flags, srcs, kernel=peer_kernel)
        assert np.array_equal(d, do), diff_report(d, do, W)
        assert np.array_equal(tiles, to), diff_report(tiles, to, W)


def test_peer_coupled_protocol_with_a_single_rank(oracle, peer_kernel):
    """world size 1: no neighbour to write to, but the update still runs through the peer-coupled round protocol (barrier
    epochs, totals in the rank's own block, termination on the global queue length), twice on the same peer block"""
    from mazetd_b200 import striped

    W, H = 320, 200
    elev, flags = synth.random_grid(W, H, seed_k=12, wall_pct=30)
    srcs = synth.pick_open_tiles(flags, 3, 5)
    st = striped.GpuStripe(0, W, H, 0, H, elev, flags, srcs)
    st.ctx.debug_set(st.ctx.DEBUG_BIG_ASYNC, 1 if peer_kernel == "dataflow" else 0)
    st.ctx.stripe_peer_connect(0, 1, [st.ctx.stripe_peer_export()])
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    for _ in range(2):
        tiles, d, rounds = striped.solve_peer(st)
        d = d.cpu().numpy().view(np.uint32).reshape(-1)
        tiles = tiles.cpu().numpy().view(np.uint32).reshape(-1)
        assert np.array_equal(d, do), diff_report(d, do, W)
        assert np.array_equal(tiles, to), diff_report(tiles, to, W)
        assert st.ctx.big_stats()["dataflow"] == (peer_kernel == "dataflow")
        assert rounds >= 1 or peer_kernel == "dataflow"
    st.close()


def test_peer_coupled_stripes_on_two_gpus(oracle):
    """the stripes of two ranks run as ONE kernel each, write their boundary rows into each other's halo copies over
    NVLink and meet at device-side round barriers (tests/peer_stripes_worker.py): the production set-up, one process
    per GPU with the peer blocks mapped through CUDA IPC. The one-GPU tests above cover the same kernels in one process."""
    import subprocess
    import sys

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs: gpurun --gpus 2 -- python -m pytest tests -m gpu -k two_gpus")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", "29641",
           os.path.join(root, "tests", "peer_stripes_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0 and "PEER-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("budget", [1, 3])
def test_stripes_with_a_round_budget_swap_rows_before_local_convergence(oracle, budget):
    W, H = 512, 512
    elev, flags = synth.random_grid(W, H, seed_k=21, wall_pct=30)
    srcs = synth.pick_open_tiles(flags, 3, 8)
    tiles, d, sweeps, _ = _striped_on_one_gpu(W, H, [0, 128, 256, 512], elev, flags, srcs, round_budget=budget)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do), diff_report(d, do, W)
    assert np.array_equal(tiles, to), diff_report(tiles, to, W)
    assert sweeps > 4


def test_stripes_maze_crosses_the_cut_many_times(oracle):
    """a maze makes shortest paths wander across the cuts: many sweeps, same fixpoint"""
    W, H = 320, 256
    elev, flags = synth.maze_grid(W, H, seed_k=4)
    srcs = synth.pick_open_tiles(flags, 1, 1)
    tiles, d, sweeps, _ = _striped_on_one_gpu(W, H, [0, 64, 128, 192, 256], elev, flags, srcs)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do), diff_report(d, do, W)
    assert np.array_equal(tiles, to), diff_report(tiles, to, W)
    assert sweeps > 2


def test_stripes_sources_on_cut_rows_and_blocked_source(oracle):
    W, H = 192, 192
    elev, flags = synth.random_grid(W, H, seed_k=9, wall_pct=25)
    flags = flags.copy()
    srcs = np.array([63 * W + 10, 64 * W + 100, 127 * W + 191, 128 * W + 0], np.uint32)
    flags[srcs] &= ~np.uint16(synth.FLAG_BUILDING | 1)
    flags[srcs[1]] |= synth.FLAG_BUILDING  # a blocked source still expands (paths.cpp:63-67)
    tiles, d, _, _ = _striped_on_one_gpu(W, H, [0, 64, 128, 192], elev, flags, srcs)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do), diff_report(d, do, W)
    assert np.array_equal(tiles, to), diff_report(tiles, to, W)


def test_stripe_context_rejects_whole_grid_calls():
    from mazetd_b200 import capi

    c = capi.Context(0)
    W, H = 128, 128
    elev, flags = synth.random_grid(W, H, seed_k=1)
    c.stripe_upload(W, H, 0, 64, elev.reshape(H, W)[0:65], flags.reshape(H, W)[0:65])
    with pytest.raises(capi.MtdError):
        c.directions_update(np.array([5], np.uint32))
    c.close()


def test_directions_layers_dev_matches_oracle_per_layer(ctx, oracle):
    """mtd_directions_layers_dev: n independent single-target layers on a tiled grid (the unit sharded across GPUs in C5)"""
    import torch

    W, H = 320, 272
    elev, flags = synth.random_grid(W, H, seed_k=6, wall_pct=30)
    targets = synth.pick_open_tiles(flags, 5, 4)
    ctx.grid_upload(W, H, elev, flags)
    dt = torch.empty((5, W * H), dtype=torch.int32, device="cuda")
    dd = torch.empty((5, W * H), dtype=torch.int32, device="cuda")
    ctx.layers_update_dev(targets, dt.data_ptr(), dd.data_ptr())
    ctx.synchronize()
    for l in range(5):
        to, do = oracle.directions_update(W, H, elev, flags, targets[l:l + 1])
        assert np.array_equal(dd[l].cpu().numpy().view(np.uint32), do), f"layer {l}: " + diff_report(dd[l].cpu().numpy().view(np.uint32), do, W)
        assert np.array_equal(dt[l].cpu().numpy().view(np.uint32), to)


# ---------------------------------------------------------------------------------------------------
# M5-M7 without the noise: newGrid stages on caller-supplied elevations (pinned by the compiled reference's integer code)
# ---------------------------------------------------------------------------------------------------
def test_grid_from_elevations_synthetic_and_golden(ctx, oracle):
    rng = np.random.default_rng(12)
    W = H = 101
    maps = []
    # smooth random terrains (sum of a few sinusoids) so slopes straddle the 89 threshold and several components appear
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float64)
    for k in range(6):
        z = np.zeros((H, W))
        for _ in range(5):
            fx, fy, ph, amp = rng.uniform(0.02, 0.35), rng.uniform(0.02, 0.35), rng.uniform(0, 6.28), rng.uniform(100, 900)
            z += amp * np.sin(fx * xx + fy * yy + ph)
        maps.append(np.clip(z - 300 * k, -25500, 25500).astype(np.int16).reshape(-1))
    maps.append(rng.integers(-3000, 3000, W * H).astype(np.int16))  # white noise: nearly everything Invalid
    maps.append(np.zeros(W * H, np.int16))                          # flat: one disc-shaped component
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "mapgen_k0_k7.npz"))
    if "elev" in g.files:
        maps.extend(list(np.asarray(g["elev"]).reshape(-1, W * H)[:2]))
    elev = np.stack(maps)
    got = ctx.grid_from_elevations(elev)
    for k in range(len(maps)):
        want = oracle.grid_stages(W, H, 50, 50, 45, elev[k])
        assert np.array_equal(got[k], want), f"map {k}: " + diff_report(got[k], want, W)


def test_grid_from_elevations_ragged_size_and_ties(ctx, oracle):
    """non-square grid, radius test disabled, two equal-sized components: the first discovered (lowest index) wins"""
    W, H = 37, 23
    elev = np.zeros(W * H, np.int16).reshape(H, W)
    elev[:, 18] = 5000  # a wall of slope splits the grid into two 18-column halves (columns 17..19 become Invalid)
    got = ctx.grid_from_elevations(elev.reshape(1, -1), W=W, H=H, ox=0, oy=0, radius=-1)[0]
    want = oracle.grid_stages(W, H, 0, 0, -1, elev.reshape(-1))
    assert np.array_equal(got, want), diff_report(got, want, W)
    valid = (got & synth.FLAG_INVALID) == 0
    assert valid.reshape(H, W)[:, :17].all() and not valid.reshape(H, W)[:, 19:].any()


# ---------------------------------------------------------------------------------------------------
# C3 at reduced size: scripted tower place/remove events with monsters steering along the flow field, every piece of
# state compared with the oracle after the script
# ---------------------------------------------------------------------------------------------------
from oracle_lib import OracleGame as _OracleGame  # noqa: E402  (tests/golden/make_golden.py imports it from here)


def test_scripted_place_remove_events_with_monsters(ctx, oracle):
    from mazetd_b200 import game_events

    o = oracle.mapgen(KAT_SEED[0], KAT_SEED[1])
    W = H = 101
    ctx.grid_upload(W, H, o["elev"], o["flags"], 50, 50)
    ctx.waypoints_update_dev(o["wp_tiles"])
    sp = int(o["agg"][2])
    mon = game_events.initial_monsters(3000, o["fullpath"][sp][: int(o["fullpath_len"][sp])], o["wp_tiles"], ctx.waypoints_find)
    og = _OracleGame(oracle, W, H, o["elev"], o["flags"], o["wp_tiles"], {k: v.copy() for k, v in mon.items()})
    gg = game_events.GpuGame(ctx, W, H, o["elev"], o["flags"], o["wp_tiles"])
    ctx.monsters_upload(mon)
    sg = game_events.run_events(gg, 400)
    so = game_events.run_events(og, 400)
    assert sg == so and sg["placed"] > 20 and sg["cleared"] >= 1, (sg, so)
    assert np.array_equal(gg.flags, og.flags)
    got = ctx.monsters_download()
    for k in got:
        assert np.array_equal(got[k], og.a[k]), k
    w = ctx.waypoints_update(o["wp_tiles"], cap=2048)  # same grid state -> same layers as the oracle's last rebuild
    for key in ("dist", "tiles", "full_distance", "agg"):
        assert np.array_equal(w[key], og.w[key]), key


@pytest.mark.parametrize("delta", [1, 700, 20000, 0xFFFFFFFF])
def test_tiled_priority_window_never_changes_the_result(oracle, delta):
    """the tile-level priority (which queued tiles a round visits) is a schedule, not part of the answer"""
    from mazetd_b200 import capi

    c = capi.Context(0)
    c.debug_set(c.DEBUG_TILE_DELTA, delta)
    W, H = 448, 352
    elev, flags = synth.random_grid(W, H, seed_k=31, wall_pct=33)
    srcs = synth.pick_open_tiles(flags, 3, 6)
    c.grid_upload(W, H, elev, flags)
    t, d = c.directions_update(srcs)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do), diff_report(d, do, W)
    assert np.array_equal(t, to), diff_report(t, to, W)
    w = c.waypoints_update(srcs, cap=64)
    wo = oracle.waypoints_update(W, H, elev, flags, srcs, cap=64)
    assert np.array_equal(w["dist"], wo["dist"]) and np.array_equal(w["tiles"], wo["tiles"])
    c.close()


# ---------------------------------------------------------------------------------------------------
# terrain SDF over the marching-cubes lattice (headless buffer; meshing stays out of scope)
# ---------------------------------------------------------------------------------------------------
def test_sdf_lattice_bit_exact(ctx, oracle):
    o = oracle.mapgen(KAT_SEED[0] + 4, KAT_SEED[1])
    for res in ((220, 60, 220), (33, 9, 47)):
        got = ctx.sdf_lattice(KAT_SEED[0] + 4, KAT_SEED[1], o["flags"], res=res)
        want = oracle.sdf_lattice(KAT_SEED[0] + 4, KAT_SEED[1], 101, 101, 50, 50, o["flags"], res=res)
        bad = np.flatnonzero(got.view(np.uint32).reshape(-1) != want.view(np.uint32).reshape(-1))
        assert bad.size == 0, f"{bad.size} of {got.size} lattice values differ; first at {int(bad[0])}: {got.reshape(-1)[bad[0]]} vs {want.reshape(-1)[bad[0]]}"
    # the cliff term only shows on invalid land tiles below y = -6.5: it must be present somewhere
    flat = ctx.sdf_lattice(KAT_SEED[0] + 4, KAT_SEED[1], np.zeros(101 * 101, np.uint16))
    assert (got.shape == (47, 9, 33)) and not np.array_equal(flat, ctx.sdf_lattice(KAT_SEED[0] + 4, KAT_SEED[1], o["flags"]))


def test_sdf_lattice_vs_reference_code(ctx, ref):
    sr, flags = ref.sdf_lattice(KAT_SEED[0] + 9, KAT_SEED[1])
    got = ctx.sdf_lattice(KAT_SEED[0] + 9, KAT_SEED[1], flags)
    assert np.array_equal(got.view(np.uint32), sr.view(np.uint32))


# ---------------------------------------------------------------------------------------------------
# damage-over-time tick on the monster SoA
# ---------------------------------------------------------------------------------------------------
def test_dots_tick_bit_exact_and_feeds_the_steering_bits(ctx, oracle):
    o = oracle.mapgen(KAT_SEED[0] + 2, KAT_SEED[1])
    W = H = 101
    ctx.grid_upload(W, H, o["elev"], o["flags"], 50, 50)
    ctx.waypoints_update(o["wp_tiles"])
    wv = oracle.waypoints_view(W, H, o["elev"], o["flags"], o["wp_tiles"], o["tiles"], o["dist"], 50, 50)
    n = 4099
    rng = np.random.default_rng(17)
    sp = int(o["agg"][2])
    path = o["fullpath"][sp][: int(o["fullpath_len"][sp])]
    tile = rng.choice(path[:-5], n).astype(np.uint32)
    a = dict(tile=tile.copy(), tileStart=tile.copy(), tileEnd=tile.copy(), timeStart=np.zeros(n, np.uint32), timeEnd=np.zeros(n, np.uint32),
             bits=np.full(n, 1 << sp, np.uint32), timeToArrive=np.full(n, 5000, np.uint32), damage=np.ones(n, np.uint32), alive=(rng.integers(0, 20, n) > 0).astype(np.uint8))
    dur = rng.integers(0, 40, (5, n)).astype(np.uint32) * (rng.integers(0, 3, (5, n)) > 0)
    dmg = rng.integers(0, 3000, (5, n)).astype(np.uint32) * (rng.integers(0, 3, (5, n)) > 0)
    dmg[0, :7] = 0xFFFFFFF0  # the x2 of a strengthened dot wraps in uint32, as in the reference
    dmg[3, :7] = 5
    life = rng.integers(1, 4000, n).astype(np.int64)
    resist = rng.integers(0, 32, n).astype(np.uint8)
    money = rng.integers(0, 100, n).astype(np.uint32)
    ctx.monsters_upload(a)
    ctx.monsters_dots_upload(dur, dmg, life, resist, money)
    a["dotflags"] = np.zeros(n, np.uint8)
    total_g = total_o = 0
    for t in range(1, 120):
        # the steering tick reads affected(Fire) / affected(Water): derive them from the oracle's dots like the device does
        a["dotflags"][:] = (dmg[1] > 0).astype(np.uint8) | ((dmg[2] > 0).astype(np.uint8) << 1)
        ctx.monsters_tick(t)
        oracle.monsters_tick(wv, t, a)
        mg, kg = ctx.monsters_dots_tick()
        mo, ko = oracle.monsters_dots_tick(dur, dmg, life, a["alive"], resist, money)
        assert (mg, kg) == (mo, ko), t
        total_g += kg
        if t % 17 == 0 or t == 119:
            d = ctx.monsters_dots_download()
            assert np.array_equal(d["duration"], dur) and np.array_equal(d["damage"], dmg), t
            assert np.array_equal(d["life"], life) and np.array_equal(d["alive"], a["alive"]), t
            got = ctx.monsters_download()
            for k in got:
                assert np.array_equal(got[k], a[k]), (t, k)
    assert total_g > 100


def test_sdf_and_game_events_against_committed_golden(ctx):
    """no oracle in the loop: the device against fixtures written from the compiled reference (tests/golden/make_golden.py)"""
    from mazetd_b200 import game_events

    gold = os.path.join(os.path.dirname(__file__), "golden")
    g = np.load(os.path.join(gold, "sdf_lattice.npz"))
    got = ctx.sdf_lattice(KAT_SEED[0] + int(g["seed_k"]), KAT_SEED[1], g["flags"], res=tuple(int(v) for v in g["res"]))
    assert np.array_equal(got.view(np.uint32), g["sdf"].view(np.uint32))

    g = np.load(os.path.join(gold, "game_events.npz"))
    m = ctx.mapgen_batch(KAT_SEED[0], KAT_SEED[1], 0, 1, cap=2048)
    mon = {k[5:]: g[k].copy() for k in g.files if k.startswith("mon0_")}
    game = game_events.GpuGame(ctx, 101, 101, m["elev"][0], m["flags"][0], m["wp_tiles"][0])
    ctx.monsters_upload(mon)
    stats = game_events.run_events(game, 300)
    assert [stats[k] for k in ("placed", "cleared", "rejected_flags", "rejected_trial")] == list(g["stats"])
    assert np.array_equal(game.flags, g["flags"])
    got = ctx.monsters_download()
    for k in got:
        assert np.array_equal(got[k], g["mon_" + k]), k
    w = ctx.waypoints_update(m["wp_tiles"][0], cap=4096)
    assert np.array_equal(w["agg"], g["agg"]) and np.array_equal(w["full_distance"], g["full_distance"])
    assert [int(w["dist"][l].astype(np.uint64).sum() & 0xFFFFFFFF) for l in range(4)] == list(g["dist_sum"])


# ---------------------------------------------------------------------------------------------------
# monster spatial index + tower attack pass
# ---------------------------------------------------------------------------------------------------
def _random_towers(rng, n, flags, W):
    open_tiles = np.flatnonzero((flags & synth.FLAG_INVALID) == 0)
    tw = dict(
        tile=rng.choice(open_tiles, n).astype(np.uint32),
        damage=rng.integers(1, 400, n).astype(np.uint32),
        overTime=np.zeros(n, np.uint32),
        firingPeriod=rng.integers(1, 40, n).astype(np.uint32),
        firingRange=rng.choice(np.array([0.5, 1.0, 2.0, 3.5, 5.0, 9.0], np.float32), n),
        splashRadius=rng.choice(np.array([0.0, 0.0, 1.0, 2.5], np.float32), n),
        manaCost=rng.choice(np.array([0, 5, 12, 40], np.uint32), n),
        invalidClasses=rng.choice(np.array([0, 0, 1, 2, 3], np.uint8), n),
        element=rng.integers(0, 5, n).astype(np.uint8),
        bonus=rng.integers(0, 7, n).astype(np.uint8),
        targeting=rng.integers(0, 7, n).astype(np.uint8),
        firingDelay=rng.integers(0, 4, n).astype(np.uint32),
        mana=rng.integers(0, 400, n).astype(np.uint32),
    )
    tw["overTime"][tw["element"] == 0] = rng.integers(0, 30, int((tw["element"] == 0).sum())).astype(np.uint32)
    return tw


@pytest.mark.parametrize("nmon,ntow", [(3000, 257), (60000, 1000)])
def test_towers_attack_pass_matches_oracle(ctx, oracle, nmon, ntow):
    """spatial rebuild + attacks (sources/game/spatial.cpp:24-34, attacks.cpp:132-160) on moving monsters, interleaved with
    steering ticks: picked targets, hit counts, tower delays / mana and the monsters' damage-over-time state, bit for bit"""
    from mazetd_b200 import game_events

    o = oracle.mapgen(KAT_SEED[0] + 3, KAT_SEED[1])
    W = H = 101
    ctx.grid_upload(W, H, o["elev"], o["flags"], 50, 50)
    ctx.waypoints_update(o["wp_tiles"])
    wv = oracle.waypoints_view(W, H, o["elev"], o["flags"], o["wp_tiles"], o["tiles"], o["dist"], 50, 50)
    sp = int(o["agg"][2])
    path = o["fullpath"][sp][: int(o["fullpath_len"][sp])]
    rng = np.random.default_rng(nmon)
    a = game_events.initial_monsters(nmon, path, o["wp_tiles"], ctx.waypoints_find, speed_range=(0.05, 0.2))
    a["alive"] = (rng.integers(0, 25, nmon) > 0).astype(np.uint8)
    # distinct picker values wherever the reference's own order would decide (ties go to the lowest index on both sides anyway)
    life = rng.integers(1, 1 << 40, nmon).astype(np.int64)
    cls = rng.choice(np.array([0, 0, 0, 1, 2], np.uint8), nmon)
    dur = np.zeros((5, nmon), np.uint32)
    dmg = np.zeros((5, nmon), np.uint32)
    tw = _random_towers(rng, ntow, o["flags"], W)
    ctx.monsters_upload(a)
    ctx.monsters_dots_upload(dur, dmg, life)
    ctx.monsters_class_upload(cls)
    ctx.towers_upload(tw)
    a["dotflags"] = np.zeros(nmon, np.uint8)
    fired = 0
    for t in range(1, 13):
        ctx.monsters_tick(t)
        oracle.monsters_tick(wv, t, a)
        words = rng.integers(0, 1 << 62, ntow).astype(np.uint64)
        ctx.towers_attack_dev(t, words)
        target, hits = oracle.towers_attack(W, H, 50, 50, o["elev"], t, tw, a, cls, life, dur, dmg, random_words=words)
        got = ctx.towers_download()
        assert np.array_equal(got["target"], target), (t, np.flatnonzero(got["target"] != target)[:5], tw["targeting"][np.flatnonzero(got["target"] != target)[:5]])
        assert np.array_equal(got["hits"], hits), t
        assert np.array_equal(got["firingDelay"], tw["firingDelay"]) and np.array_equal(got["mana"], tw["mana"]), t
        fired += int((target >= 0).sum())
        d = ctx.monsters_dots_download()
        assert np.array_equal(d["damage"], dmg) and np.array_equal(d["duration"], dur), t
        a["dotflags"][:] = (dmg[1] > 0).astype(np.uint8) | ((dmg[2] > 0).astype(np.uint8) << 1)
        if t % 4 == 0:
            # the damage pass of the same tick, then steering reads the refreshed fire / water bits
            ctx.monsters_dots_tick()
            oracle.monsters_dots_tick(dur, dmg, life, a["alive"])
    assert fired > ntow // 4  # towers really fired
    assert (hits > 1).any()  # and splashes hit several monsters


# ---------------------------------------------------------------------------------------------------
# incremental re-pathing (one blocking flip -> the resident layers are repaired)
# ---------------------------------------------------------------------------------------------------
def test_incremental_repathing_matches_full_updates(ctx, oracle):
    """random single place / remove flips on a generated map, each followed by Waypoints::update through the repair path
    (the statistics prove it was taken) and compared with the oracle's from-scratch result: all layers, next hops, paths,
    aggregates; plus the trial of placingBuildingBlocksMonsters, which reuses layer 0 (buildingsPlacement.cpp:18-42)"""
    from mazetd_b200 import game_events

    o = oracle.mapgen(KAT_SEED[0] + 5, KAT_SEED[1], cap=2048)
    W = H = 101
    flags = o["flags"].copy()
    wp = o["wp_tiles"]
    ctx.debug_set(ctx.DEBUG_INCREMENTAL, 1)
    ctx.grid_upload(W, H, o["elev"], flags, 50, 50)
    ctx.waypoints_update(wp, cap=2048)
    rng = np.random.default_rng(11)
    sp = int(o["agg"][2])
    path = o["fullpath"][sp][: int(o["fullpath_len"][sp])]
    mon = game_events.initial_monsters(500, path, wp, ctx.waypoints_find)
    ctx.monsters_upload(mon)
    passable = np.flatnonzero((flags & (synth.FLAG_INVALID | synth.FLAG_WAYPOINT)) == 0)
    on_path = [int(t) for t in path[3:-3:7] if (flags[t] & synth.FLAG_WAYPOINT) == 0]
    cands = on_path[:12] + [int(t) for t in rng.choice(passable, 40)]
    base = ctx.big_stats()
    placed = []
    checked_trials = 0
    for step, t in enumerate(cands):
        if flags[t] & synth.FLAG_BUILDING:
            continue
        # trial on the current grid (layers are up to date: the repair path of layer 0)
        got = ctx.placing_blocks(t)
        want = oracle.placing_blocks(W, H, o["elev"], flags, wp, t, mon["tile"], mon["alive"], 50, 50)
        assert bool(got) == bool(want), (step, t)
        checked_trials += 1
        if want:
            continue
        flags[t] |= synth.FLAG_BUILDING
        ctx.grid_set_flags(t, int(flags[t]))
        placed.append(t)
        w = ctx.waypoints_update(wp, cap=2048)
        wo = oracle.waypoints_update(W, H, o["elev"], flags, wp, 50, 50, cap=2048)
        for key in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
            assert np.array_equal(w[key], wo[key]), (step, t, key)
        if step % 3 == 2 and placed:
            # sell an earlier tower again: labels only drop, behind the freed tile
            r = placed.pop(int(rng.integers(0, len(placed))))
            flags[r] &= ~np.uint16(synth.FLAG_BUILDING)
            ctx.grid_set_flags(r, int(flags[r]))
            w = ctx.waypoints_update(wp, cap=2048)
            wo = oracle.waypoints_update(W, H, o["elev"], flags, wp, 50, 50, cap=2048)
            for key in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
                assert np.array_equal(w[key], wo[key]), (step, "sell", r, key)
    st = ctx.big_stats()
    assert st["rounds"] - base["rounds"] >= 20, st  # incremental updates taken
    assert st["visits"] - base["visits"] >= 10, st  # incremental trials taken
    assert checked_trials >= 30
    # the drop-in adapter's pattern: the whole grid is uploaded again with one flag changed -> still a repair
    t = int(passable[len(passable) // 2])
    if not (flags[t] & synth.FLAG_BUILDING):
        flags[t] |= synth.FLAG_BUILDING
        ctx.grid_upload(W, H, o["elev"], flags, 50, 50)
        before = ctx.big_stats()["rounds"]
        w = ctx.waypoints_update(wp, cap=2048)
        assert ctx.big_stats()["rounds"] == before + 1
        wo = oracle.waypoints_update(W, H, o["elev"], flags, wp, 50, 50, cap=2048)
        for key in ("dist", "tiles", "fullpath_len", "full_distance", "agg", "fullpath"):
            assert np.array_equal(w[key], wo[key]), ("re-upload", key)
    # two flips between updates: from scratch, same answer
    a, b = int(passable[7]), int(passable[len(passable) - 9])
    for t in (a, b):
        flags[t] ^= np.uint16(synth.FLAG_BUILDING)
        ctx.grid_set_flags(t, int(flags[t]))
    before = ctx.big_stats()["rounds"]
    w = ctx.waypoints_update(wp, cap=2048)
    assert ctx.big_stats()["rounds"] == before
    wo = oracle.waypoints_update(W, H, o["elev"], flags, wp, 50, 50, cap=2048)
    for key in ("dist", "tiles", "agg"):
        assert np.array_equal(w[key], wo[key]), ("two flips", key)


def test_scripted_events_at_a_larger_size(ctx, oracle):
    """the C3 script two orders of magnitude above the golden fixture's size: 1 500 events x 200 000 monsters (3e8 monster
    ticks, several hundred rebuilds and trials through the incremental path), every integer of the final state against the
    oracle; the bench runs 10 000 x 1 000 000 of the same script"""
    from mazetd_b200 import game_events
    from oracle_lib import OracleGame

    o = oracle.mapgen(KAT_SEED[0], KAT_SEED[1])
    W = H = 101
    ctx.grid_upload(W, H, o["elev"], o["flags"], 50, 50)
    ctx.waypoints_update_dev(o["wp_tiles"])
    sp = int(o["agg"][2])
    mon = game_events.initial_monsters(200000, o["fullpath"][sp][: int(o["fullpath_len"][sp])], o["wp_tiles"], ctx.waypoints_find, speed_range=(0.03, 0.3))
    og = OracleGame(oracle, W, H, o["elev"], o["flags"], o["wp_tiles"], {k: v.copy() for k, v in mon.items()})
    gg = game_events.GpuGame(ctx, W, H, o["elev"], o["flags"], o["wp_tiles"])
    ctx.monsters_upload(mon)
    sg = game_events.run_events(gg, 1500)
    so = game_events.run_events(og, 1500)
    assert sg == so and sg["placed"] > 100 and sg["cleared"] >= 5, (sg, so)
    assert np.array_equal(gg.flags, og.flags)
    got = ctx.monsters_download()
    for k in got:
        assert np.array_equal(got[k], og.a[k]), k
    w = ctx.waypoints_update(o["wp_tiles"], cap=2048)
    for key in ("dist", "tiles", "full_distance", "agg"):
        assert np.array_equal(w[key], og.w[key]), key
