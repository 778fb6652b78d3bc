"""Host-side multi-GPU logic on CPU: partition functions, the row-stripe protocol in one process, and the same protocol /
the per-rank map ranges over torch.distributed with the gloo backend at world_size 2 and 3."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mazetd_b200 import shard, striped, synth

from stripe_cpu_solver import CpuStripe


def test_map_range_partitions_everything_in_order():
    for nmaps in (0, 1, 7, 4096, 4099):
        for world in (1, 2, 3, 4, 8):
            spans = [shard.map_range(r, world, nmaps) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == nmaps
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_layer_assignment_covers_each_layer_once():
    for nlayers in (1, 4, 8, 13):
        for world in (1, 2, 4, 8):
            got = sorted(l for r in range(world) for l in shard.layer_assignment(r, world, nlayers))
            assert got == list(range(nlayers))


def test_stripe_rows_cover_grid_and_align_to_tiles():
    for H in (2, 63, 64, 320, 4096, 16384, 1000):
        for world in (1, 2, 3, 4, 8):
            if H < world:
                with pytest.raises(ValueError):
                    shard.stripe_rows(0, world, H)
                continue
            spans = [shard.stripe_rows(r, world, H) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == H
            assert all(a[1] == b[0] and a[0] < a[1] for a, b in zip(spans, spans[1:] + [(H, H + 1)]))
            if H // 64 >= world:
                assert all(a[0] % 64 == 0 for a in spans)
    assert shard.stripe_halo_rows(0, 64, 128) == (0, 65)
    assert shard.stripe_halo_rows(64, 128, 128) == (63, 128)
    with pytest.raises(ValueError):
        shard.map_range(2, 2, 10)


def _case(W, H, seed_k, nsrc, wall_pct=30):
    elev, flags = synth.random_grid(W, H, seed_k=seed_k, wall_pct=wall_pct)
    srcs = synth.pick_open_tiles(flags, nsrc, seed_k)
    return elev, flags, srcs


def _stack(results):
    tiles = np.concatenate([t.numpy().view(np.uint32).reshape(-1) for t, _ in results])
    d = np.concatenate([x.numpy().view(np.uint32).reshape(-1) for _, x in results])
    return tiles, d


@pytest.mark.parametrize("W,H,world,nsrc", [(40, 37, 2, 1), (33, 50, 3, 4), (24, 64, 4, 2), (17, 9, 8, 1)])
def test_stripe_protocol_in_one_process_matches_whole_grid(oracle, W, H, world, nsrc):
    elev, flags, srcs = _case(W, H, 11 + world, nsrc)
    solvers = []
    for r in range(world):
        y0, y1 = shard.stripe_rows(r, world, H, align=8)
        solvers.append(CpuStripe(oracle, W, H, y0, y1, elev, flags, srcs))
    results, sweeps = striped.solve_local(solvers)
    tiles, d = _stack(results)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do)
    assert np.array_equal(tiles, to)
    assert sweeps >= 1


def test_stripe_protocol_source_on_cut_and_blocked_source(oracle):
    W, H = 20, 16
    elev, flags, _ = _case(W, H, 5, 1)
    flags = flags.copy()
    # one source on the last row of stripe 0, one (blocked) source on the first row of stripe 1
    srcs = np.array([7 * W + 3, 8 * W + 15], np.uint32)
    flags[srcs[1]] |= synth.FLAG_BUILDING
    flags[srcs[0]] &= ~np.uint16(synth.FLAG_BUILDING | 1)
    solvers = [CpuStripe(oracle, W, H, 0, 8, elev, flags, srcs), CpuStripe(oracle, W, H, 8, 16, elev, flags, srcs)]
    results, _ = striped.solve_local(solvers)
    tiles, d = _stack(results)
    to, do = oracle.directions_update(W, H, elev, flags, srcs)
    assert np.array_equal(d, do) and np.array_equal(tiles, to)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, W, H, nsrc, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle_lib

        orc = oracle_lib.Oracle()
        elev, flags, srcs = _case(W, H, 3, nsrc)
        y0, y1 = shard.stripe_rows(rank, world, H, align=8)
        tiles, d, sweeps = striped.solve_distributed(CpuStripe(orc, W, H, y0, y1, elev, flags, srcs), rank, world)
        # rank 0 gathers the stripes and compares with one Dijkstra over the whole grid
        parts = [None] * world
        dist.gather_object((y0, tiles.numpy().view(np.uint32).reshape(-1), d.numpy().view(np.uint32).reshape(-1), sweeps), parts if rank == 0 else None, dst=0)
        # independent maps: every rank fingerprints its own block of map indices, rank 0 checks the union
        a, b = shard.map_range(rank, world, 10)
        mine = [(k, int(orc.mapgen(oracle_lib.KAT_SEED[0] + k, oracle_lib.KAT_SEED[1], cap=16)["agg"].sum())) for k in range(a, b)]
        allmaps = [None] * world
        dist.gather_object(mine, allmaps if rank == 0 else None, dst=0)
        if rank == 0:
            parts.sort(key=lambda p: p[0])
            to, do = orc.directions_update(W, H, elev, flags, srcs)
            ok = np.array_equal(np.concatenate([p[2] for p in parts]), do) and np.array_equal(np.concatenate([p[1] for p in parts]), to)
            flat = [m for part in allmaps for m in part]
            ok_maps = [k for k, _ in flat] == list(range(10))
            out.put((bool(ok), bool(ok_maps), parts[0][3]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_stripe_protocol_and_map_ranges_over_gloo(oracle, world):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 30, 41, 3, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    ok, ok_maps, sweeps = out.get(timeout=10)
    assert ok, "striped distances / next hops differ from the whole-grid oracle"
    assert ok_maps
    assert sweeps >= 1


# ---- C3 over several ranks: monsters sharded by index, layers broadcast per accepted event ----
def _c3_inputs(orc, nmon):
    import oracle_lib
    from mazetd_b200 import game_events

    o = orc.mapgen(oracle_lib.KAT_SEED[0], oracle_lib.KAT_SEED[1])
    sp = int(o["agg"][2])
    view = orc.waypoints_view(101, 101, o["elev"], o["flags"], o["wp_tiles"], o["tiles"], o["dist"], 50, 50)
    find = lambda pos, bits: tuple(np.array(x, np.uint32) for x in zip(*[orc.find(view, int(p), int(b)) for p, b in zip(pos, bits)]))  # noqa: E731
    mon = game_events.initial_monsters(nmon, o["fullpath"][sp][: int(o["fullpath_len"][sp])], o["wp_tiles"], find)
    return o, mon


def _c3_worker(rank, world, port, mode, nmon, nevents, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle_lib
        from mazetd_b200 import game_events, sharded_game

        orc = oracle_lib.Oracle()
        o, mon = _c3_inputs(orc, nmon)
        mine = sharded_game.shard_monsters(mon, rank, world)
        local = oracle_lib.OracleGame(orc, 101, 101, o["elev"], o["flags"], o["wp_tiles"], mine)
        game = sharded_game.ShardedGame(local, rank, world, mode=mode)
        stats = game_events.run_events(game, nevents)
        parts = [None] * world
        dist.gather_object((rank, {k: v for k, v in mine.items()}, stats, local.flags.copy(), game.collectives), parts if rank == 0 else None, dst=0)
        if rank == 0:
            parts.sort(key=lambda p: p[0])
            merged = {k: np.concatenate([p[1][k] for p in parts]) for k in mine}
            out.put((merged, [p[2] for p in parts], [p[3] for p in parts], parts[0][4]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["broadcast", "replicate"])
def test_c3_monsters_sharded_over_gloo_match_one_process(oracle, mode):
    """two ranks, each ticking half of the monsters; trial verdicts OR-ed, layers broadcast from rank 0 (or rebuilt on
    every rank): the union of the shards equals one process running all monsters, event for event"""
    import oracle_lib
    from mazetd_b200 import game_events

    nmon, nevents, world = 1201, 70, 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_c3_worker, args=(r, world, port, mode, nmon, nevents, out)) for r in range(world)]
    for p in procs:
        p.start()
    merged, stats, flags, collectives = out.get(timeout=240)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    o, mon = _c3_inputs(oracle, nmon)
    one = oracle_lib.OracleGame(oracle, 101, 101, o["elev"], o["flags"], o["wp_tiles"], mon)
    want = game_events.run_events(one, nevents)
    assert all(s == want for s in stats), (stats, want)
    assert want["placed"] > 3
    assert all(np.array_equal(f, one.flags) for f in flags)
    for k in ("tile", "tileStart", "tileEnd", "timeStart", "timeEnd", "bits", "alive"):
        assert np.array_equal(merged[k], one.a[k]), k
    assert collectives > 0
