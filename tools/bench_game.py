"""C1 and C3 of BASELINE.json on one GPU (latency-bound configurations):
    C1  default generated map (seed k = 0): trial + full rebuild after a scripted tower placement, through the host C ABI
    C3  10 000 place/remove events on the default map with 1 000 000 monsters ticking between events
Usage: python tools/bench_game.py [events] [monsters]   -> JSON lines"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mazetd_b200 import capi, game_events, synth  # noqa: E402


def main():
    nevents = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
    nmon = int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
    ctx = capi.Context(0)
    s1, s2 = synth.KAT_SEED
    g = ctx.mapgen_batch(s1, s2, 0, 1, cap=4096)
    W = H = 101
    elev, flags, wp = g["elev"][0], g["flags"][0], g["wp_tiles"][0]
    sp = int(g["agg"][0][2])
    path = g["fullpath"][0][sp][: int(g["fullpath_len"][0][sp])]

    # ---- C1: first tile (ascending index) on the min spawner's path that passes the flags test of canPlaceBuilding ----
    ctx.grid_upload(W, H, elev, flags, 50, 50)
    ctx.waypoints_update_dev(wp)
    on_path = np.zeros(W * H, bool)
    on_path[path] = True
    cand = int(np.flatnonzero(on_path & ((flags & game_events.UNBUILDABLE) == 0))[0])
    lat = []
    for rep in range(200):
        f2 = flags.copy()
        t0 = time.perf_counter()
        blocked = ctx.placing_blocks(cand)                   # P5: one trial Dijkstra
        f2[cand] |= synth.FLAG_BUILDING
        ctx.grid_set_flags(cand, int(f2[cand]))              # P6: flag delta H2D
        w = ctx.waypoints_update(wp, cap=4096)               # P2: 4 layers + walks + aggregates, D2H of everything
        lat.append(time.perf_counter() - t0)
        ctx.grid_set_flags(cand, int(flags[cand]))
    lat = np.array(lat[20:]) * 1e6
    print(json.dumps(dict(config="C1", candidate=cand, trial_blocked=bool(blocked), us_median=float(np.median(lat)), us_p90=float(np.percentile(lat, 90)),
                          note="placing_blocks + grid_set_flags + waypoints_update with host outputs (326 KB D2H), wall clock around the synchronous C-ABI calls")), flush=True)
    lat = []
    for rep in range(200):
        t0 = time.perf_counter()
        ctx.waypoints_update_dev(wp)
        ctx.synchronize()
        lat.append(time.perf_counter() - t0)
    lat = np.array(lat[20:]) * 1e6
    print(json.dumps(dict(config="C1-resident", us_median=float(np.median(lat)), note="waypoints_update_dev + synchronize: layers stay on the device")), flush=True)

    # ---- C3 ----
    ctx.grid_upload(W, H, elev, flags, 50, 50)
    ctx.waypoints_update_dev(wp)
    # speeds 10x below SURVEY.md 8d's [0.3, 0.6]: with that range most monsters reach the last waypoint within ~7 000 ticks and
    # the later ticks would run on an almost empty population; this range keeps every monster alive for the whole script
    mon = game_events.initial_monsters(nmon, path, wp, ctx.waypoints_find, speed_range=(0.03, 0.06))
    game = game_events.GpuGame(ctx, W, H, elev, flags, wp)
    ctx.monsters_upload(mon)
    l0 = ctx.launches
    ctx.synchronize()
    t0 = time.perf_counter()
    stats = game_events.run_events(game, nevents)
    dmg, alive = ctx.monsters_tick(nevents + 1)
    wall = time.perf_counter() - t0
    print(json.dumps(dict(config="C3", events=nevents, monsters=nmon, wall_s=wall, events_per_s=nevents / wall, monster_ticks_per_s=nmon * nevents / wall,
                          rebuilds=stats["placed"] + stats["cleared"], trials=stats["placed"] + stats["rejected_trial"], stats=stats, alive=int(alive), launches=ctx.launches - l0,
                          steering_algorithmic_GBps=28.0 * nmon * nevents / wall / 1e9)), flush=True)


if __name__ == "__main__":
    main()
