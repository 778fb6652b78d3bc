"""Generates tests/golden/*.npz from the REFERENCE'S OWN compiled code (oracle/_ref/libmazetd_ref.so, built by
`make -C oracle ref` from /root/reference in place). Run in the authoring container only:

    python tests/golden/make_golden.py

The fixtures are small (full int16/uint16 grids, checksums of the 4-byte layers) and travel to the GPU box, where
/root/reference does not exist. TEST INFRASTRUCTURE.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle_lib  # noqa: E402
from mazetd_b200 import synth  # noqa: E402

KAT = oracle_lib.KAT_SEED


def mapgen_golden(ref, n=8, cap=1024):
    keys = ("elev", "flags", "wp_tiles", "agg", "full_distance", "fullpath_len")
    acc = {k: [] for k in keys}
    dist_xor, dist_sum, tiles_sum, fullpath = [], [], [], []
    for k in range(n):
        r = ref.mapgen(KAT[0] + k, KAT[1], cap=cap)
        for key in keys:
            acc[key].append(r[key])
        dist_xor.append([int(np.bitwise_xor.reduce(r["dist"][l])) for l in range(4)])
        dist_sum.append([int(r["dist"][l].astype(np.uint64).sum() & 0xFFFFFFFF) for l in range(4)])
        tiles_sum.append([int(r["tiles"][l].astype(np.uint64).sum() & 0xFFFFFFFF) for l in range(4)])
        fullpath.append(r["fullpath"])
    out = {k: np.stack(v) for k, v in acc.items()}
    out["dist_xor"] = np.array(dist_xor, np.uint32)
    out["dist_sum"] = np.array(dist_sum, np.uint32)
    out["tiles_sum"] = np.array(tiles_sum, np.uint32)
    out["fullpath"] = np.stack(fullpath)
    return out


def directions_golden(ref):
    """Directions::update of the reference on synthetic grids (sources/map/paths.cpp:24-78): full layers of a
    67x45 walls grid and a maze, checksums + sampled values of a 300x260 grid (tiled path)."""
    out = {}
    for name, (W, H, elev, flags) in {
        "walls": (67, 45) + synth.random_grid(67, 45, seed_k=11, wall_pct=30, elev_mod=512),
        "maze": (67, 45) + synth.maze_grid(67, 45, 3),
    }.items():
        src = int(synth.pick_open_tiles(flags, 1, 2)[0])
        t, d = ref.directions_update(W, H, elev, flags, src)
        out[name + "_src"] = np.uint32(src)
        out[name + "_tiles"] = t
        out[name + "_dist"] = d
    W, H = 300, 260
    elev, flags = synth.random_grid(W, H, seed_k=21, wall_pct=30, elev_mod=512)
    src = int(synth.pick_open_tiles(flags, 1, 5)[0])
    t, d = ref.directions_update(W, H, elev, flags, src)
    out["big_src"] = np.uint32(src)
    out["big_dist_xor"] = np.uint32(np.bitwise_xor.reduce(d))
    out["big_dist_sum"] = np.uint32(d.astype(np.uint64).sum() & 0xFFFFFFFF)
    out["big_tiles_sum"] = np.uint32(t.astype(np.uint64).sum() & 0xFFFFFFFF)
    out["big_dist_row100"] = d[100 * W : 101 * W].copy()
    out["big_tiles_row100"] = t[100 * W : 101 * W].copy()
    return out


def sdf_golden(ref):
    """Maker::sdf over a reduced marching-cubes lattice from the compiled reference pieces (ref_sdf_lattice)"""
    res = (44, 10, 40)
    sdf, flags = ref.sdf_lattice(KAT[0] + 6, KAT[1], res=res)
    return dict(res=np.array(res, np.int32), seed_k=np.int32(6), sdf=sdf, flags=flags)


def game_golden(ref, oracle):
    """the scripted place/remove workload (mazetd_b200/game_events.py) at reduced size: map from the reference, event loop
    over the oracle backend of tests/test_gpu_parity.py (which test_oracle_vs_ref.py pins to the reference): final state"""
    from mazetd_b200 import game_events
    from test_gpu_parity import _OracleGame

    r = ref.mapgen(KAT[0], KAT[1])
    sp = int(r["agg"][2])
    wv = oracle.waypoints_view(101, 101, r["elev"], r["flags"], r["wp_tiles"], r["tiles"], r["dist"], 50, 50)
    find = lambda pos, bits: tuple(np.array(x, np.uint32) for x in zip(*[oracle.find(wv, int(p), int(b)) for p, b in zip(pos, bits)]))
    mon = game_events.initial_monsters(600, r["fullpath"][sp][: int(r["fullpath_len"][sp])], r["wp_tiles"], find)
    og = _OracleGame(oracle, 101, 101, r["elev"], r["flags"], r["wp_tiles"], {k: v.copy() for k, v in mon.items()})
    stats = game_events.run_events(og, 300)
    out = {"mon0_" + k: v for k, v in mon.items()}
    out.update({"mon_" + k: og.a[k] for k in ("tile", "tileStart", "tileEnd", "timeStart", "timeEnd", "bits", "alive")})
    out.update(flags=og.flags, stats=np.array([stats[k] for k in ("placed", "cleared", "rejected_flags", "rejected_trial")], np.int32),
               dist_sum=np.array([int(og.w["dist"][l].astype(np.uint64).sum() & 0xFFFFFFFF) for l in range(4)], np.uint32), agg=og.w["agg"], full_distance=og.w["full_distance"])
    return out


def main():
    oracle_lib.build_ref()
    ref = oracle_lib.Ref()
    np.savez_compressed(os.path.join(HERE, "mapgen_k0_k7.npz"), **mapgen_golden(ref))
    np.savez_compressed(os.path.join(HERE, "directions.npz"), **directions_golden(ref))
    np.savez_compressed(os.path.join(HERE, "sdf_lattice.npz"), **sdf_golden(ref))
    oracle_lib.build_oracle()
    np.savez_compressed(os.path.join(HERE, "game_events.npz"), **game_golden(ref, oracle_lib.Oracle()))
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
