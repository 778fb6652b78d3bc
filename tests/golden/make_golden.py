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


def main():
    oracle_lib.build_ref()
    ref = oracle_lib.Ref()
    np.savez_compressed(os.path.join(HERE, "mapgen_k0_k7.npz"), **mapgen_golden(ref))
    np.savez_compressed(os.path.join(HERE, "directions.npz"), **directions_golden(ref))
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
