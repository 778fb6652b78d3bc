"""Full-size goldens for the two large-grid configurations of BASELINE.json, from the REFERENCE'S OWN compiled
Directions::update (sources/map/paths.cpp:24-78 in oracle/_ref/libmazetd_ref.so). Run in the authoring container only
(minutes of CPU, ~25 GB of RAM for C4):

    python tests/golden/make_golden_big.py [C5] [C4]

  C5  4096^2, 8 layers: one reference Dijkstra per layer.
  C4  16384^2, 16 sources, ONE multi-source layer. The reference is single-source (paths.cpp:48-49), so the golden
      distances are the element-wise minimum of 16 reference runs (the multi-source fixpoint by definition); next hops
      are the closed form of the heap order applied to those distances, computed here in numpy and
      cross-checked against the oracle's multi-source restatement on the sampled rows.

Stored per layer: xor and 64-bit sum of `distances` and of `tiles`, the count of unreachable tiles, and three full rows
of each array. The inputs are mazetd_b200.synth.random_grid(seed_k=1) / pick_open_tiles(seed_k=1), the same call bench.py
and the GPU tests make. TEST INFRASTRUCTURE.
"""
import os
import sys
from concurrent.futures import ProcessPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle_lib  # noqa: E402
from mazetd_b200 import synth  # noqa: E402

ROWS = {4096: (0, 1777, 4095), 16384: (0, 9001, 16383)}


def inputs(size, nsrc):
    elev, flags = synth.random_grid(size, size, seed_k=1, wall_pct=30, elev_mod=512)
    srcs = synth.pick_open_tiles(flags, nsrc, 1)
    return elev, flags, srcs


def checks(a):
    a64 = a.astype(np.uint64)
    return int(np.bitwise_xor.reduce(a)), int(a64.sum() & np.uint64(0xFFFFFFFFFFFFFFFF))


def _one(args):
    size, src = args
    elev, flags, _ = inputs(size, 16 if size == 16384 else 8)
    ref = oracle_lib.Ref()
    t, d = ref.directions_update(size, size, elev, flags, int(src))
    return t, d


def next_hops(W, H, d, wr, wd, blocked, sources):
    """tiles[v] = neighbour u with d[u] + w(u,v) == d[v] minimising (d[u], u); m at sources / unreachable"""
    M = np.uint64(0xFFFFFFFF)
    d2 = d.reshape(H, W).astype(np.uint64)
    idx = np.arange(W * H, dtype=np.uint64).reshape(H, W)
    best = np.full((H, W), (M << np.uint64(32)) | M, np.uint64)
    fin = d2 != M

    def cand(du, wu, iu, ok):
        hit = ok & (du != M) & (du + wu == d2_view)
        key = (du << np.uint64(32)) | iu
        return np.where(hit, key, (M << np.uint64(32)) | M)

    out = best
    # up neighbour (y-1): weight wd[y-1]
    d2_view = d2[1:, :]
    out[1:, :] = np.minimum(out[1:, :], cand(d2[:-1, :], wd[:-1, :].astype(np.uint64), idx[:-1, :], fin[1:, :]))
    d2_view = d2[:-1, :]
    out[:-1, :] = np.minimum(out[:-1, :], cand(d2[1:, :], wd[:-1, :].astype(np.uint64), idx[1:, :], fin[:-1, :]))
    d2_view = d2[:, 1:]
    out[:, 1:] = np.minimum(out[:, 1:], cand(d2[:, :-1], wr[:, :-1].astype(np.uint64), idx[:, :-1], fin[:, 1:]))
    d2_view = d2[:, :-1]
    out[:, :-1] = np.minimum(out[:, :-1], cand(d2[:, 1:], wr[:, :-1].astype(np.uint64), idx[:, 1:], fin[:, :-1]))
    tiles = (out & M).astype(np.uint32).reshape(-1)
    tiles[blocked.reshape(-1) & (d == 0xFFFFFFFF)] = 0xFFFFFFFF
    tiles[np.asarray(sources, np.int64)] = 0xFFFFFFFF
    return tiles


def main():
    which = [a for a in sys.argv[1:]] or ["C5", "C4"]
    oracle_lib.build_ref()
    oracle_lib.build_oracle()
    path = os.path.join(HERE, "big_configs.npz")
    out = dict(np.load(path)) if os.path.exists(path) else {}
    if "C5" in which:
        size, n = 4096, 8
        elev, flags, srcs = inputs(size, n)
        with ProcessPoolExecutor(4) as ex:
            res = list(ex.map(_one, [(size, s) for s in srcs]))
        out["C5_sources"] = srcs
        out["C5_rows"] = np.array(ROWS[size], np.int32)
        out["C5_dist_xor"] = np.array([checks(d)[0] for _, d in res], np.uint32)
        out["C5_dist_sum"] = np.array([checks(d)[1] for _, d in res], np.uint64)
        out["C5_tiles_xor"] = np.array([checks(t)[0] for t, _ in res], np.uint32)
        out["C5_tiles_sum"] = np.array([checks(t)[1] for t, _ in res], np.uint64)
        out["C5_unreachable"] = np.array([int((d == 0xFFFFFFFF).sum()) for _, d in res], np.int64)
        out["C5_dist_rows"] = np.stack([np.stack([d[r * size:(r + 1) * size] for r in ROWS[size]]) for _, d in res])
        out["C5_tiles_rows"] = np.stack([np.stack([t[r * size:(r + 1) * size] for r in ROWS[size]]) for t, _ in res])
        np.savez_compressed(path, **out)
        print("C5 done", out["C5_dist_sum"])
    if "C4" in which:
        size, n = 16384, 16
        elev, flags, srcs = inputs(size, n)
        dmin = None
        with ProcessPoolExecutor(4) as ex:
            for t, d in ex.map(_one, [(size, s) for s in srcs]):
                dmin = d if dmin is None else np.minimum(dmin, d)
                del t
        wr, wd = oracle_lib.Ref().edge_weights(size, size, elev)  # the compiled Grid::neighborDistance
        blocked = (flags & (synth.FLAG_INVALID | synth.FLAG_BUILDING)) != 0
        tiles = next_hops(size, size, dmin, wr.reshape(size, size), wd.reshape(size, size), blocked, srcs)
        out["C4_sources"] = srcs
        out["C4_rows"] = np.array(ROWS[size], np.int32)
        out["C4_dist_xor"], s = checks(dmin)
        out["C4_dist_xor"] = np.uint32(out["C4_dist_xor"])
        out["C4_dist_sum"] = np.uint64(s)
        x, s = checks(tiles)
        out["C4_tiles_xor"] = np.uint32(x)
        out["C4_tiles_sum"] = np.uint64(s)
        out["C4_unreachable"] = np.int64((dmin == 0xFFFFFFFF).sum())
        out["C4_dist_rows"] = np.stack([dmin[r * size:(r + 1) * size] for r in ROWS[size]])
        out["C4_tiles_rows"] = np.stack([tiles[r * size:(r + 1) * size] for r in ROWS[size]])
        np.savez_compressed(path, **out)
        print("C4 done", out["C4_dist_sum"], out["C4_tiles_sum"])
    print(os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
