"""scratch: one peer-coupled case side by side on one GPU, mismatch map against the oracle"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from mazetd_b200 import capi, shard, striped, synth
import oracle_lib
oracle = oracle_lib.Oracle()
W, H, cuts, nsrc, pct = 384, 320, [0, 128, 320], 1, 30
if len(sys.argv) > 1:
    W, H, nsrc, pct = [int(x) for x in sys.argv[1:5]]; cuts = [int(x) for x in sys.argv[5].split(",")]
elev, flags = synth.random_grid(W, H, seed_k=W + len(cuts), wall_pct=pct)
srcs = synth.pick_open_tiles(flags, nsrc, 2)
print("sources", [(int(s) % W, int(s) // W) for s in srcs])
to, do = oracle.directions_update(W, H, elev, flags, srcs)
# warm both kernels alone
w0 = striped.GpuStripe(0, 64, 64, 0, 64, *synth.random_grid(64, 64, seed_k=1, wall_pct=10), synth.pick_open_tiles(synth.random_grid(64, 64, seed_k=1, wall_pct=10)[1], 1, 1))
striped.solve_peer_local([w0]); w0.close()
stripes = []
for y0, y1 in zip(cuts[:-1], cuts[1:]):
    ya, yb = shard.stripe_halo_rows(y0, y1, H)
    stripes.append(striped.GpuStripe(0, W, H, y0, y1, elev.reshape(H, W)[ya:yb], flags.reshape(H, W)[ya:yb], srcs))
for rep in range(3):
    results, rounds = striped.solve_peer_local(stripes, ctas=64)
    d = np.concatenate([x.cpu().numpy().view(np.uint32).reshape(-1) for _, x in results])
    t = np.concatenate([x.cpu().numpy().view(np.uint32).reshape(-1) for x, _ in results])
    badt = np.nonzero(t != to)[0]
    if len(badt):
        print("rep", rep, "NEXT-HOP mismatches", len(badt), "rows", sorted(set((badt // W).tolist()))[:20])
    bad = np.nonzero(d != do)[0]
    print("rep", rep, "mismatches", len(bad), "next-hop mismatches", len(badt), "visits", [s.ctx.big_stats() for s in stripes])
    if len(bad):
        ys, xs = bad // W, bad % W
        print("  y range", ys.min(), ys.max(), "x range", xs.min(), xs.max(), "got INF:", int((d[bad] == 0xFFFFFFFF).sum()), "got smaller:", int((d[bad] < do[bad]).sum()), "got larger finite:", int(((d[bad] > do[bad]) & (d[bad] != 0xFFFFFFFF)).sum()))
        tiles = sorted(set(zip((xs // 32).tolist(), (ys // 32).tolist())))
        print("  tiles (tx, ty global):", tiles[:40])
