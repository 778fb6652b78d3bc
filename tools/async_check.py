"""scratch: dataflow scheduler (k_tile_async) against the oracle and the round-based kernel, then timings
    python tools/async_check.py [quick]"""
import sys, time, json, os
import numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import oracle_lib
from mazetd_b200 import capi, synth
import torch

orc = oracle_lib.Oracle()
ctx = capi.Context(0)
ok = True
for (W, H, nsrc, wall) in [(96, 70, 1, 30), (300, 260, 3, 30), (700, 500, 5, 30), (1024, 1024, 2, 35), (333, 777, 4, 0)]:
    elev, flags = synth.random_grid(W, H, seed_k=W, wall_pct=wall, elev_mod=512)
    srcs = synth.pick_open_tiles(flags, nsrc, 2)
    to, do = orc.directions_update(W, H, elev, flags, srcs)
    ctx.grid_upload(W, H, elev, flags)
    for mode in (3, 1, 0):
        ctx.debug_set(ctx.DEBUG_BIG_ASYNC, 1 if mode else 0)
        ctx.debug_set(ctx.DEBUG_BIG_TMA, 1 if mode == 3 else 0)
        t, d = ctx.directions_update(srcs)
        st = ctx.big_stats()
        e = np.array_equal(d, do) and np.array_equal(t, to)
        ok &= e
        print(W, H, nsrc, {3: "async+tma", 1: "async", 0: "rounds"}[mode], "OK" if e else "MISMATCH %d" % int((d != do).sum()), st, flush=True)
elev, flags = synth.maze_grid(200, 160, 3)
srcs = synth.pick_open_tiles(flags, 1, 2)
to, do = orc.directions_update(200, 160, elev, flags, srcs)
ctx.grid_upload(200, 160, elev, flags)
ctx.debug_set(ctx.DEBUG_BIG_ASYNC, 1)
ctx.debug_set(ctx.DEBUG_BIG_TMA, 1)
t, d = ctx.directions_update(srcs)
print("maze", "OK" if np.array_equal(d, do) and np.array_equal(t, to) else "MISMATCH", ctx.big_stats(), flush=True)
print("PARITY", "GREEN" if ok else "RED", flush=True)

def timing(size, nsrc, layers, modes):
    W = H = size
    elev, flags = synth.random_grid(W, H, seed_k=1, wall_pct=30, elev_mod=512)
    srcs = synth.pick_open_tiles(flags, max(nsrc, layers), 1)
    ctx.profile_enable(True)
    ctx.grid_upload(W, H, elev, flags)
    n = max(layers, 1)
    dt = torch.empty((n, W * H), dtype=torch.int32, device='cuda')
    dd = torch.empty((n, W * H), dtype=torch.int32, device='cuda')
    ref = None
    for (mode, delta) in modes:
        ctx.debug_set(ctx.DEBUG_BIG_ASYNC, 1 if mode else 0)
        ctx.debug_set(ctx.DEBUG_BIG_TMA, 1 if mode == 3 else 0)
        if mode:
            ctx.debug_set(ctx.DEBUG_ASYNC_DELTA, delta)
        else:
            ctx.debug_set(ctx.DEBUG_TILE_DELTA, delta)
        for it in range(3):
            if layers:
                ctx.layers_update_dev(srcs[:layers], dt.data_ptr(), dd.data_ptr())
            else:
                ctx.directions_update_dev(srcs[:nsrc], dt.data_ptr(), dd.data_ptr())
            ctx.synchronize()
            relax, _ = ctx.profile_read(ctx.K_BIG_RELAX)
            dirs, _ = ctx.profile_read(ctx.K_BIG_DIRECTIONS)
            prep, _ = ctx.profile_read(ctx.K_BIG_PREPARE)
        cs = int((dd.to(torch.int64) & 0xFFFFFFFF).sum().item()), int((dt.to(torch.int64) & 0xFFFFFFFF).sum().item())
        if ref is None:
            ref = cs
        print(json.dumps(dict(size=size, nsrc=nsrc, layers=layers, mode={3: "async+tma", 1: "async", 0: "rounds"}[mode], delta=delta, relax_ms=round(relax, 3), dir_ms=round(dirs, 3), prep_ms=round(prep, 3), same=cs == ref, **ctx.big_stats())), flush=True)

quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
timing(4096, 1, 8, [(0, 98304), (1, 32768), (3, 32768), (3, 16384), (3, 65536), (3, 8192)])
if not quick:
    timing(16384, 16, 0, [(0, 98304), (1, 32768), (3, 32768), (3, 16384), (3, 65536), (3, 131072)])
