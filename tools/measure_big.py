"""scratch: time the tiled large-grid path (C4/C5-like) on the GPU
    python tools/measure_big.py <size> <nsrc> [layers]"""
import sys, time, json
import numpy as np
sys.path.insert(0, '/root/repo')
from mazetd_b200 import capi, synth
import torch
size = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nsrc = int(sys.argv[2]) if len(sys.argv) > 2 else 1
layers = int(sys.argv[3]) if len(sys.argv) > 3 else 0
W = H = size
t0 = time.time()
elev, flags = synth.random_grid(W, H, seed_k=1, wall_pct=30, elev_mod=512)
srcs = synth.pick_open_tiles(flags, max(nsrc, layers), 1)
print("gen", time.time() - t0, flush=True)
ctx = capi.Context(0)
ctx.profile_enable(True)
import os
if os.environ.get('MTD_TILE_DELTA'):
    ctx.debug_set(ctx.DEBUG_TILE_DELTA, int(os.environ['MTD_TILE_DELTA']))
ctx.grid_upload(W, H, elev, flags)
if layers:
    for it in range(3):
        l0 = ctx.launches
        t0 = time.time()
        ctx.waypoints_update_dev(srcs[:layers])
        ctx.synchronize()
        wall = time.time() - t0
        relax, _ = ctx.profile_read(ctx.K_BIG_RELAX)
        dirs, _ = ctx.profile_read(ctx.K_BIG_DIRECTIONS)
        prep, _ = ctx.profile_read(ctx.K_BIG_PREPARE)
        print(json.dumps(dict(size=size, layers=layers, wall_ms=wall * 1e3, relax_ms=relax, dir_ms=dirs, prep_ms=prep, launches=ctx.launches - l0, gbps_alg=W * H * (4 + 8 * layers) / (relax + dirs) / 1e6)), flush=True)
else:
    dt = torch.empty(W * H, dtype=torch.int32, device='cuda')
    dd = torch.empty(W * H, dtype=torch.int32, device='cuda')
    for it in range(3):
        l0 = ctx.launches
        t0 = time.time()
        ctx.directions_update_dev(srcs[:nsrc], dt.data_ptr(), dd.data_ptr())
        ctx.synchronize()
        wall = time.time() - t0
        relax, _ = ctx.profile_read(ctx.K_BIG_RELAX)
        dirs, _ = ctx.profile_read(ctx.K_BIG_DIRECTIONS)
        prep, _ = ctx.profile_read(ctx.K_BIG_PREPARE)
        print(json.dumps(dict(size=size, nsrc=nsrc, wall_ms=wall * 1e3, relax_ms=relax, dir_ms=dirs, prep_ms=prep, launches=ctx.launches - l0, gbps_alg=W * H * 12 / (relax + dirs) / 1e6)), flush=True)
