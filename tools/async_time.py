"""scratch: time the dataflow scheduler on C5 / C4   python tools/async_time.py C5|C4 delta[,delta...] [tma]"""
import sys, json
sys.path.insert(0, '/root/repo')
from mazetd_b200 import capi, synth
import torch
cfg = sys.argv[1]
deltas = [int(x) for x in sys.argv[2].split(",")]
tma = int(sys.argv[3]) if len(sys.argv) > 3 else 1
size, nsrc, layers = (4096, 1, 8) if cfg == "C5" else (16384, 16, 0)
W = H = size
dev = torch.device("cuda", 0)
elev, flags = synth.random_grid_torch(W, H, seed_k=1, wall_pct=30, elev_mod=512, device=dev)
srcs = synth.pick_open_tiles_torch(flags, max(nsrc, layers), 1)
ctx = capi.Context(0)
ctx.profile_enable(True)
ctx.grid_upload_dev(W, H, elev.data_ptr(), flags.data_ptr())
n = max(layers, 1)
dt = torch.empty((n, W * H), dtype=torch.int32, device=dev)
dd = torch.empty((n, W * H), dtype=torch.int32, device=dev)
ctx.debug_set(ctx.DEBUG_BIG_TMA, tma)
import os
nb = int(os.environ.get('MTD_NB', '-1'))
depth = int(os.environ.get('MTD_DEPTH', '2'))
ctx.debug_set(ctx.DEBUG_ASYNC_NB_RULE, nb)
ctx.debug_set(ctx.DEBUG_ASYNC_DEPTH, depth)
ctx.debug_set(ctx.DEBUG_ASYNC_WORKERS, int(os.environ.get('MTD_WORKERS', '16')))
looks = int(os.environ.get('MTD_LOOKS', '1'))
contv = int(os.environ.get('MTD_CONT', '0'))
try:
    ctx.debug_set(ctx.DEBUG_ASYNC_CONT, contv)
except Exception:
    contv = -1
try:
    ctx.debug_set(ctx.DEBUG_ASYNC_LATE_LOOKS, looks)
except Exception:
    looks = -1  # a library built before the knob existed
for delta in deltas:
    ctx.debug_set(ctx.DEBUG_ASYNC_DELTA, delta)
    for it in range(2):
        if layers:
            ctx.layers_update_dev(srcs[:layers], dt.data_ptr(), dd.data_ptr())
        else:
            ctx.directions_update_dev(srcs[:nsrc], dt.data_ptr(), dd.data_ptr())
        ctx.synchronize()
        relax, _ = ctx.profile_read(ctx.K_BIG_RELAX)
    print(json.dumps(dict(cfg=cfg, delta=delta, nb=nb, depth=depth, looks=looks, cont=contv, workers=int(os.environ.get('MTD_WORKERS', '16')), relax_ms=round(relax, 3), **ctx.big_stats())), flush=True)
