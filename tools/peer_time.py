"""scratch: C4 (16384^2, 16 sources) as row stripes on N ranks, per-rank kernel times.  torchrun --nproc-per-node N tools/peer_time.py [size] [delta,delta..]"""
import json, os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mazetd_b200 import capi, shard, striped, synth
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
size = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
deltas = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [65536]
W = H = size
elev, flags = synth.random_grid_torch(W, H, seed_k=1, wall_pct=30, elev_mod=512, device=dev)
srcs = synth.pick_open_tiles_torch(flags, 16, 1)
y0, y1 = shard.stripe_rows(rank, world, H)
ya, yb = shard.stripe_halo_rows(y0, y1, H)
e = elev.view(H, W)[ya:yb].contiguous(); f = flags.view(H, W)[ya:yb].contiguous()
del elev, flags
st = striped.GpuStripe(local, W, H, y0, y1, e, f, srcs)
striped.connect_peers(st, rank, world)
st.ctx.profile_enable(True)
for kernel in (1, 0):
    st.ctx.debug_set(st.ctx.DEBUG_BIG_ASYNC, kernel)
    for delta in (deltas if kernel else [0]):
        if kernel:
            st.ctx.debug_set(st.ctx.DEBUG_ASYNC_DELTA, delta)
        res = []
        for it in range(3):
            dist.barrier(); torch.cuda.synchronize()
            st.ctx.profile_read(st.ctx.K_BIG_RELAX)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); striped.solve_peer(st); e1.record(); torch.cuda.synchronize()
            relax, _ = st.ctx.profile_read(st.ctx.K_BIG_RELAX)
            res.append((round(e0.elapsed_time(e1), 3), round(relax, 3)))
        bs = st.ctx.big_stats()
        rec = dict(rank=rank, world=world, kernel="dataflow" if kernel else "rounds", delta=delta, whole_ms_relax_ms=res, visits=bs["visits"], polls=bs["polls"], rounds=bs["rounds"], rows=[y0, y1], sources_here=int(sum(1 for s in srcs if y0 <= int(s) // W < y1)))
        out = [None] * world
        dist.all_gather_object(out, rec)
        if rank == 0:
            for r in out:
                print(json.dumps(r), flush=True)
dist.destroy_process_group()
