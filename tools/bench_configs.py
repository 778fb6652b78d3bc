"""Timing of the two large-grid configurations of BASELINE.json next to bench.py's headline (C2):
    C4  one 16384^2 grid, 16 sources, one layer, row-striped over the ranks (halo rows swapped per sweep)
    C4P the same with peer-coupled stripes: one kernel run per rank, rows and keys written into the neighbours' memory
    C5  one 4096^2 grid, 8 waypoint layers, layers dealt round-robin to the ranks (no collective)
Run:  python tools/bench_configs.py C4 [size] [reps]                       (1 GPU)
      python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
             tools/bench_configs.py C4 [size] [reps]                      (N GPUs, one rank per GPU, NCCL)
Rank 0 prints one JSON line per repetition. Device time between two CUDA events on the stream every kernel is launched
on, max over ranks. The checksums are identical for every N (the result does not depend on the partitioning)."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mazetd_b200 import capi, shard, striped, synth  # noqa: E402


def checksum(t, world):
    """(sum of finite labels, number of finite labels) over all ranks; t = int32 tensor of uint32 bit patterns"""
    v = t.reshape(-1).to(torch.int64) & 0xFFFFFFFF
    fin = v != 0xFFFFFFFF
    s = torch.stack([(v * fin).sum(), fin.sum()])
    if world > 1:
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    return [int(s[0].item()), int(s[1].item())]


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
    size = int(sys.argv[2]) if len(sys.argv) > 2 else (16384 if cfg in ("C4", "C4P") else 4096)
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    budget = int(sys.argv[4]) if len(sys.argv) > 4 else 16  # C4: relaxation rounds between halo swaps (0 = local convergence)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    W = H = size
    elev, flags = synth.random_grid(W, H, seed_k=1, wall_pct=30, elev_mod=512)
    nsrc = 16 if cfg in ("C4", "C4P") else 8
    srcs = synth.pick_open_tiles(flags, nsrc, 1)

    def timed(fn):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        out = fn()
        e1.record(stream)
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), out

    if cfg in ("C4", "C4P"):
        peer = cfg == "C4P" and world > 1
        st = striped.make_gpu_stripe(local, rank, world, W, H, elev, flags, srcs, stream=stream.cuda_stream, round_budget=budget if world > 1 and not peer else 0)
        del elev, flags
        if os.environ.get("MTD_TILE_DELTA"):  # priority window of the tile scheduler (experiments)
            st.ctx.debug_set(st.ctx.DEBUG_TILE_DELTA, int(os.environ["MTD_TILE_DELTA"]))
        if peer:
            striped.connect_peers(st, rank, world)
        for rep in range(reps):
            l0 = st.ctx.launches
            ms, (tiles, d, sweeps) = timed((lambda: striped.solve_peer(st)) if peer else (lambda: striped.solve_distributed(st, rank, world)))
            line = dict(config=cfg, grid=[W, H], sources=nsrc, n_gpus=world, rep=rep, ms=ms, sweeps=sweeps, round_budget=0 if peer else budget,
                        tile_relaxations_per_s=W * H / (ms * 1e-3), algorithmic_GBps=W * H * 12 / (ms * 1e-3) / 1e9,
                        launches_rank0=st.ctx.launches - l0, dist_checksum=checksum(d, world), tiles_checksum=checksum(tiles, world))
            if rank == 0:
                print(json.dumps(line), flush=True)
    else:
        mine = shard.layer_assignment(rank, world, nsrc)
        ctx = capi.Context(local, stream=stream.cuda_stream)
        ctx.grid_upload(W, H, elev, flags)
        del elev, flags
        n = len(mine)
        dt = torch.empty((max(n, 1), W * H), dtype=torch.int32, device=dev)
        dd = torch.empty((max(n, 1), W * H), dtype=torch.int32, device=dev)

        def run():
            # one Directions::update per owned layer (independent tasks in the reference: paths.cpp:82-85, 119)
            ctx.layers_update_dev(srcs[mine], dt.data_ptr(), dd.data_ptr()) if n else None

        for rep in range(reps):
            l0 = ctx.launches
            ms, _ = timed(run)
            line = dict(config="C5", grid=[W, H], layers=nsrc, n_gpus=world, rep=rep, ms=ms,
                        tile_relaxations_per_s=W * H * nsrc / (ms * 1e-3), algorithmic_GBps=W * H * (4 * world + 8 * nsrc) / (ms * 1e-3) / 1e9,
                        launches_rank0=ctx.launches - l0, dist_checksum=checksum(dd[:n], world), tiles_checksum=checksum(dt[:n], world))
            if rank == 0:
                print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
