"""Worker of test_peer_coupled_stripes_on_two_gpus (run under torch.distributed.run, one rank per GPU): every rank solves
its row stripe of a few grids with the peer-coupled protocol (mazetd_b200/striped.py: connect_peers + solve_peer) and
compares its rows with the oracle's Directions::update of the whole grid. Prints "PEER-OK <cases>" on rank 0."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from mazetd_b200 import shard, striped, synth  # noqa: E402
import oracle_lib  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    oracle = oracle_lib.Oracle()
    cases = [
        # W, H, nsrc, wall %, maze, align (stripe cut alignment)
        (384, 320, 1, 30, False, 64),
        (200, 301, 3, 35, False, 1),     # ragged width, cuts inside tiles
        (1024, 1024, 16, 30, False, 64),
        (320, 256, 1, 0, True, 32),      # a maze wanders across the cuts
        (2048, 2048, 4, 30, False, 64),
    ]
    bad = 0
    for ci, (W, H, nsrc, pct, maze, align) in enumerate(cases):
        elev, flags = synth.maze_grid(W, H, seed_k=4) if maze else synth.random_grid(W, H, seed_k=W + ci, wall_pct=pct)
        srcs = synth.pick_open_tiles(flags, nsrc, 2)
        y0, y1 = shard.stripe_rows(rank, world, H, align=align)
        ya, yb = shard.stripe_halo_rows(y0, y1, H)
        st = striped.GpuStripe(local, W, H, y0, y1, elev.reshape(H, W)[ya:yb], flags.reshape(H, W)[ya:yb], srcs)
        striped.connect_peers(st, rank, world)
        for rep in range(4):  # reps 0, 1: dataflow kernel (the default), 2, 3: round-based kernel; later runs reuse the peer blocks
            st.ctx.debug_set(st.ctx.DEBUG_BIG_ASYNC, 1 if rep < 2 else 0)
            tiles, d, rounds = striped.solve_peer(st)
            if st.ctx.big_stats()["dataflow"] != (rep < 2):
                print(f"rank {rank} case {ci} rep {rep}: wrong kernel ran", flush=True)
                bad += 1
            if rep == 0:
                to, do = oracle.directions_update(W, H, elev, flags, srcs)
            mine_d = d.cpu().numpy().view(np.uint32).reshape(-1)
            mine_t = tiles.cpu().numpy().view(np.uint32).reshape(-1)
            ok = np.array_equal(mine_d, do[y0 * W:y1 * W]) and np.array_equal(mine_t, to[y0 * W:y1 * W])
            if not ok:
                nd = int((mine_d != do[y0 * W:y1 * W]).sum())
                nt = int((mine_t != to[y0 * W:y1 * W]).sum())
                print(f"rank {rank} case {ci} rep {rep}: MISMATCH dist {nd} tiles {nt} rounds {rounds}", flush=True)
                bad += 1
        dist.barrier()
        st.close()
    t = torch.tensor([bad], device="cuda")
    dist.all_reduce(t)
    if rank == 0:
        print(("PEER-OK %d" % len(cases)) if int(t.item()) == 0 else "PEER-FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(1 if int(t.item()) else 0)


if __name__ == "__main__":
    main()
