"""BASELINE.json configs[2] ("C3") over several GPUs, as SURVEY.md 8e lays it out: the 101 x 101 map itself never shards
(10 201 tiles fit one SM: every rank keeps a replica of the grid), the MONSTERS are sharded by index, and after every
accepted tower event the 326 KB of distance / direction layers go from rank 0 to the others by one broadcast
(torch.distributed: NCCL over NVLink on the GPU box, gloo in the CPU tests). The only other exchange is one integer per
trial: placingBuildingBlocksMonsters (sources/player/buildingsPlacement.cpp:34-41) rejects a tile if ANY monster would
be cut off, so the ranks' local verdicts are OR-ed (all-reduce MAX). Damage to the player is summed on the host.

`ShardedGame` wraps any local backend with the four operations of mazetd_b200.game_events (GpuGame on a GPU, the
oracle backend in the tests) plus `export_layers()` / `import_layers(t, d)`; every rank runs the same event script.
"""
import numpy as np

from . import shard


def monster_range(rank, world, n):
    """[first, last) of the monster indices rank owns (contiguous blocks, sizes differ by at most one)"""
    return shard.map_range(rank, world, n)


def shard_monsters(mon, rank, world):
    a, b = monster_range(rank, world, len(mon["tile"]))
    return {k: np.ascontiguousarray(v[a:b]) for k, v in mon.items()}


class ShardedGame:
    """mode "broadcast": rank 0 rebuilds, its layers are broadcast; mode "replicate": every rank rebuilds its own copy
    (no collective: the rebuild costs less than the broadcast's latency on one box, kept for comparison)"""

    def __init__(self, local, rank, world, mode="broadcast", group=None, device=None):
        import torch

        self.local, self.rank, self.world, self.mode, self.group = local, rank, world, mode, group
        self.torch = torch
        self.device = device
        self.flags = local.flags
        self.collectives = 0

    def trial_blocks(self, tile):
        mine = 1 if self.local.trial_blocks(tile) else 0
        if self.world == 1:
            return bool(mine)
        import torch.distributed as dist

        t = self.torch.tensor([mine], dtype=self.torch.int32, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        self.collectives += 1
        return bool(int(t.item()))

    def set_flags(self, tile, value):
        self.local.set_flags(tile, value)

    def rebuild(self):
        if self.world == 1 or self.mode == "replicate":
            self.local.rebuild()
            return
        import torch.distributed as dist

        if self.rank == 0:
            self.local.rebuild()
        tiles, d = self.local.export_layers()  # tensors that ALIAS the backend's layer storage where it can (zero copy on the GPU)
        dist.broadcast(tiles, src=0, group=self.group)
        dist.broadcast(d, src=0, group=self.group)
        self.collectives += 2
        if self.rank != 0:
            self.local.import_layers(tiles, d)

    def tick(self, t):
        self.local.tick(t)


class _DevView:
    """zero-copy torch view of device memory owned by the library (the resident layers)"""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (int(ptr), False), "version": 2}


class GpuShardBackend:
    """game_events.GpuGame plus the two layer hooks, on this rank's monster shard"""

    def __init__(self, ctx, W, H, elev, flags, wp_tiles, device, ox=50, oy=50):
        import torch

        from . import game_events

        self.g = game_events.GpuGame(ctx, W, H, elev, flags, wp_tiles, ox, oy)
        self.ctx, self.flags = ctx, self.g.flags
        n = len(wp_tiles) * W * H
        pt, pd = ctx.layers_dev()
        self.tiles = torch.as_tensor(_DevView(pt, n), device=device)
        self.dist = torch.as_tensor(_DevView(pd, n), device=device)
        self.trial_blocks, self.set_flags, self.rebuild, self.tick = self.g.trial_blocks, self.g.set_flags, self.g.rebuild, self.g.tick

    def export_layers(self):
        self.ctx.synchronize()  # the rebuild ran on the library's stream, the broadcast runs on torch's
        return self.tiles, self.dist

    def import_layers(self, tiles, d):
        import torch

        torch.cuda.current_stream().synchronize()
        self.ctx.layers_refresh()
