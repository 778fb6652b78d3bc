"""The scripted tower place/remove workload of BASELINE.json configs[2] ("C3") over any backend that
offers the four map operations of the game loop. The product backend is GpuGame (libmazetd_b200 through the C ABI);
the tests run the same script over an oracle backend and compare every piece of state.

Per event j (RNG = the reference's xoroshiro128**, seed (s1 + 424242, s2)):
    tile = randomRange(0, T)
    Building set            -> clearStructure path  (sources/player/buildingsPlacement.cpp:143-164): clear bit, rebuild
    else canPlaceBuilding() -> placeStructure path  (buildingsPlacement.cpp:46-54, 120-130): flags test, trial
                               (placingBuildingBlocksMonsters, :18-42), set bit, rebuild
    else                    -> no-op
    one "move monsters" tick (sources/game/monsters.cpp:55-99) at gameTime = j + 1
"""
import numpy as np

from . import synth

UNBUILDABLE = synth.FLAG_INVALID | synth.FLAG_WAYPOINT | synth.FLAG_BUILDING | synth.FLAG_TRAP | synth.FLAG_WATER | synth.FLAG_SNOW


def run_events(game, nevents, seed_k=424242, first_time=1):
    """game: .flags (host uint16 mirror), .trial_blocks(tile) -> bool, .set_flags(tile, value), .rebuild(), .tick(t).
    Returns counters {placed, cleared, rejected_flags, rejected_trial}."""
    g = synth.Xoroshiro(synth.KAT_SEED[0] + seed_k, synth.KAT_SEED[1])
    T = game.flags.size
    stats = dict(placed=0, cleared=0, rejected_flags=0, rejected_trial=0)
    for j in range(nevents):
        tile = g.range_u32(0, T)
        f = int(game.flags[tile])
        if f & synth.FLAG_BUILDING:
            game.flags[tile] = f & ~synth.FLAG_BUILDING
            game.set_flags(tile, int(game.flags[tile]))
            game.rebuild()
            stats["cleared"] += 1
        elif f & UNBUILDABLE:
            stats["rejected_flags"] += 1
        elif game.trial_blocks(tile):
            stats["rejected_trial"] += 1
        else:
            game.flags[tile] = f | synth.FLAG_BUILDING
            game.set_flags(tile, int(game.flags[tile]))
            game.rebuild()
            stats["placed"] += 1
        game.tick(first_time + j)
    return stats


def initial_monsters(n, path, wp_tiles, find, seed=7, game_time=0, speed_range=(0.3, 0.6)):
    """n monsters at uniformly random indices along `path` (the min spawner's fullPath) with the steering state
    SpawningGroup::spawnOne gives them (sources/game/monsterSpawning.cpp:205-216), speed uniform in speed_range.
    find(pos[], bits[]) -> (tile[], distance[]) is Waypoints::find evaluated in batch."""
    rng = np.random.default_rng(seed)
    path = np.asarray(path, np.uint32)
    seen = np.zeros(len(path), np.uint32)
    for i, t in enumerate(wp_tiles):
        seen |= np.where(path == t, np.uint32(1 << i), np.uint32(0))
    seen = np.bitwise_or.accumulate(seen)
    idx = rng.integers(0, max(1, len(path) - 1), n)
    tile = path[idx]
    bits = seen[idx]
    speed = rng.uniform(speed_range[0], speed_range[1], n).astype(np.float32)
    _, dist = find(tile, bits)
    tta = np.float32(30.0) * (dist.astype(np.float32) * np.float32(1.0 / 255)) / speed
    a = dict(tile=tile.copy(), tileStart=tile.copy(), tileEnd=tile.copy(), timeStart=np.full(n, game_time, np.uint32), timeEnd=np.full(n, game_time, np.uint32),
             bits=bits.astype(np.uint32), timeToArrive=(np.uint32(game_time) + tta.astype(np.uint32)).astype(np.uint32), dotflags=np.zeros(n, np.uint8),
             damage=np.ones(n, np.uint32), alive=np.ones(n, np.uint8))
    return a


class GpuGame:
    """the four operations on a capi.Context holding the map and the monsters"""

    def __init__(self, ctx, W, H, elev, flags, wp_tiles, ox=50, oy=50):
        self.ctx, self.wp = ctx, np.ascontiguousarray(wp_tiles, np.uint32)
        self.flags = np.array(flags, np.uint16)
        ctx.grid_upload(W, H, elev, self.flags, ox, oy)
        ctx.waypoints_update_dev(self.wp)

    def trial_blocks(self, tile):
        return self.ctx.placing_blocks(int(tile))

    def set_flags(self, tile, value):
        self.ctx.grid_set_flags(int(tile), int(value))

    def rebuild(self):
        self.ctx.waypoints_update_dev(self.wp)

    def tick(self, t):
        self.ctx.monsters_tick_dev(int(t))
