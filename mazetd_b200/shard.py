"""Host-side partitioning of the path across the GPUs of one box. Pure functions: no torch, no CUDA.

Only partitions that exist in the domain are offered:
  * independent maps    (C2: batch of generated maps)          -> contiguous block of map indices per rank, no collective
  * independent layers  (C5: one Directions layer per waypoint) -> round-robin layers per rank, inputs replicated
  * row stripes         (C4: one very large grid)               -> contiguous rows per rank + 1 halo row, exchanged per sweep
The 101 x 101 game map itself never shards ("replicas only"): 10 201 tiles fit one SM's shared memory.
"""


def map_range(rank, world, nmaps):
    """[first, last) of the map indices rank owns when nmaps independent maps are split over world ranks;
    the first nmaps % world ranks hold one extra map (sizes differ by at most one, order preserved)."""
    _check(rank, world)
    if nmaps < 0:
        raise ValueError("nmaps must be >= 0")
    base, extra = divmod(nmaps, world)
    first = rank * base + min(rank, extra)
    return first, first + base + (1 if rank < extra else 0)


def layer_assignment(rank, world, nlayers):
    """layer l (one waypoint's Directions, sources/map/paths.cpp:82-85 runs them as independent tasks) -> rank l % world"""
    _check(rank, world)
    return [l for l in range(nlayers) if l % world == rank]


def stripe_rows(rank, world, H, align=64):
    """[y0, y1) of the grid rows rank owns. Cuts are multiples of `align` (the kernels' tile height) whenever H allows it,
    so no tile straddles two ranks; every rank owns at least one row when H >= world."""
    _check(rank, world)
    if H < world:
        raise ValueError(f"cannot split {H} rows over {world} ranks")
    units = H // align
    if units >= world:
        # whole tile rows per rank, remainder rows (H % align) go to the last rank
        a, _ = map_range(rank, world, units)
        _, b = map_range(rank, world, units)
        y0, y1 = a * align, b * align
        if rank == world - 1:
            y1 = H
        return y0, y1
    return map_range(rank, world, H)


def stripe_halo_rows(y0, y1, H):
    """rows [ya, yb) a stripe has to see: its own plus one halo row on each inner side (edge weights across the cut,
    sources/map/grid.cpp:141-145, need the elevation of both end points)"""
    return max(y0 - 1, 0), min(y1 + 1, H)


def _check(rank, world):
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
