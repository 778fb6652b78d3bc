"""Synthetic fixed-seed inputs for the bench configs and parity tests.

All streams come from the reference PRNG, xoroshiro128** (externals/cage/sources/libcore/math/random.cpp:27-38),
re-stated here in vectorised numpy so CPU and GPU harnesses see identical data. Base seed is the one the
reference's own known-answer test pins (externals/cage/sources/test-core/random.cpp:47); run k uses (s1+k, s2).
"""
import numpy as np

KAT_SEED = (1346519564496, 42245614964156)
M = 0xFFFFFFFF

FLAG_INVALID = 1 << 0
FLAG_BORDER = 1 << 1
FLAG_WAYPOINT = 1 << 2
FLAG_WATER = 1 << 5
FLAG_SUN = 1 << 6
FLAG_WIND = 1 << 7
FLAG_SNOW = 1 << 8
FLAG_MANA = 1 << 13
FLAG_BUILDING = 1 << 14
FLAG_TRAP = 1 << 15

_U64 = np.uint64


def _rotl(x, k):
    return (x << _U64(k)) | (x >> _U64(64 - k))


class Xoroshiro:
    """Scalar generator, bit-identical with cage::RandomGenerator."""

    def __init__(self, s1, s2):
        self.s0 = int(s1) & (2**64 - 1)
        self.s1 = int(s2) & (2**64 - 1)

    def next(self):
        mask = 2**64 - 1
        s0, s1 = self.s0, self.s1
        r = (s0 * 5) & mask
        r = ((r << 7) | (r >> 57)) & mask
        r = (r * 9) & mask
        s1 ^= s0
        self.s0 = (((s0 << 24) | (s0 >> 40)) & mask) ^ s1 ^ ((s1 << 16) & mask)
        self.s1 = ((s1 << 37) | (s1 >> 27)) & mask
        return r

    def range_u32(self, lo, hi):
        if lo == hi:
            return lo
        return self.next() % (hi - lo) + lo

    def chance(self):
        return np.float32((self.next() >> 11) * 2.0**-53)


def stream(s1, s2, n, lanes=4096):
    """n pseudo-random uint64 from `lanes` interleaved xoroshiro128** generators seeded (s1 + lane, s2).

    Element i comes from lane i % lanes, draw i // lanes. (A single sequential generator would need a
    python-level loop of n steps; interleaving keeps it vectorised and still fully determined by the seed.)
    """
    lanes = int(min(lanes, max(1, n)))
    with np.errstate(over="ignore"):
        s0 = np.arange(lanes, dtype=_U64) + _U64(s1)
        s1v = np.full(lanes, s2, dtype=_U64)
        steps = (n + lanes - 1) // lanes
        out = np.empty((steps, lanes), dtype=_U64)
        for t in range(steps):
            out[t] = _rotl(s0 * _U64(5), 7) * _U64(9)
            x = s1v ^ s0
            s0 = _rotl(s0, 24) ^ x ^ (x << _U64(16))
            s1v = _rotl(x, 37)
    return out.reshape(-1)[:n]


def random_grid(W, H, seed_k=0, wall_pct=30, elev_mod=512, lanes=None):
    """C4/C5 distributions: blocked[i] = (r % 100) < wall_pct (flag Invalid), elev[i] = int16(r' % elev_mod) - elev_mod/2."""
    T = W * H
    lanes = lanes or min(1 << 16, T)
    r = stream(KAT_SEED[0] + seed_k, KAT_SEED[1], 2 * T, lanes=lanes)
    blocked = (r[:T] % _U64(100)) < _U64(wall_pct)
    elev = ((r[T:] % _U64(elev_mod)).astype(np.int64) - elev_mod // 2).astype(np.int16)
    flags = np.where(blocked, FLAG_INVALID, FLAG_SUN | FLAG_MANA).astype(np.uint16)
    return elev, flags


def pick_open_tiles(flags, n, seed_k=0):
    """first n distinct unblocked draws of randomRange(0, T) from generator (s1 + 7777 + k, s2)"""
    g = Xoroshiro(KAT_SEED[0] + 7777 + seed_k, KAT_SEED[1])
    T = flags.size
    out = []
    while len(out) < n:
        i = g.range_u32(0, T)
        if (int(flags.reshape(-1)[i]) & (FLAG_INVALID | FLAG_BUILDING)) == 0 and i not in out:
            out.append(i)
    return np.array(out, dtype=np.uint32)


def maze_grid(W, H, seed_k=0, elev_amp=60):
    """A player-style serpentine maze: open field with Building walls every other row leaving one gap,
    used to exercise very long, direction-reversing shortest paths."""
    elev, _ = random_grid(W, H, seed_k, wall_pct=0, elev_mod=max(2, elev_amp))
    flags = np.full((H, W), FLAG_SUN | FLAG_MANA, np.uint16)
    for y in range(1, H, 2):
        flags[y, :] |= FLAG_BUILDING
        gap = (W - 1) if ((y // 2) % 2 == 0) else 0
        flags[y, gap] &= ~np.uint16(FLAG_BUILDING)
    return elev, flags.reshape(-1)


def random_grid_torch(W, H, seed_k=0, wall_pct=30, elev_mod=512, lanes=None, device="cuda"):
    """`random_grid` computed with torch integer ops on `device` (bit-identical output, as int16 / int16-viewed-uint16
    tensors): the 16384^2 grid of C4 takes 20 s in numpy on the host and well under a second on the GPU. Input
    generation only -- the hot path never sees this function."""
    import torch

    T = W * H
    lanes = int(lanes or min(1 << 16, T))
    lanes = int(min(lanes, max(1, 2 * T)))
    steps = (2 * T + lanes - 1) // lanes
    i64 = torch.int64

    def c(v):  # python int -> wrapped int64 constant
        v &= (1 << 64) - 1
        return v - (1 << 64) if v >= (1 << 63) else v

    def lrs(x, k):
        return (x >> k) & ((1 << (64 - k)) - 1)

    def rotl(x, k):
        return (x << k) | lrs(x, 64 - k)

    def umod(x, m):  # unsigned 64-bit x mod small m
        return ((lrs(x, 1) % m) * 2 + (x & 1)) % m

    s0 = torch.arange(lanes, dtype=i64, device=device) + c(KAT_SEED[0] + seed_k)
    s1 = torch.full((lanes,), c(KAT_SEED[1]), dtype=i64, device=device)
    blocked = torch.empty(T, dtype=torch.bool, device=device)
    elev = torch.empty(T, dtype=torch.int16, device=device)
    for t in range(steps):
        r = rotl(s0 * 5, 7) * 9
        x = s1 ^ s0
        s0 = rotl(s0, 24) ^ x ^ (x << 16)
        s1 = rotl(x, 37)
        a, b = t * lanes, min((t + 1) * lanes, 2 * T)
        if a < T:  # first T draws: walls
            n = min(b, T) - a
            blocked[a:a + n] = umod(r[:n], 100) < wall_pct
        if b > T:  # second T draws: elevations
            lo = max(a, T)
            elev[lo - T:b - T] = (umod(r[lo - a:b - a], elev_mod) - elev_mod // 2).to(torch.int16)
    flags = torch.where(blocked, torch.tensor(FLAG_INVALID, dtype=torch.int16, device=device), torch.tensor(FLAG_SUN | FLAG_MANA, dtype=torch.int16, device=device))
    return elev, flags


def pick_open_tiles_torch(flags_t, n, seed_k=0):
    """`pick_open_tiles` on a device tensor of flags: only the drawn tiles are read back"""
    g = Xoroshiro(KAT_SEED[0] + 7777 + seed_k, KAT_SEED[1])
    T = flags_t.numel()
    out = []
    while len(out) < n:
        i = g.range_u32(0, T)
        if (int(flags_t.view(-1)[i].item()) & (FLAG_INVALID | FLAG_BUILDING)) == 0 and i not in out:
            out.append(i)
    return np.array(out, dtype=np.uint32)
