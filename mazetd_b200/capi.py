"""ctypes binding of libmazetd_b200.so (include/mazetd_b200.h). No torch here: plain pointers and sizes.

The library is the product; this module only loads it and fails loudly when it is missing or when the
CUDA device is absent (there is no CPU fallback by design)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MTD_LIB") or os.path.join(HERE, "libmazetd_b200.so")  # MTD_LIB: developer probe builds only

M = 0xFFFFFFFF

EXPORTS = [
    "mtd_create", "mtd_destroy", "mtd_last_error", "mtd_version", "mtd_synchronize", "mtd_launch_count",
    "mtd_profile_enable", "mtd_profile_read", "mtd_debug_set", "mtd_big_stats",
    "mtd_grid_upload", "mtd_grid_upload_dev", "mtd_grid_set_flags", "mtd_neighbor_distances", "mtd_edge_weights",
    "mtd_directions_update", "mtd_directions_update_dev", "mtd_directions_layers_dev",
    "mtd_stripe_upload", "mtd_stripe_upload_dev", "mtd_stripe_begin_dev", "mtd_stripe_exchange_dev", "mtd_stripe_finalize_dev", "mtd_stripe_round_budget", "mtd_stripe_pending",
    "mtd_stripe_peer_export", "mtd_stripe_peer_connect", "mtd_stripe_peer_connect_local", "mtd_stripe_peer_launch_dev", "mtd_stripe_peer_finish",
    "mtd_waypoints_update", "mtd_waypoints_update_dev", "mtd_waypoints_find", "mtd_waypoints_layers_dev",
    "mtd_placing_blocks",
    "mtd_mapgen_batch", "mtd_mapgen_batch_dev", "mtd_grid_from_elevations", "mtd_elevation_samples", "mtd_sdf_lattice",
    "mtd_monsters_upload", "mtd_monsters_download", "mtd_monsters_tick_dev", "mtd_monsters_tick",
    "mtd_monsters_positions_dev", "mtd_monsters_positions", "mtd_monster_spawn",
    "mtd_monsters_dots_upload", "mtd_monsters_dots_tick_dev", "mtd_monsters_dots_tick", "mtd_monsters_dots_download",
    "mtd_monsters_class_upload", "mtd_towers_upload", "mtd_towers_attack_dev", "mtd_towers_download", "mtd_waypoints_layers_refresh",
    "mtd_waypoints_paths_noalloc",
]


class MtdError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"libmazetd_b200 error {code}: {text}")
        self.code = code


class MapgenParams(C.Structure):
    _fields_ = [("W", C.c_int32), ("H", C.c_int32), ("ox", C.c_int32), ("oy", C.c_int32), ("radius", C.c_int32), ("nwp", C.c_int32), ("fullpath_cap", C.c_uint32)]


_lib = None


def load():
    """dlopen the in-tree library (built by mazetd_b200/build.py); raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} is missing: run `python -m mazetd_b200.build` (needs nvcc). There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, u32, u64, f32 = C.c_void_p, C.c_int32, C.c_uint32, C.c_uint64, C.c_float
    L.mtd_create.argtypes = [C.POINTER(vp), C.c_int, vp]
    L.mtd_destroy.argtypes = [vp]
    L.mtd_destroy.restype = None
    L.mtd_last_error.argtypes = [vp]
    L.mtd_last_error.restype = C.c_char_p
    L.mtd_synchronize.argtypes = [vp]
    L.mtd_launch_count.argtypes = [vp]
    L.mtd_launch_count.restype = u64
    L.mtd_profile_enable.argtypes = [vp, C.c_int]
    L.mtd_profile_read.argtypes = [vp, C.c_int, C.POINTER(C.c_double), C.POINTER(u64)]
    L.mtd_debug_set.argtypes = [vp, C.c_int, C.c_int64]
    L.mtd_big_stats.argtypes = [vp, C.c_void_p]
    L.mtd_grid_upload.argtypes = [vp, i32, i32, i32, i32, vp, vp]
    L.mtd_grid_upload_dev.argtypes = [vp, i32, i32, i32, i32, vp, vp]
    L.mtd_grid_set_flags.argtypes = [vp, u32, C.c_uint16]
    L.mtd_neighbor_distances.argtypes = [vp, vp, vp, i32, vp]
    L.mtd_edge_weights.argtypes = [vp, vp, vp]
    L.mtd_directions_update.argtypes = [vp, vp, i32, u32, vp, vp]
    L.mtd_directions_update_dev.argtypes = [vp, vp, i32, u32, vp, vp]
    L.mtd_directions_layers_dev.argtypes = [vp, vp, i32, vp, vp]
    L.mtd_stripe_upload.argtypes = [vp, i32, i32, i32, i32, vp, vp]
    L.mtd_stripe_upload_dev.argtypes = [vp, i32, i32, i32, i32, vp, vp]
    L.mtd_stripe_begin_dev.argtypes = [vp, vp, i32, vp]
    L.mtd_stripe_exchange_dev.argtypes = [vp, vp, vp, C.POINTER(u32)]
    L.mtd_stripe_finalize_dev.argtypes = [vp, vp]
    L.mtd_stripe_round_budget.argtypes = [vp, i32]
    L.mtd_stripe_pending.argtypes = [vp, C.POINTER(u32)]
    L.mtd_stripe_peer_export.argtypes = [vp, vp]
    L.mtd_stripe_peer_connect.argtypes = [vp, i32, i32, vp]
    L.mtd_stripe_peer_connect_local.argtypes = [vp, i32, i32, C.POINTER(vp), i32]
    L.mtd_stripe_peer_launch_dev.argtypes = [vp, vp, i32, vp, vp]
    L.mtd_stripe_peer_finish.argtypes = [vp, C.POINTER(u32)]
    L.mtd_waypoints_update.argtypes = [vp, vp, i32, vp, vp, vp, u32, vp, vp, vp]
    L.mtd_waypoints_update_dev.argtypes = [vp, vp, i32]
    L.mtd_waypoints_find.argtypes = [vp, vp, vp, i32, vp, vp]
    L.mtd_waypoints_layers_dev.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    L.mtd_placing_blocks.argtypes = [vp, u32, C.POINTER(i32)]
    L.mtd_mapgen_batch.argtypes = [vp, C.POINTER(MapgenParams), u64, u64, i32, i32] + [vp] * 9
    L.mtd_mapgen_batch_dev.argtypes = [vp, C.POINTER(MapgenParams), u64, u64, i32, i32] + [vp] * 9
    L.mtd_grid_from_elevations.argtypes = [vp, C.POINTER(MapgenParams), i32, vp, vp]
    L.mtd_waypoints_paths_noalloc.argtypes = [vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp]
    L.mtd_sdf_lattice.argtypes = [vp, u64, u64, i32, i32, i32, i32, vp, vp, vp, vp, vp]
    L.mtd_elevation_samples.argtypes = [vp, u64, u64, vp, i32, vp]
    L.mtd_monsters_upload.argtypes = [vp, i32] + [vp] * 10
    L.mtd_monsters_download.argtypes = [vp] + [vp] * 7
    L.mtd_monsters_tick_dev.argtypes = [vp, u32]
    L.mtd_monsters_tick.argtypes = [vp, u32, C.POINTER(C.c_int64), C.POINTER(u32)]
    L.mtd_monsters_positions_dev.argtypes = [vp, u32, vp]
    L.mtd_monsters_positions.argtypes = [vp, u32, vp]
    L.mtd_monster_spawn.argtypes = [vp, u32, f32, vp]
    L.mtd_monsters_dots_upload.argtypes = [vp, vp, vp, vp, vp, vp]
    L.mtd_monsters_dots_tick_dev.argtypes = [vp]
    L.mtd_monsters_dots_tick.argtypes = [vp, C.POINTER(C.c_int64), C.POINTER(u32)]
    L.mtd_monsters_dots_download.argtypes = [vp, vp, vp, vp, vp]
    L.mtd_monsters_class_upload.argtypes = [vp, vp]
    L.mtd_towers_upload.argtypes = [vp, i32] + [vp] * 13
    L.mtd_towers_attack_dev.argtypes = [vp, u32, vp]
    L.mtd_towers_download.argtypes = [vp, vp, vp, vp, vp]
    L.mtd_waypoints_layers_refresh.argtypes = [vp]
    _lib = L
    return L


def _p(a):
    """host numpy array (or None, or raw int device pointer) -> void*"""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    assert a.flags["C_CONTIGUOUS"], "arrays crossing the C ABI must be contiguous"
    return C.c_void_p(a.ctypes.data)


class Context:
    """One mtd_ctx: a device, a stream, and the device-resident grid / layers / monsters."""

    def __init__(self, device=0, stream=None):
        self.L = load()
        h = C.c_void_p()
        rc = self.L.mtd_create(C.byref(h), int(device), C.c_void_p(stream) if stream else None)
        if rc != 0:
            raise MtdError(rc, "mtd_create failed (no CUDA device? the library has no CPU fallback)")
        self.h = h
        # developer probe: MTD_DEBUG="key=value,key=value" applies mtd_debug_set to every new context (performance knobs only)
        for kv in filter(None, os.environ.get("MTD_DEBUG", "").split(",")):
            k, v = kv.split("=")
            if self.L.mtd_debug_set(self.h, int(k), int(v)) != 0:
                raise MtdError(-1, "MTD_DEBUG: key %s rejected" % k)
        self.W = self.H = self.T = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.mtd_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    DEBUG_QSEG_CAP = 1
    DEBUG_GROUP_WARPS = 2
    DEBUG_WALKER = 3
    DEBUG_TILE_DELTA = 4
    DEBUG_BIG_ASYNC = 5
    DEBUG_ASYNC_DELTA = 6
    DEBUG_BIG_TMA = 7
    DEBUG_INCREMENTAL = 10
    DEBUG_FUSED_MAPGEN = 11
    DEBUG_GEN_WARPS = 12
    DEBUG_ASYNC_WORKERS = 13
    DEBUG_ASYNC_LATE_LOOKS = 14
    DEBUG_ASYNC_CONT = 15
    DEBUG_ASYNC_NB_RULE = 8
    DEBUG_ASYNC_DEPTH = 9

    def big_stats(self):
        """rounds, tile visits, scheduler polls, dataflow flag of the last tiled update"""
        out = (C.c_uint32 * 4)()
        self._ck(self.L.mtd_big_stats(self.h, out))
        return dict(rounds=int(out[0]), visits=int(out[1]), polls=int(out[2]), dataflow=bool(out[3] & 1), tma=bool(out[3] & 2))

    def debug_set(self, key, value):
        self._ck(self.L.mtd_debug_set(self.h, int(key), int(value)))

    def _ck(self, rc):
        if rc != 0:
            raise MtdError(rc, self.L.mtd_last_error(self.h).decode())

    def synchronize(self):
        self._ck(self.L.mtd_synchronize(self.h))

    @property
    def launches(self):
        return int(self.L.mtd_launch_count(self.h))

    # ---- kernel timing probe ----
    K_WAYPOINTS_SMALL, K_MAPGEN_GRID, K_BIG_RELAX, K_BIG_DIRECTIONS, K_MONSTERS_TICK, K_BIG_PREPARE = range(6)

    def profile_enable(self, on=True):
        self._ck(self.L.mtd_profile_enable(self.h, int(bool(on))))

    def profile_read(self, kind):
        """(total device ms, bracketed regions) of `kind` since the last read; synchronises the stream"""
        ms, n = C.c_double(0), C.c_uint64(0)
        self._ck(self.L.mtd_profile_read(self.h, int(kind), C.byref(ms), C.byref(n)))
        return ms.value, int(n.value)

    # ---- grid ----
    def grid_upload(self, W, H, elev, flags, ox=0, oy=0):
        elev = np.ascontiguousarray(elev, np.int16)
        flags = np.ascontiguousarray(flags, np.uint16)
        assert elev.size == W * H and flags.size == W * H
        self._ck(self.L.mtd_grid_upload(self.h, W, H, ox, oy, _p(elev), _p(flags)))
        self.W, self.H, self.T = W, H, W * H

    def grid_upload_dev(self, W, H, d_elev, d_flags, ox=0, oy=0):
        self._ck(self.L.mtd_grid_upload_dev(self.h, W, H, ox, oy, _p(d_elev), _p(d_flags)))
        self.W, self.H, self.T = W, H, W * H

    def grid_set_flags(self, tile, flags):
        self._ck(self.L.mtd_grid_set_flags(self.h, int(tile), int(flags)))

    def neighbor_distances(self, a, b):
        a = np.ascontiguousarray(a, np.uint32)
        b = np.ascontiguousarray(b, np.uint32)
        out = np.empty(len(a), np.uint32)
        self._ck(self.L.mtd_neighbor_distances(self.h, _p(a), _p(b), len(a), _p(out)))
        return out

    def edge_weights(self):
        r = np.empty(self.T, np.uint32)
        d = np.empty(self.T, np.uint32)
        self._ck(self.L.mtd_edge_weights(self.h, _p(r), _p(d)))
        return r, d

    # ---- Directions::update ----
    def directions_update(self, sources, extra_blocked=M, want_tiles=True):
        src = np.ascontiguousarray(np.atleast_1d(sources), np.uint32)
        tiles = np.empty(self.T, np.uint32) if want_tiles else None
        dist = np.empty(self.T, np.uint32)
        self._ck(self.L.mtd_directions_update(self.h, _p(src), len(src), int(extra_blocked), _p(tiles), _p(dist)))
        return tiles, dist

    def directions_update_dev(self, sources, d_tiles, d_dist, extra_blocked=M):
        src = np.ascontiguousarray(np.atleast_1d(sources), np.uint32)
        self._ck(self.L.mtd_directions_update_dev(self.h, _p(src), len(src), int(extra_blocked), _p(d_tiles), _p(d_dist)))

    def layers_update_dev(self, targets, d_tiles, d_dist):
        """n independent single-target Directions layers, device outputs [n][T]"""
        t = np.ascontiguousarray(np.atleast_1d(targets), np.uint32)
        self._ck(self.L.mtd_directions_layers_dev(self.h, _p(t), len(t), _p(d_tiles), _p(d_dist)))

    # ---- row stripes of one large grid (one Context per stripe / GPU) ----
    def stripe_upload(self, W, H, y0, y1, elev_rows, flags_rows):
        """elev_rows / flags_rows: host arrays of rows [max(y0-1,0), min(y1+1,H)) of the full grid"""
        elev_rows = np.ascontiguousarray(elev_rows, np.int16)
        flags_rows = np.ascontiguousarray(flags_rows, np.uint16)
        rows = (y1 - y0) + (1 if y0 > 0 else 0) + (1 if y1 < H else 0)
        assert elev_rows.size == rows * W and flags_rows.size == rows * W
        self._ck(self.L.mtd_stripe_upload(self.h, W, H, y0, y1, _p(elev_rows), _p(flags_rows)))
        self.W, self.H, self.T = W, y1 - y0, W * (y1 - y0)

    def stripe_upload_dev(self, W, H, y0, y1, d_elev_rows, d_flags_rows):
        self._ck(self.L.mtd_stripe_upload_dev(self.h, W, H, y0, y1, _p(d_elev_rows), _p(d_flags_rows)))
        self.W, self.H, self.T = W, y1 - y0, W * (y1 - y0)

    def stripe_begin_dev(self, sources, d_dist):
        src = np.ascontiguousarray(np.atleast_1d(sources), np.uint32)
        self._ck(self.L.mtd_stripe_begin_dev(self.h, _p(src), len(src), _p(d_dist)))

    def stripe_exchange_dev(self, d_row_above, d_row_below):
        """merge the neighbours' boundary rows (device pointers or None) and relax again; returns the number of
        halo cells whose label dropped"""
        n = C.c_uint32(0)
        self._ck(self.L.mtd_stripe_exchange_dev(self.h, _p(d_row_above), _p(d_row_below), C.byref(n)))
        return int(n.value)

    def stripe_round_budget(self, rounds):
        self._ck(self.L.mtd_stripe_round_budget(self.h, int(rounds)))

    def stripe_pending(self):
        n = C.c_uint32(0)
        self._ck(self.L.mtd_stripe_pending(self.h, C.byref(n)))
        return int(n.value)

    # ---- peer-coupled stripes ----
    def stripe_peer_export(self):
        """this rank's peer block as a CUDA IPC handle (64 bytes)"""
        h = (C.c_uint8 * 64)()
        self._ck(self.L.mtd_stripe_peer_export(self.h, h))
        return bytes(h)

    def stripe_peer_connect(self, rank, world, handles):
        """handles: the 64-byte handles of all ranks, in rank order"""
        buf = b"".join(handles)
        assert len(buf) == 64 * world
        self._ck(self.L.mtd_stripe_peer_connect(self.h, int(rank), int(world), buf))

    def stripe_peer_connect_local(self, rank, world, contexts, ctas):
        """contexts: the Context of every rank (this process, this device), in rank order"""
        arr = (C.c_void_p * world)(*[c.h.value for c in contexts])
        self._ck(self.L.mtd_stripe_peer_connect_local(self.h, int(rank), int(world), arr, int(ctas)))

    def stripe_peer_launch_dev(self, sources, d_dist, d_tiles):
        src = np.ascontiguousarray(np.atleast_1d(sources), np.uint32)
        self._ck(self.L.mtd_stripe_peer_launch_dev(self.h, src.ctypes.data, len(src), _p(d_dist), _p(d_tiles)))

    def stripe_peer_finish(self):
        n = C.c_uint32(0)
        self._ck(self.L.mtd_stripe_peer_finish(self.h, C.byref(n)))
        return int(n.value)

    def stripe_finalize_dev(self, d_tiles):
        self._ck(self.L.mtd_stripe_finalize_dev(self.h, _p(d_tiles)))

    # ---- Waypoints::update ----
    def waypoints_update(self, wp_tiles, cap=None, want_layers=True):
        wp = np.ascontiguousarray(wp_tiles, np.uint32)
        n = len(wp)
        cap = self.T if cap is None else cap
        out = dict(
            tiles=np.empty((n, self.T), np.uint32) if want_layers else None,
            dist=np.empty((n, self.T), np.uint32) if want_layers else None,
            fullpath=np.full((n, cap), M, np.uint32) if cap else None,
            fullpath_len=np.zeros(n, np.uint32),
            full_distance=np.zeros(n, np.uint32),
            agg=np.zeros(3, np.uint32),
        )
        self._ck(self.L.mtd_waypoints_update(self.h, _p(wp), n, _p(out["tiles"]), _p(out["dist"]), _p(out["fullpath"]), cap, _p(out["fullpath_len"]), _p(out["full_distance"]), _p(out["agg"])))
        return out

    def waypoints_update_dev(self, wp_tiles):
        wp = np.ascontiguousarray(wp_tiles, np.uint32)
        self._ck(self.L.mtd_waypoints_update_dev(self.h, _p(wp), len(wp)))

    def waypoints_find(self, pos, bits):
        pos = np.ascontiguousarray(np.atleast_1d(pos), np.uint32)
        bits = np.ascontiguousarray(np.atleast_1d(bits), np.uint32)
        t = np.empty(len(pos), np.uint32)
        d = np.empty(len(pos), np.uint32)
        self._ck(self.L.mtd_waypoints_find(self.h, _p(pos), _p(bits), len(pos), _p(t), _p(d)))
        return t, d

    def layers_dev(self):
        a, b = C.c_void_p(), C.c_void_p()
        self._ck(self.L.mtd_waypoints_layers_dev(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def layers_refresh(self):
        self._ck(self.L.mtd_waypoints_layers_refresh(self.h))

    def placing_blocks(self, candidate):
        out = C.c_int32(0)
        self._ck(self.L.mtd_placing_blocks(self.h, int(candidate), C.byref(out)))
        return bool(out.value)

    # ---- generator ----
    def mapgen_batch(self, s1, s2, first, nmaps, W=101, H=101, ox=50, oy=50, radius=45, nwp=4, cap=1024, want_layers=True):
        T = W * H
        p = MapgenParams(W, H, ox, oy, radius, nwp, cap)
        out = dict(
            elev=np.empty((nmaps, T), np.int16),
            flags=np.empty((nmaps, T), np.uint16),
            wp_tiles=np.empty((nmaps, nwp), np.uint32),
            tiles=np.empty((nmaps, nwp, T), np.uint32) if want_layers else None,
            dist=np.empty((nmaps, nwp, T), np.uint32) if want_layers else None,
            fullpath=np.empty((nmaps, nwp, cap), np.uint32) if cap else None,
            fullpath_len=np.zeros((nmaps, nwp), np.uint32),
            full_distance=np.zeros((nmaps, nwp), np.uint32),
            agg=np.zeros((nmaps, 3), np.uint32),
        )
        self._ck(self.L.mtd_mapgen_batch(self.h, C.byref(p), s1, s2, first, nmaps, *[_p(out[k]) for k in ("elev", "flags", "wp_tiles", "tiles", "dist", "fullpath", "fullpath_len", "full_distance", "agg")]))
        return out

    def mapgen_batch_dev(self, s1, s2, first, nmaps, ptrs, W=101, H=101, ox=50, oy=50, radius=45, nwp=4, cap=1024):
        """ptrs: dict of raw device pointers (ints) for elev, flags, wp_tiles, tiles, dist, fullpath, fullpath_len, full_distance, agg"""
        p = MapgenParams(W, H, ox, oy, radius, nwp, cap)
        self._ck(self.L.mtd_mapgen_batch_dev(self.h, C.byref(p), s1, s2, first, nmaps, *[_p(ptrs.get(k)) for k in ("elev", "flags", "wp_tiles", "tiles", "dist", "fullpath", "fullpath_len", "full_distance", "agg")]))

    NOALLOC_SCRATCH_BYTES = 4 << 20

    def waypoints_paths_noalloc(self, W, H, nmaps, n, d_elev, d_flags, d_wp_tiles, d_fullpath_len, d_full_distance, d_agg):
        """mtd_waypoints_paths_noalloc: raw device pointers (ints) in and out; asynchronous; allocates nothing"""
        self._ck(self.L.mtd_waypoints_paths_noalloc(self.h, W, H, nmaps, n, _p(d_elev), _p(d_flags), _p(d_wp_tiles), _p(d_fullpath_len), _p(d_full_distance), _p(d_agg)))

    def grid_from_elevations(self, elev, W=101, H=101, ox=50, oy=50, radius=45):
        """newGrid after the elevation sampling on the caller's int16 elevations [nmaps][T] -> flags [nmaps][T]"""
        elev = np.ascontiguousarray(elev, np.int16).reshape(-1, W * H)
        flags = np.empty(elev.shape, np.uint16)
        p = MapgenParams(W, H, ox, oy, radius, 0, 0)
        self._ck(self.L.mtd_grid_from_elevations(self.h, C.byref(p), elev.shape[0], _p(elev), _p(flags)))
        return flags

    def sdf_lattice(self, s1, s2, flags, W=101, H=101, ox=50, oy=50, box_a=(-55, -15, -55), box_b=(55, 15, 55), res=(220, 60, 220)):
        """Maker::sdf on the marching-cubes lattice -> float32 [rz][ry][rx]"""
        flags = np.ascontiguousarray(flags, np.uint16)
        a, b, r = np.array(box_a, np.float32), np.array(box_b, np.float32), np.array(res, np.int32)
        out = np.empty((int(r[2]), int(r[1]), int(r[0])), np.float32)
        self._ck(self.L.mtd_sdf_lattice(self.h, s1, s2, W, H, ox, oy, _p(flags), _p(a), _p(b), _p(r), _p(out)))
        return out

    def elevation_samples(self, s1, s2, xy):
        xy = np.ascontiguousarray(xy, np.float32)
        out = np.empty(len(xy), np.float32)
        self._ck(self.L.mtd_elevation_samples(self.h, s1, s2, _p(xy), len(xy), _p(out)))
        return out

    # ---- monsters ----
    def monsters_upload(self, a):
        n = len(a["tile"])
        g = lambda k: _p(a[k]) if a.get(k) is not None else None
        self._ck(self.L.mtd_monsters_upload(self.h, n, g("tile"), g("tileStart"), g("tileEnd"), g("timeStart"), g("timeEnd"), g("bits"), g("timeToArrive"), g("dotflags"), g("damage"), g("alive")))
        self.nmon = n

    def monsters_download(self):
        n = self.nmon
        a = {k: np.empty(n, np.uint32) for k in ("tile", "tileStart", "tileEnd", "timeStart", "timeEnd", "bits")}
        a["alive"] = np.empty(n, np.uint8)
        self._ck(self.L.mtd_monsters_download(self.h, *[_p(a[k]) for k in ("tile", "tileStart", "tileEnd", "timeStart", "timeEnd", "bits", "alive")]))
        return a

    def monsters_tick(self, game_time):
        dmg, alive = C.c_int64(0), C.c_uint32(0)
        self._ck(self.L.mtd_monsters_tick(self.h, int(game_time), C.byref(dmg), C.byref(alive)))
        return dmg.value, alive.value

    def monsters_tick_dev(self, game_time):
        self._ck(self.L.mtd_monsters_tick_dev(self.h, int(game_time)))

    def monsters_positions(self, game_time):
        out = np.empty((self.nmon, 3), np.float32)
        self._ck(self.L.mtd_monsters_positions(self.h, int(game_time), _p(out)))
        return out

    def monsters_positions_dev(self, game_time, d_out):
        self._ck(self.L.mtd_monsters_positions_dev(self.h, int(game_time), _p(d_out)))

    # ---- damage over time ----
    def monsters_dots_upload(self, duration, damage, life, resist=None, money=None):
        """duration / damage: uint32 [5][n] (Physical, Fire, Water, Poison, Magic); life int64 [n]; resist uint8 [n]; money uint32 [n]"""
        self._ck(self.L.mtd_monsters_dots_upload(self.h, _p(np.ascontiguousarray(duration, np.uint32)), _p(np.ascontiguousarray(damage, np.uint32)),
                                                 _p(None if resist is None else np.ascontiguousarray(resist, np.uint8)), _p(np.ascontiguousarray(life, np.int64)),
                                                 _p(None if money is None else np.ascontiguousarray(money, np.uint32))))

    def monsters_dots_tick(self):
        money, killed = C.c_int64(0), C.c_uint32(0)
        self._ck(self.L.mtd_monsters_dots_tick(self.h, C.byref(money), C.byref(killed)))
        return money.value, killed.value

    def monsters_dots_tick_dev(self):
        self._ck(self.L.mtd_monsters_dots_tick_dev(self.h))

    def monsters_dots_download(self):
        n = self.nmon
        out = dict(duration=np.empty((5, n), np.uint32), damage=np.empty((5, n), np.uint32), life=np.empty(n, np.int64), alive=np.empty(n, np.uint8))
        self._ck(self.L.mtd_monsters_dots_download(self.h, _p(out["duration"]), _p(out["damage"]), _p(out["life"]), _p(out["alive"])))
        return out

    # ---- towers: spatial index + attack pass ----
    TOWER_U32 = ("tile", "damage", "overTime", "firingPeriod")
    TOWER_FIELDS = ("tile", "damage", "overTime", "firingPeriod", "firingRange", "splashRadius", "manaCost", "invalidClasses", "element", "bonus", "targeting", "firingDelay", "mana")

    def monsters_class_upload(self, monster_class):
        self._ck(self.L.mtd_monsters_class_upload(self.h, _p(None if monster_class is None else np.ascontiguousarray(monster_class, np.uint8))))

    def towers_upload(self, tw):
        """tw: dict of arrays named like the reference's components (TOWER_FIELDS)"""
        dt = dict(firingRange=np.float32, splashRadius=np.float32, invalidClasses=np.uint8, element=np.uint8, bonus=np.uint8, targeting=np.uint8)
        arrs = [np.ascontiguousarray(tw[k], dt.get(k, np.uint32)) for k in self.TOWER_FIELDS]
        self.ntowers = len(arrs[0])
        self._ck(self.L.mtd_towers_upload(self.h, self.ntowers, *[_p(a) for a in arrs]))

    def towers_attack_dev(self, game_time, random_words=None):
        rw = None if random_words is None else np.ascontiguousarray(random_words, np.uint64)
        self._ck(self.L.mtd_towers_attack_dev(self.h, int(game_time), _p(rw)))
        self._rw_keep = rw  # the copy is asynchronous

    def towers_download(self):
        n = self.ntowers
        out = dict(firingDelay=np.empty(n, np.uint32), mana=np.empty(n, np.uint32), target=np.empty(n, np.int32), hits=np.empty(n, np.uint32))
        self._ck(self.L.mtd_towers_download(self.h, _p(out["firingDelay"]), _p(out["mana"]), _p(out["target"]), _p(out["hits"])))
        return out

    def monster_spawn(self, game_time, speed):
        out = np.zeros(3, np.uint32)
        self._ck(self.L.mtd_monster_spawn(self.h, int(game_time), float(speed), _p(out)))
        return out
