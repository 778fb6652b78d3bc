"""ctypes bindings for the two CPU checkers (TEST INFRASTRUCTURE, never imported by mazetd_b200/):

* ``Oracle`` -- oracle/libmazetd_oracle.so, the plain-C restatement (oracle/mazetd_oracle.c).
* ``Ref``    -- oracle/_ref/libmazetd_ref.so, the reference's own sources compiled in place
                (oracle/Makefile `ref`); present here and, prebuilt, on the GPU box.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libmazetd_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libmazetd_ref.so")
ADAPTER_SO = os.path.join(ROOT, "oracle", "_ref", "libmazetd_adapter.so")

M = 0xFFFFFFFF
KAT_SEED = (1346519564496, 42245614964156)  # externals/cage/sources/test-core/random.cpp:47

u8p = C.POINTER(C.c_uint8)
u16p = C.POINTER(C.c_uint16)
i16p = C.POINTER(C.c_int16)
u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)
f32p = C.POINTER(C.c_float)


def ptr(a, t):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(t)


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"])


def build_adapter():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "adapter"])


def build_ref():
    if os.path.isdir("/root/reference/sources/map"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref"])


class MoRng(C.Structure):
    _fields_ = [("s", C.c_uint64 * 2)]


class MoGrid(C.Structure):
    _fields_ = [("W", C.c_int32), ("H", C.c_int32), ("ox", C.c_int32), ("oy", C.c_int32), ("elev", i16p), ("flags", u16p)]


class MoWaypoints(C.Structure):
    _fields_ = [("grid", C.POINTER(MoGrid)), ("n", C.c_int), ("wp_tiles", u32p), ("tiles", u32p), ("dist", u32p)]


class MoProcedural(C.Structure):
    _fields_ = [("off_x", C.c_float), ("off_y", C.c_float), ("seeds", C.c_uint32 * 5), ("fractal_bounding", C.c_float)]


class MoTowers(C.Structure):
    _fields_ = [("n", C.c_int)] + [(k, C.c_void_p) for k in ("tile", "damage", "overTime", "firingPeriod", "manaCost", "firingRange", "splashRadius", "invalidClasses", "element", "bonus", "targeting",
                                                              "firingDelay", "mana", "random_words", "target_out", "hits_out")]


class MoMonsters(C.Structure):
    _fields_ = [
        ("n", C.c_int),
        ("tile", u32p),
        ("tileStart", u32p),
        ("tileEnd", u32p),
        ("timeStart", u32p),
        ("timeEnd", u32p),
        ("bits", u32p),
        ("timeToArrive", u32p),
        ("dotflags", u8p),
        ("damage", u32p),
        ("alive", u8p),
    ]


class Oracle:
    """Thin numpy wrapper over oracle/libmazetd_oracle.so."""

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        L = self.L = C.CDLL(ORACLE_SO)
        L.mo_rng_next.restype = C.c_uint64
        L.mo_rng_chance_double.restype = C.c_double
        L.mo_rng_chance.restype = C.c_float
        L.mo_rng_range_u32.restype = C.c_uint32
        L.mo_rng_range_i32.restype = C.c_int32
        L.mo_rng_range_f.restype = C.c_float
        L.mo_rng_range_f.argtypes = [C.c_void_p, C.c_float, C.c_float]
        L.mo_rtos16.restype = C.c_int16
        L.mo_rtos16.argtypes = [C.c_float]
        L.mo_rtos32.restype = C.c_int32
        L.mo_rtos32.argtypes = [C.c_float]
        L.mo_stor.restype = C.c_float
        L.mo_stor.argtypes = [C.c_int32]
        L.mo_grid_index_i.restype = C.c_uint32
        L.mo_grid_index_f.restype = C.c_uint32
        L.mo_grid_index_f.argtypes = [C.c_void_p, C.c_float, C.c_float]
        L.mo_neighbor_distance.restype = C.c_uint32
        L.mo_elevation.restype = C.c_float
        L.mo_elevation.argtypes = [C.c_void_p, C.c_float, C.c_float]
        L.mo_waypoints_fullpath.restype = C.c_uint32
        L.mo_add_mana.restype = C.c_uint32
        L.mo_mapgen_batch_checksum.restype = C.c_uint64
        L.mo_mapgen_batch_checksum.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_int]
        L.mo_mapgen.argtypes = [C.c_uint64, C.c_uint64] + [C.c_int32] * 5 + [C.c_int] + [C.c_void_p] * 6 + [C.c_uint32] + [C.c_void_p] * 3
        L.mo_monster_spawn.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_float, C.c_void_p]
        L.mo_monsters_tick.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]

    # -- helpers -------------------------------------------------------------------------------
    @staticmethod
    def grid(W, H, ox, oy, elev, flags):
        g = MoGrid(W, H, ox, oy, ptr(elev, i16p), ptr(flags, u16p))
        g._keep = (elev, flags)
        return g

    def rng(self, s1, s2):
        r = MoRng()
        r.s[0], r.s[1] = s1, s2
        return r

    def edge_weights(self, W, H, elev):
        g = self.grid(W, H, 0, 0, elev, None)
        r = np.zeros(W * H, np.uint32)
        d = np.zeros(W * H, np.uint32)
        self.L.mo_edge_weights(C.byref(g), ptr(r, u32p), ptr(d, u32p))
        return r, d

    def directions_update(self, W, H, elev, flags, sources, ox=0, oy=0):
        g = self.grid(W, H, ox, oy, elev, flags)
        src = np.ascontiguousarray(np.atleast_1d(sources), np.uint32)
        tiles = np.empty(W * H, np.uint32)
        dist = np.empty(W * H, np.uint32)
        rc = self.L.mo_directions_update(C.byref(g), ptr(src, u32p), len(src), ptr(tiles, u32p), ptr(dist, u32p))
        assert rc == 0
        return tiles, dist

    def waypoints_update(self, W, H, elev, flags, wp_tiles, ox=0, oy=0, threads=1, cap=None):
        g = self.grid(W, H, ox, oy, elev, flags)
        wp = np.ascontiguousarray(wp_tiles, np.uint32)
        n = len(wp)
        T = W * H
        cap = cap or T
        out = dict(
            tiles=np.empty((n, T), np.uint32),
            dist=np.empty((n, T), np.uint32),
            fullpath=np.full((n, cap), M, np.uint32),
            fullpath_len=np.zeros(n, np.uint32),
            full_distance=np.zeros(n, np.uint32),
            agg=np.zeros(3, np.uint32),
        )
        rc = self.L.mo_waypoints_update(C.byref(g), ptr(wp, u32p), n, threads, ptr(out["tiles"], u32p), ptr(out["dist"], u32p), ptr(out["fullpath"], u32p), C.c_uint32(cap), ptr(out["fullpath_len"], u32p), ptr(out["full_distance"], u32p), ptr(out["agg"], u32p))
        assert rc == 0
        return out

    def waypoints_view(self, W, H, elev, flags, wp_tiles, tiles, dist, ox=0, oy=0):
        g = self.grid(W, H, ox, oy, elev, flags)
        wp = np.ascontiguousarray(wp_tiles, np.uint32)
        w = MoWaypoints(C.pointer(g), len(wp), ptr(wp, u32p), ptr(tiles, u32p), ptr(dist, u32p))
        w._keep = (g, wp, tiles, dist)
        return w

    def find(self, wview, pos, bits):
        out = np.zeros(2, np.uint32)
        self.L.mo_waypoints_find(C.byref(wview), C.c_uint32(pos), C.c_uint32(bits), ptr(out, u32p))
        return int(out[0]), int(out[1])

    def mapgen(self, s1, s2, W=101, H=101, ox=50, oy=50, radius=45, nwp=4, cap=None):
        T = W * H
        cap = cap or T
        out = dict(
            elev=np.zeros(T, np.int16),
            flags=np.zeros(T, np.uint16),
            wp_tiles=np.zeros(nwp, np.uint32),
            tiles=np.empty((nwp, T), np.uint32),
            dist=np.empty((nwp, T), np.uint32),
            fullpath=np.full((nwp, cap), M, np.uint32),
            fullpath_len=np.zeros(nwp, np.uint32),
            full_distance=np.zeros(nwp, np.uint32),
            agg=np.zeros(3, np.uint32),
        )
        rc = self.L.mo_mapgen(s1, s2, W, H, ox, oy, radius, nwp, *[out[k].ctypes.data for k in ("elev", "flags", "wp_tiles", "tiles", "dist", "fullpath")], cap, *[out[k].ctypes.data for k in ("fullpath_len", "full_distance", "agg")])
        assert rc == 0
        return out

    def grid_stages(self, W, H, ox, oy, radius, elev):
        """integer stages of newGrid on given elevations -> flags"""
        flags = np.zeros(W * H, np.uint16)
        self.L.mo_grid_flags_from_elev(W, H, ox, oy, radius, ptr(elev, i16p), ptr(flags, u16p))
        self.L.mo_fill_unreachable(W, H, ptr(flags, u16p))
        self.L.mo_find_border(W, H, ptr(flags, u16p))
        self.L.mo_add_mana(W, H, ptr(flags, u16p))
        return flags

    def sdf_lattice(self, s1, s2, W, H, ox, oy, flags, box_a=(-55, -15, -55), box_b=(55, 15, 55), res=(220, 60, 220)):
        """Maker::sdf on the marching-cubes lattice for the Procedural of seed (s1, s2) and the given grid flags"""
        r = self.rng(s1, s2)
        pr = MoProcedural()
        self.L.mo_procedural_init(C.byref(pr), C.byref(r))
        g = self.grid(W, H, ox, oy, None, np.ascontiguousarray(flags, np.uint16))
        a, b, rs = np.array(box_a, np.float32), np.array(box_b, np.float32), np.array(res, np.int32)
        out = np.empty((int(rs[2]), int(rs[1]), int(rs[0])), np.float32)
        self.L.mo_sdf_lattice(C.byref(pr), C.byref(g), C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data), C.c_void_p(rs.ctypes.data), C.c_void_p(out.ctypes.data))
        return out

    def monsters(self, arrs):
        ms = MoMonsters(
            len(arrs["tile"]),
            ptr(arrs["tile"], u32p),
            ptr(arrs["tileStart"], u32p),
            ptr(arrs["tileEnd"], u32p),
            ptr(arrs["timeStart"], u32p),
            ptr(arrs["timeEnd"], u32p),
            ptr(arrs["bits"], u32p),
            ptr(arrs["timeToArrive"], u32p),
            ptr(arrs.get("dotflags"), u8p),
            ptr(arrs.get("damage"), u32p),
            ptr(arrs["alive"], u8p),
        )
        ms._keep = arrs
        return ms

    def monsters_tick(self, wview, game_time, arrs):
        ms = self.monsters(arrs)
        dmg = C.c_int64(0)
        self.L.mo_monsters_tick(C.addressof(wview), game_time, C.addressof(ms), C.addressof(dmg))
        return dmg.value

    def monsters_dots_tick(self, duration, damage, life, alive, resist=None, money=None):
        """in place on uint32 [5][n] duration / damage, int64 life, uint8 alive; returns (money of the killed, killed)"""
        n = len(life)
        ms, kl = C.c_int64(0), C.c_uint32(0)
        vp = lambda a: None if a is None else C.c_void_p(a.ctypes.data)
        self.L.mo_monsters_dots_tick(n, vp(duration), vp(damage), vp(resist), vp(life), vp(money), vp(alive), C.byref(ms), C.byref(kl))
        return ms.value, kl.value

    TOWER_DTYPES = dict(firingRange=np.float32, splashRadius=np.float32, invalidClasses=np.uint8, element=np.uint8, bonus=np.uint8, targeting=np.uint8)

    def towers_attack(self, W, H, ox, oy, elev, game_time, tw, mon, monster_class, life, dot_duration, dot_damage, random_words=None):
        """one "attacks" tick (sources/game/attacks.cpp:132-160). tw: dict of tower arrays (firingDelay and mana are updated in
        place), mon: monster SoA dict, dots uint32 [5][n] updated in place. Returns (target per tower, hits per tower)."""
        g = self.grid(W, H, ox, oy, elev, None)
        n = len(tw["tile"])
        for k in tw:
            want = self.TOWER_DTYPES.get(k, np.uint32)
            assert tw[k].dtype == want and tw[k].flags["C_CONTIGUOUS"], k
        target = np.full(n, -1, np.int32)
        hits = np.zeros(n, np.uint32)
        rw = None if random_words is None else np.ascontiguousarray(random_words, np.uint64)
        t = MoTowers()
        t.n = n
        for k in ("tile", "damage", "overTime", "firingPeriod", "manaCost", "firingRange", "splashRadius", "invalidClasses", "element", "bonus", "targeting", "firingDelay", "mana"):
            setattr(t, k, tw[k].ctypes.data)
        t.random_words = None if rw is None else rw.ctypes.data
        t.target_out = target.ctypes.data
        t.hits_out = hits.ctypes.data
        vp = lambda a: None if a is None else C.c_void_p(a.ctypes.data)
        self.L.mo_towers_attack(C.byref(g), C.c_uint32(game_time), C.byref(t), len(mon["tileStart"]), vp(mon["tileStart"]), vp(mon["tileEnd"]), vp(mon["timeStart"]), vp(mon["timeEnd"]), vp(mon["alive"]),
                                vp(monster_class), vp(mon["timeToArrive"]), vp(life), vp(dot_duration), vp(dot_damage))
        return target, hits

    def monster_spawn(self, wview, spawner, game_time, speed):
        out = np.zeros(3, np.uint32)
        self.L.mo_monster_spawn(C.addressof(wview), spawner, game_time, speed, out.ctypes.data)
        return out

    def monsters_position(self, W, H, ox, oy, elev, game_time, tileStart, tileEnd, timeStart, timeEnd):
        g = self.grid(W, H, ox, oy, elev, None)
        n = len(tileStart)
        out = np.zeros((n, 3), np.float32)
        self.L.mo_monsters_position(C.byref(g), C.c_uint32(game_time), n, ptr(tileStart, u32p), ptr(tileEnd, u32p), ptr(timeStart, u32p), ptr(timeEnd, u32p), ptr(out, f32p))
        return out

    def placing_blocks(self, W, H, elev, flags, wp_tiles, candidate, monster_tiles=None, monster_alive=None, ox=0, oy=0):
        g = self.grid(W, H, ox, oy, elev, flags)
        wp = np.ascontiguousarray(wp_tiles, np.uint32)
        st = np.empty(W * H, np.uint32)
        sd = np.empty(W * H, np.uint32)
        nm = 0 if monster_tiles is None else len(monster_tiles)
        return self.L.mo_placing_blocks(C.byref(g), ptr(wp, u32p), len(wp), C.c_uint32(candidate), ptr(monster_tiles, u32p), ptr(monster_alive, u8p), nm, ptr(st, u32p), ptr(sd, u32p))


class Ref:
    """Thin numpy wrapper over oracle/_ref/libmazetd_ref.so (the reference's own code)."""

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def __init__(self, path=None):
        """path: another library exporting the same ref_* entry layer -- oracle/_ref/libmazetd_adapter.so is the product's
        host adapter (mazetd_b200/host) linked under it instead of the reference's paths.cpp / grid.cpp"""
        L = self.L = C.CDLL(path or REF_SO)
        L.ref_rng_next.restype = C.c_uint64
        L.ref_rng_chance.restype = C.c_float
        L.ref_rng_range_u32.restype = C.c_uint32
        L.ref_rng_seed.argtypes = [C.c_uint64, C.c_uint64]
        L.ref_rng_kat_range4i.argtypes = [C.c_uint64, C.c_uint64, C.c_int32, C.c_int32, C.c_void_p]
        L.ref_rng_kat_range4.argtypes = [C.c_uint64, C.c_uint64, C.c_float, C.c_float, C.c_void_p]
        L.ref_rtos32.restype = C.c_int32
        L.ref_rtos32.argtypes = [C.c_float]
        L.ref_rtos16.restype = C.c_int16
        L.ref_rtos16.argtypes = [C.c_float]
        L.ref_stor.restype = C.c_float
        L.ref_stor.argtypes = [C.c_int32]
        L.ref_grid_index_i.restype = C.c_uint32
        L.ref_grid_index_f.restype = C.c_uint32
        L.ref_grid_index_f.argtypes = [C.c_int] * 4 + [C.c_float, C.c_float]
        L.ref_map_create.restype = C.c_void_p
        L.ref_map_create.argtypes = [C.c_int] * 4 + [C.c_void_p] * 3 + [C.c_int]
        L.ref_mapgen.restype = C.c_void_p
        L.ref_mapgen.argtypes = [C.c_uint64, C.c_uint64]
        L.ref_mapgen_grid_only.restype = C.c_void_p
        L.ref_mapgen_grid_only.argtypes = [C.c_uint64, C.c_uint64]
        L.ref_elevation_samples.argtypes = [C.c_uint64, C.c_uint64, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_waypoints_fullpath.restype = C.c_uint32
        L.ref_waypoints_fullpath.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_void_p]
        for f in ("ref_map_destroy", "ref_waypoints_update", "ref_map_waypoint_count"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.ref_map_set_flags.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_map_get_grid.argtypes = [C.c_void_p] * 4
        L.ref_map_waypoint_tiles.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_waypoints_get_layer.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_waypoints_aggregates.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_waypoints_find.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        L.ref_monsters_tick.argtypes = [C.c_void_p, C.c_uint32, C.c_int] + [C.c_void_p] * 11
        L.ref_monster_spawn.argtypes = [C.c_void_p, C.c_uint32, C.c_float, C.c_void_p]
        L.ref_monsters_position.argtypes = [C.c_void_p, C.c_uint32, C.c_int] + [C.c_void_p] * 5

    def edge_weights(self, W, H, elev):
        r = np.zeros(W * H, np.uint32)
        d = np.zeros(W * H, np.uint32)
        self.L.ref_edge_weights(W, H, ptr(elev, i16p), ptr(r, u32p), ptr(d, u32p))
        return r, d

    def directions_update(self, W, H, elev, flags, tile, ox=0, oy=0):
        tiles = np.empty(W * H, np.uint32)
        dist = np.empty(W * H, np.uint32)
        self.L.ref_directions_update(W, H, ox, oy, ptr(elev, i16p), ptr(flags, u16p), C.c_uint32(int(tile)), ptr(tiles, u32p), ptr(dist, u32p))
        return tiles, dist

    def map_create(self, W, H, elev, flags, wp_tiles, ox=0, oy=0):
        wp = np.ascontiguousarray(wp_tiles, np.uint32)
        return self.L.ref_map_create(W, H, ox, oy, elev.ctypes.data, flags.ctypes.data, wp.ctypes.data, len(wp))

    def map_collect(self, h, T, cap=None):
        """run nothing; just read everything a Waypoints holds"""
        L = self.L
        n = L.ref_map_waypoint_count(h)
        cap = cap or T
        out = dict(
            wp_tiles=np.zeros(n, np.uint32),
            tiles=np.empty((n, T), np.uint32),
            dist=np.empty((n, T), np.uint32),
            fullpath=np.full((n, cap), M, np.uint32),
            fullpath_len=np.zeros(n, np.uint32),
            full_distance=np.zeros(n, np.uint32),
            agg=np.zeros(3, np.uint32),
        )
        L.ref_map_waypoint_tiles(h, out["wp_tiles"].ctypes.data)
        for i in range(n):
            L.ref_waypoints_get_layer(h, i, out["tiles"][i].ctypes.data, out["dist"][i].ctypes.data)
            fd = C.c_uint32(0)
            out["fullpath_len"][i] = L.ref_waypoints_fullpath(h, i, out["fullpath"][i].ctypes.data, cap, C.addressof(fd))
            out["full_distance"][i] = fd.value
        L.ref_waypoints_aggregates(h, out["agg"].ctypes.data)
        return out

    def waypoints_update(self, W, H, elev, flags, wp_tiles, ox=0, oy=0, cap=None):
        h = self.map_create(W, H, elev, flags, wp_tiles, ox, oy)
        self.L.ref_waypoints_update(h)
        out = self.map_collect(h, W * H, cap)
        self.L.ref_map_destroy(h)
        return out

    def mapgen(self, s1, s2, cap=None):
        h = self.L.ref_mapgen(s1, s2)
        T = 101 * 101
        out = self.map_collect(h, T, cap)
        out["elev"] = np.zeros(T, np.int16)
        out["flags"] = np.zeros(T, np.uint16)
        self.L.ref_map_get_grid(h, None, out["elev"].ctypes.data, out["flags"].ctypes.data)
        self.L.ref_map_destroy(h)
        return out

    def sdf_lattice(self, s1, s2, box_a=(-55, -15, -55), box_b=(55, 15, 55), res=(220, 60, 220)):
        """(sdf [rz][ry][rx], flags of the generated grid) from the compiled reference pieces"""
        a, b, rs = np.array(box_a, np.float32), np.array(box_b, np.float32), np.array(res, np.int32)
        out = np.empty((int(rs[2]), int(rs[1]), int(rs[0])), np.float32)
        flags = np.zeros(101 * 101, np.uint16)
        self.L.ref_sdf_lattice(C.c_uint64(s1), C.c_uint64(s2), C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data), C.c_void_p(rs.ctypes.data), C.c_void_p(out.ctypes.data), C.c_void_p(flags.ctypes.data))
        return out, flags

    def find(self, h, pos, bits):
        out = np.zeros(2, np.uint32)
        self.L.ref_waypoints_find(h, pos, bits, out.ctypes.data)
        return int(out[0]), int(out[1])

    def monsters_tick(self, h, game_time, arrs):
        dmg = C.c_int64(0)
        n = len(arrs["tile"])
        g = lambda k: None if arrs.get(k) is None else arrs[k].ctypes.data
        self.L.ref_monsters_tick(h, game_time, n, g("tile"), g("tileStart"), g("tileEnd"), g("timeStart"), g("timeEnd"), g("bits"), g("timeToArrive"), g("dotflags"), g("damage"), g("alive"), C.addressof(dmg))
        return dmg.value

    def monster_spawn(self, h, game_time, speed):
        out = np.zeros(3, np.uint32)
        self.L.ref_monster_spawn(h, game_time, speed, out.ctypes.data)
        return out

    def monsters_position(self, h, game_time, tileStart, tileEnd, timeStart, timeEnd):
        n = len(tileStart)
        out = np.zeros((n, 3), np.float32)
        self.L.ref_monsters_position(h, game_time, n, tileStart.ctypes.data, tileEnd.ctypes.data, timeStart.ctypes.data, timeEnd.ctypes.data, out.ctypes.data)
        return out


class OracleGame:
    """the four map operations of mazetd_b200.game_events over the CPU oracle (scripted place/remove workload, C3)"""
    def __init__(self, oracle, W, H, elev, flags, wp, monsters):
        self.o, self.W, self.H, self.elev, self.wp, self.a = oracle, W, H, elev, wp, monsters
        self.flags = np.array(flags, np.uint16)
        self.rebuild()

    def trial_blocks(self, tile):
        return bool(self.o.placing_blocks(self.W, self.H, self.elev, self.flags, self.wp, int(tile), self.a["tile"], self.a["alive"], 50, 50))

    def set_flags(self, tile, value):
        pass  # self.flags is the grid

    def rebuild(self):
        self.w = self.o.waypoints_update(self.W, self.H, self.elev, self.flags, self.wp, 50, 50)
        self.view = self.o.waypoints_view(self.W, self.H, self.elev, self.flags, self.wp, self.w["tiles"], self.w["dist"], 50, 50)

    def tick(self, t):
        self.o.monsters_tick(self.view, t, self.a)

    # the two layer hooks of mazetd_b200.sharded_game (C3 over several ranks)
    def export_layers(self):
        import torch

        return torch.from_numpy(self.w["tiles"].view(np.int32).reshape(-1)), torch.from_numpy(self.w["dist"].view(np.int32).reshape(-1))

    def import_layers(self, tiles, d):
        # the broadcast wrote into self.w's arrays (the tensors alias them); the view holds pointers to the same arrays
        self.view = self.o.waypoints_view(self.W, self.H, self.elev, self.flags, self.wp, self.w["tiles"], self.w["dist"], 50, 50)
