#!/usr/bin/env python
"""bench.py -- the driver's measurement contract for the mazetd flow-field hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c4|c5|c1]

One "step" = one pass of the hot path over one batch of synthetic fixed-seed input. Default workload (N=1) is
BASELINE.json configs[1] ("C2"): a batch of 4096 procedurally generated 101x101 maps per GPU -- newProcedural ->
newGrid -> newWaypoints incl. Waypoints::update (4 distance + direction layers, full paths, aggregates).
Metric = tile relaxations / s (one tile finalised in one layer), map rebuilds / s is reported beside it.

* `value`   : device-resident throughput (seeds in, all outputs written to HBM), CUDA events, max over ranks.
* `e2e`     : the same work through the host-buffer C-ABI call mtd_mapgen_batch (pinned host outputs, every byte the
              reference's Grid/Waypoints hold is copied device->host inside the timed region).
* `roofline`: dominant kernel timed live with CUDA events on the library's stream (mtd_profile_*), algorithmic
              bytes per DESIGN.md, peak from MEASURED_PEAKS.json.
* `cpu_baseline` / `--impl reference`: the reference's own sources compiled in place (oracle/_ref) on this box's
              host cores; falls back to the plain-C oracle port when that library did not travel.

Under torchrun (N > 1) every rank owns one GPU and an independent batch of maps (weak scaling, no data-path
collective); only the timing is reduced (max over ranks).
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

KAT_SEED = (1346519564496, 42245614964156)  # externals/cage/sources/test-core/random.cpp:47
W = H = 101
T = W * H
NWP = 4
FULLPATH_CAP = 512  # tiles per Waypoint::fullPath buffer; the longest path of the batch is 281 (a longer one is an error, not a truncation)


# --------------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi fields via NVML), runs during the timed region
# --------------------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index):
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thr = None
        self.h = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None

    def _loop(self):
        nv = self.nv
        names = {
            0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
            0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting", 0x10: "sync_boost",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def start(self):
        if self.h is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join()
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# --------------------------------------------------------------------------------------------------------
# CPU checkers used ONLY as timed baselines here (never on the product path)
# --------------------------------------------------------------------------------------------------------
def cpu_mapgen_fn():
    """returns (kind, fn(first, nmaps, threads) -> checksum)"""
    ref_so = os.path.join(ROOT, "oracle", "_ref", "libmazetd_ref.so")
    if os.path.exists(ref_so):
        L = C.CDLL(ref_so)
        if hasattr(L, "ref_mapgen_batch"):
            L.ref_mapgen_batch.restype = C.c_uint64
            L.ref_mapgen_batch.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int]
            return "reference", lambda first, n, thr: L.ref_mapgen_batch(KAT_SEED[0], KAT_SEED[1], first, n, thr)
    ora_so = os.path.join(ROOT, "oracle", "libmazetd_oracle.so")
    if not os.path.exists(ora_so):
        import subprocess

        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    L = C.CDLL(ora_so)
    L.mo_mapgen_batch_checksum.restype = C.c_uint64
    L.mo_mapgen_batch_checksum.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_int]
    return "port", lambda first, n, thr: L.mo_mapgen_batch_checksum(KAT_SEED[0] + first, KAT_SEED[1], n, thr)


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def time_cpu_sample(target_s=12.0):
    """time the CPU implementation on a bounded sample of the C2 workload (all host cores, one map per thread)"""
    kind, fn = cpu_mapgen_fn()
    cores = host_cores()
    probe = max(2 * cores, 16)
    t0 = time.perf_counter()
    fn(0, probe, cores)
    dt = time.perf_counter() - t0
    n = int(min(4096, max(probe, probe * target_s / max(dt, 1e-6))))
    t0 = time.perf_counter()
    fn(0, n, cores)
    dt = time.perf_counter() - t0
    return {
        "value": n * T * NWP / dt,
        "unit": "tile relaxations/s",
        "maps_per_s": n / dt,
        "cores": cores,
        "kind": kind,
        "cpu": cpu_model(),
        "sample": f"{n} of 4096 maps (seeds k=0..{n - 1}), newProcedural+newGrid+newWaypoints incl. Waypoints::update, one map per thread, {dt:.2f} s wall",
    }


# --------------------------------------------------------------------------------------------------------
def run_reference_other(args):
    """CPU reference figures for the configurations that are not the bench headline: C1 = trial +
    Waypoints::update on the default map after the scripted placement, C5 = 8 Directions::update on a 4096^2 grid.
    The reference's own compiled code (oracle/_ref) when it travelled, else the oracle port."""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    from mazetd_b200 import game_events, synth

    cores = host_cores()
    use_ref = oracle_lib.Ref.available()
    if not use_ref:
        oracle_lib.build_oracle()
    eng = oracle_lib.Ref() if use_ref else oracle_lib.Oracle()
    kind = "reference" if use_ref else "port"
    if args.workload == "C1":
        W = H = 101
        g = eng.mapgen(KAT_SEED[0], KAT_SEED[1])
        sp = int(g["agg"][2])
        path = g["fullpath"][sp][: int(g["fullpath_len"][sp])]
        on_path = np.zeros(W * H, bool)
        on_path[path] = True
        cand = int(np.flatnonzero(on_path & ((g["flags"] & game_events.UNBUILDABLE) == 0))[0])
        f2 = g["flags"].copy()
        f2[cand] |= synth.FLAG_BUILDING
        if use_ref:
            eng.L.ref_set_task_threads(min(4, cores))
        lat = []
        h = eng.map_create(W, H, g["elev"], f2, g["wp_tiles"], 50, 50) if use_ref else None
        for _ in range(max(args.steps, 10) + 3):
            t0 = time.perf_counter()
            if use_ref:
                eng.directions_update(W, H, g["elev"], f2, int(g["wp_tiles"][0]), 50, 50)  # the trial of placingBuildingBlocksMonsters (grid copy + one Dijkstra)
                eng.L.ref_waypoints_update(h)  # Waypoints::update on the live objects, nothing copied out
            else:
                eng.directions_update(W, H, g["elev"], f2, g["wp_tiles"][:1], 50, 50)
                eng.waypoints_update(W, H, g["elev"], f2, g["wp_tiles"], 50, 50)
            lat.append(time.perf_counter() - t0)
        if h:
            eng.L.ref_map_destroy(h)
        med = float(np.median(lat[3:]))
        line = {"impl": "reference", "workload": "C1", "metric": "us per placement (trial Dijkstra + Waypoints::update, 101x101, 4 layers)", "value": med * 1e6, "unit": "us",
                "higher_is_better": False, "tile_relaxations_per_s": W * H * NWP / med, "cpu_baseline": {"kind": kind, "cores": min(4, cores) if use_ref else 1, "cpu": cpu_model(), "sample": f"median of {len(lat) - 3} placements, candidate tile {cand}"}}
    else:
        from concurrent.futures import ThreadPoolExecutor

        W = H = args.size
        elev, flags = synth.random_grid(W, H, seed_k=1, wall_pct=30, elev_mod=512)
        srcs = synth.pick_open_tiles(flags, 8, 1)
        thr = min(8, cores)
        one = (lambda s: eng.directions_update(W, H, elev, flags, int(s))) if use_ref else (lambda s: eng.directions_update(W, H, elev, flags, np.array([s], np.uint32)))
        t0 = time.perf_counter()
        with ThreadPoolExecutor(thr) as ex:
            list(ex.map(one, srcs))
        dt = time.perf_counter() - t0
        line = {"impl": "reference", "workload": "C5", "metric": "tile relaxations/sec (8 Directions::update layers on one grid)", "value": W * H * 8 / dt, "unit": "tile relaxations/s",
                "higher_is_better": True, "ms": dt * 1e3, "config": {"grid": [W, H], "layers": 8}, "cpu_baseline": {"kind": kind, "cores": thr, "cpu": cpu_model(), "sample": "one run of 8 concurrent heap Dijkstras"}}
    print(json.dumps(line), flush=True)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    kind, fn = cpu_mapgen_fn()
    cores = host_cores()
    maps = args.maps
    # bounded step: at most ~8 s of CPU work per step
    t0 = time.perf_counter()
    probe = max(2 * cores, 16)
    fn(0, probe, cores)
    per_map = (time.perf_counter() - t0) / probe
    step_maps = int(min(maps, max(probe, 8.0 / per_map)))
    for _ in range(args.warmup):
        fn(0, step_maps, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn(0, step_maps, cores)
    dt = (time.perf_counter() - t0) / args.steps
    value = step_maps * T * NWP / dt
    sample = f"each step = {step_maps} of {maps} maps (seeds k=0..{step_maps - 1}) on {cores} host threads, one map per thread"
    line = {
        "impl": "reference",
        "metric": "tile relaxations/sec (map gen + Waypoints::update on batched 101x101 maps)",
        "value": value,
        "unit": "tile relaxations/s",
        "map_rebuilds_per_s": step_maps / dt,
        "n_gpus": args.gpus,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": dt * 1e3 * maps / step_maps,
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "u32",
        "data": "synthetic",
        "config": workload_config(args, maps),
        "cpu_baseline": {"value": value, "unit": "tile relaxations/s", "cores": cores, "kind": kind, "cpu": cpu_model(), "sample": sample},
        "e2e": {"value": value, "unit": "tile relaxations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------------
# The other BASELINE.json configurations, measured in the same run and reported under "configs" (C2 stays `value`)
# --------------------------------------------------------------------------------------------------------
def hbm_peak():
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    if "hbm_gbs" in peaks:
        return float(peaks["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    return 6650.0, "fallback 6650 GB/s (of fallback)"


def _golden_big():
    import numpy as np

    path = os.path.join(ROOT, "tests", "golden", "big_configs.npz")
    return np.load(path) if os.path.exists(path) else None


def _sum64(t):
    """64-bit wrap-around sum of the uint32 bit patterns held by an int32 device tensor"""
    import torch

    s = 0
    for part in torch.chunk(t.reshape(-1), 16):
        s = (s + int((part.to(torch.int64) & 0xFFFFFFFF).sum().item())) & 0xFFFFFFFFFFFFFFFF
    return s


def _timed_reps(fn, reps, warm, stream, world, dev):
    """device time per repetition (CUDA events on the launching stream, max over ranks)"""
    import torch

    def barrier():
        if world > 1:
            import torch.distributed as dist

            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warm):
        fn()
    ms = []
    for _ in range(reps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            import torch.distributed as dist

            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms.append(float(t.item()))
    return ms


def _big_roofline(ctx, algo_bytes, update_ms, reps, cfg=None):
    from mazetd_b200 import capi

    traffic = None
    try:  # DRAM bytes of the relaxation kernel from the committed ncu capture of this configuration (full size only)
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"k_tile_async_{cfg}_dram_bytes_per_launch") if cfg else None
    except Exception:
        pass

    peak, src = hbm_peak()
    relax_ms, relax_n = ctx.profile_read(capi.Context.K_BIG_RELAX)
    dir_ms, dir_n = ctx.profile_read(capi.Context.K_BIG_DIRECTIONS)
    prep_ms, _ = ctx.profile_read(capi.Context.K_BIG_PREPARE)
    st = ctx.big_stats()
    achieved = algo_bytes / (update_ms * 1e-3) / 1e9
    return {
        "bound": "hbm", "kernel": "k_tile_async" if st["dataflow"] else "k_tile_rounds", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": traffic if st["dataflow"] else None, "traffic_note": "relaxation kernel only, one launch (profiles/traffic.json)" if traffic and st["dataflow"] else None,
        "peak_source": src, "algorithmic_bytes_per_update": algo_bytes,
        "note": "achieved = W*H*(4 + 8*layers) bytes of the WHOLE update / its device time (init + relaxation + next-hop pass), the conservative reading of SURVEY.md 8d",
        "kernel_avg_ms": relax_ms / max(relax_n, 1), "kernel_share_of_update": (relax_ms / max(relax_n, 1)) / update_ms if update_ms > 0 else None,
        "next_hop_pass_avg_ms": dir_ms / max(dir_n, 1), "init_avg_ms": prep_ms / max(reps, 1),
        "tile_visits": st["visits"], "global_rounds": st["rounds"],
    }


def leg_c5(ctx, dev, stream, rank, world, args):
    """C5: 8 Directions::update layers on one 4096^2 grid; layers dealt round-robin to the ranks, no collective"""
    import numpy as np
    import torch

    from mazetd_b200 import shard, synth

    Wb = Hb = 4096
    L = 8
    elev, flags = synth.random_grid_torch(Wb, Hb, seed_k=1, wall_pct=30, elev_mod=512, device=dev)
    srcs = synth.pick_open_tiles_torch(flags, L, 1)
    mine = shard.layer_assignment(rank, world, L)
    ctx.grid_upload_dev(Wb, Hb, elev.data_ptr(), flags.data_ptr())
    n = len(mine)
    dt = torch.empty((max(n, 1), Wb * Hb), dtype=torch.int32, device=dev)
    dd = torch.empty((max(n, 1), Wb * Hb), dtype=torch.int32, device=dev)

    def run():
        if n:
            ctx.layers_update_dev(srcs[mine], dt.data_ptr(), dd.data_ptr())

    ctx.profile_enable(True)
    _timed_reps(run, 0, 2, stream, world, dev)
    for k in (ctx.K_BIG_RELAX, ctx.K_BIG_DIRECTIONS, ctx.K_BIG_PREPARE):
        ctx.profile_read(k)
    reps = 5
    ms = _timed_reps(run, reps, 0, stream, world, dev)
    med = float(np.median(ms))
    out = {
        "workload": "C5: 8 waypoint layers (distance + direction) on one 4096x4096 grid, 30 % walls, weights 255..570", "grid": [Wb, Hb], "layers": L, "n_gpus": world,
        "partition": "layers round-robin over the ranks, grid replicated, no collective" if world > 1 else "one GPU",
        "ms": med, "ms_all": ms, "value": Wb * Hb * L / (med * 1e-3), "unit": "tile relaxations/s",
        "roofline": _big_roofline(ctx, Wb * Hb * (4 * world + 8 * L), med, reps, "C5" if (Wb, L, world) == (4096, 8, 1) else None),
    }
    ctx.profile_enable(False)
    g = _golden_big()
    if g is not None and n:
        ok = all(_sum64(dd[i]) == int(g["C5_dist_sum"][l]) and _sum64(dt[i]) == int(g["C5_tiles_sum"][l]) for i, l in enumerate(mine))
        out["parity"] = "golden-ok" if ok else "GOLDEN-MISMATCH"
    if rank == 0 and world == 1 and not args.no_e2e:
        # e2e through the host ABI: grid H2D + 8 x mtd_directions_update with pinned host outputs
        he = elev.cpu().numpy()
        hf = flags.cpu().numpy().view(np.uint16)
        ht = torch.empty((L, Wb * Hb), dtype=torch.int32, pin_memory=True)
        hd = torch.empty((L, Wb * Hb), dtype=torch.int32, pin_memory=True)

        def e2e():
            ctx.grid_upload(Wb, Hb, he, hf)
            for l in range(L):
                ctx._ck(ctx.L.mtd_directions_update(ctx.h, C.c_void_p(srcs[l:l + 1].ctypes.data), 1, 0xFFFFFFFF, C.c_void_p(ht[l].data_ptr()), C.c_void_p(hd[l].data_ptr())))

        e2e()
        t0 = time.perf_counter()
        e2e()
        dt_s = time.perf_counter() - t0
        out["e2e"] = {"value": Wb * Hb * L / dt_s, "unit": "tile relaxations/s", "ms": dt_s * 1e3, "h2d_bytes": int(he.nbytes + hf.nbytes), "d2h_bytes": int(ht.numel() * 4 * 2),
                      "api": "mtd_grid_upload + 8 x mtd_directions_update (host buffers, pinned outputs)"}
        if g is not None:
            out["e2e"]["parity"] = "golden-ok" if all(_sum64(hd[l]) == int(g["C5_dist_sum"][l]) for l in range(L)) else "GOLDEN-MISMATCH"
    if rank == 0 and world == 1 and not args.no_cpu:
        out["cpu_baseline"] = cpu_c5(elev.cpu().numpy(), flags.cpu().numpy().view(np.uint16), srcs, Wb, Hb)
    return out


def cpu_c5(elev, flags, srcs, Wb, Hb):
    from concurrent.futures import ThreadPoolExecutor

    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib

    use_ref = oracle_lib.Ref.available()
    if not use_ref:
        oracle_lib.build_oracle()
    eng = oracle_lib.Ref() if use_ref else oracle_lib.Oracle()
    thr = min(len(srcs), host_cores())
    one = (lambda s: eng.directions_update(Wb, Hb, elev, flags, int(s))) if use_ref else (lambda s: eng.directions_update(Wb, Hb, elev, flags, np.array([s], np.uint32)))
    t0 = time.perf_counter()
    with ThreadPoolExecutor(thr) as ex:
        list(ex.map(one, srcs))
    dt = time.perf_counter() - t0
    return {"value": Wb * Hb * len(srcs) / dt, "unit": "tile relaxations/s", "ms": dt * 1e3, "cores": thr, "kind": "reference" if use_ref else "port", "cpu": cpu_model(),
            "sample": f"the whole workload once: {len(srcs)} concurrent heap Dijkstras (Directions::update) on {thr} host threads"}


def leg_c4(ctx, dev, stream, rank, world, args):
    """C4: one 16384^2 grid, 16 sources, one layer. One GPU: the whole grid. N > 1: row stripes, peer-coupled."""
    import numpy as np
    import torch

    from mazetd_b200 import shard, striped, synth

    Wb = Hb = args.c4_size
    nsrc = 16
    elev, flags = synth.random_grid_torch(Wb, Hb, seed_k=1, wall_pct=30, elev_mod=512, device=dev)
    srcs = synth.pick_open_tiles_torch(flags, nsrc, 1)
    g = _golden_big() if Wb == 16384 else None
    out = {"workload": f"C4: one {Wb}x{Hb} grid, 30 % walls, weights 255..570, 16 sources, one distance + direction layer", "grid": [Wb, Hb], "sources": nsrc, "n_gpus": world}
    reps = 3
    if world == 1:
        ctx.grid_upload_dev(Wb, Hb, elev.data_ptr(), flags.data_ptr())
        dt = torch.empty(Wb * Hb, dtype=torch.int32, device=dev)
        dd = torch.empty(Wb * Hb, dtype=torch.int32, device=dev)
        run = lambda: ctx.directions_update_dev(srcs, dt.data_ptr(), dd.data_ptr())  # noqa: E731
        ctx.profile_enable(True)
        _timed_reps(run, 0, 1, stream, world, dev)
        for k in (ctx.K_BIG_RELAX, ctx.K_BIG_DIRECTIONS, ctx.K_BIG_PREPARE):
            ctx.profile_read(k)
        ms = _timed_reps(run, reps, 0, stream, world, dev)
        med = float(np.median(ms))
        out.update(partition="one GPU", ms=med, ms_all=ms, value=Wb * Hb / (med * 1e-3), unit="tile relaxations/s", roofline=_big_roofline(ctx, Wb * Hb * 12, med, reps, "C4" if Wb == 16384 else None))
        ctx.profile_enable(False)
        if g is not None:
            out["parity"] = "golden-ok" if _sum64(dd) == int(g["C4_dist_sum"]) and _sum64(dt) == int(g["C4_tiles_sum"]) else "GOLDEN-MISMATCH"
        if not args.no_e2e:
            he = elev.cpu().numpy()
            hf = flags.cpu().numpy().view(np.uint16)
            ht = torch.empty(Wb * Hb, dtype=torch.int32, pin_memory=True)
            hd = torch.empty(Wb * Hb, dtype=torch.int32, pin_memory=True)

            def e2e():
                ctx.grid_upload(Wb, Hb, he, hf)
                ctx._ck(ctx.L.mtd_directions_update(ctx.h, C.c_void_p(srcs.ctypes.data), nsrc, 0xFFFFFFFF, C.c_void_p(ht.data_ptr()), C.c_void_p(hd.data_ptr())))

            e2e()
            t0 = time.perf_counter()
            e2e()
            dt_s = time.perf_counter() - t0
            out["e2e"] = {"value": Wb * Hb / dt_s, "unit": "tile relaxations/s", "ms": dt_s * 1e3, "h2d_bytes": int(he.nbytes + hf.nbytes), "d2h_bytes": int(ht.numel() * 8),
                          "api": "mtd_grid_upload + mtd_directions_update (host buffers, pinned outputs)"}
            if g is not None:
                out["e2e"]["parity"] = "golden-ok" if _sum64(hd) == int(g["C4_dist_sum"]) and _sum64(ht) == int(g["C4_tiles_sum"]) else "GOLDEN-MISMATCH"
            del ht, hd
        if not args.no_cpu:
            out["cpu_baseline"] = cpu_c4_sample(elev, flags, Wb, Hb)
        ctx.grid_upload(64, 64, np.zeros(4096, np.int16), np.zeros(4096, np.uint16))  # release the derived data of the big grid
    else:
        y0, y1 = shard.stripe_rows(rank, world, Hb)
        ya, yb = shard.stripe_halo_rows(y0, y1, Hb)
        e = elev.view(Hb, Wb)[ya:yb].contiguous()
        f = flags.view(Hb, Wb)[ya:yb].contiguous()
        del elev, flags
        st = striped.GpuStripe(dev.index, Wb, Hb, y0, y1, e, f, srcs, stream=stream.cuda_stream)
        striped.connect_peers(st, rank, world)
        rounds = [0]

        def run():
            rounds[0] = striped.solve_peer(st)[2]

        _timed_reps(run, 0, 1, stream, world, dev)
        ms = _timed_reps(run, reps, 0, stream, world, dev)
        med = float(np.median(ms))
        peak, src = hbm_peak()
        ach = Wb * Hb * 12 / (med * 1e-3) / 1e9
        bs = st.ctx.big_stats()
        out.update(partition=f"{world} row stripes (cuts on 64-row tile boundaries), boundary rows and keys stored into the neighbour's memory by the relaxation kernel (CUDA IPC over NVLink)",
                   ms=med, ms_all=ms, value=Wb * Hb / (med * 1e-3), unit="tile relaxations/s", global_rounds=int(rounds[0]), tile_visits_rank0=bs["visits"],
                   roofline={"bound": "hbm", "kernel": ("k_tile_async" if bs["dataflow"] else "k_tile_rounds") + " (peer-coupled)", "achieved": ach, "peak": peak * world, "unit": "GB/s", "frac": ach / (peak * world), "traffic": None, "peak_source": src + f" x {world} GPUs"})
        if g is not None:
            import torch.distributed as dist

            sums = torch.tensor([_sum64(st.dist) & 0x7FFFFFFFFFFFFFFF, _sum64(st.tiles) & 0x7FFFFFFFFFFFFFFF], dtype=torch.int64, device=dev)
            parts = [torch.zeros_like(sums) for _ in range(world)]
            dist.all_gather(parts, sums)
            tot_d = sum(int(p[0].item()) for p in parts) & 0xFFFFFFFFFFFFFFFF
            tot_t = sum(int(p[1].item()) for p in parts) & 0xFFFFFFFFFFFFFFFF
            # per-rank sums were masked to 63 bits: compare modulo 2^63
            ok_d = (tot_d & 0x7FFFFFFFFFFFFFFF) == (int(g["C4_dist_sum"]) & 0x7FFFFFFFFFFFFFFF)
            ok_t = (tot_t & 0x7FFFFFFFFFFFFFFF) == (int(g["C4_tiles_sum"]) & 0x7FFFFFFFFFFFFFFF)
            ok = ok_d and ok_t
            out["parity_detail"] = {"distances": bool(ok_d), "next_hops": bool(ok_t)}
            out["parity"] = "golden-ok" if ok else "GOLDEN-MISMATCH"
            out["peer_parity"] = "PEER-OK" if ok else "PEER-MISMATCH"
        st.close()
    return out


def cpu_c4_sample(elev_t, flags_t, Wb, Hb):
    """bounded CPU sample of C4: the top-left 4096 x 4096 block of the same grid (1/16 of the tiles) with one source,
    i.e. C4's source density, one heap Dijkstra on one host thread (a Directions::update is single-threaded)"""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    from mazetd_b200 import synth

    S = min(4096, Wb)
    e = elev_t.view(Hb, Wb)[:S, :S].contiguous().cpu().numpy().reshape(-1)
    f = flags_t.view(Hb, Wb)[:S, :S].contiguous().cpu().numpy().view(np.uint16).reshape(-1)
    src = int(synth.pick_open_tiles(f, 1, 1)[0])
    use_ref = oracle_lib.Ref.available()
    if not use_ref:
        oracle_lib.build_oracle()
    eng = oracle_lib.Ref() if use_ref else oracle_lib.Oracle()
    t0 = time.perf_counter()
    if use_ref:
        eng.directions_update(S, S, e, f, src)
    else:
        eng.directions_update(S, S, e, f, np.array([src], np.uint32))
    dt = time.perf_counter() - t0
    return {"value": S * S / dt, "unit": "tile relaxations/s", "ms_sample": dt * 1e3, "ms_extrapolated_full_grid": dt * 1e3 * (Wb * Hb) / (S * S), "cores": 1, "kind": "reference" if use_ref else "port", "cpu": cpu_model(),
            "sample": f"top-left {S}x{S} block of the grid (1/{(Wb * Hb) // (S * S)} of the tiles), one source, one Directions::update on one host thread"}


def leg_c1_c3(ctx, args, cpu):
    """C1: the default generated map, trial + flag delta + full rebuild after the scripted placement, through the host ABI.
    C3: 10 000 place/remove events on it with 1 000 000 monsters ticking between events (bounded CPU sample beside it)."""
    import numpy as np

    from mazetd_b200 import game_events, synth

    out = {}
    s1, s2 = KAT_SEED
    g = ctx.mapgen_batch(s1, s2, 0, 1, cap=4096)
    elev, flags, wp = g["elev"][0], g["flags"][0], g["wp_tiles"][0]
    sp = int(g["agg"][0][2])
    path = g["fullpath"][0][sp][: int(g["fullpath_len"][0][sp])]
    ctx.grid_upload(W, H, elev, flags, 50, 50)
    ctx.waypoints_update_dev(wp)
    on_path = np.zeros(T, bool)
    on_path[path] = True
    cand = int(np.flatnonzero(on_path & ((flags & game_events.UNBUILDABLE) == 0))[0])
    lat = []
    for _ in range(120):
        t0 = time.perf_counter()
        ctx.placing_blocks(cand)                                  # P5: the trial Dijkstra
        ctx.grid_set_flags(cand, int(flags[cand]) | synth.FLAG_BUILDING)  # P6: flag delta
        ctx.waypoints_update(wp, cap=4096)                        # P2: 4 layers + walks + aggregates, everything copied to the host
        lat.append(time.perf_counter() - t0)
        ctx.grid_set_flags(cand, int(flags[cand]))  # untimed: the tower is sold again and the layers follow (clearStructure path)
        ctx.waypoints_update_dev(wp)
        ctx.synchronize()
    med = float(np.median(lat[20:]))
    res = []
    for rep in range(120):
        ctx.grid_set_flags(cand, (int(flags[cand]) | synth.FLAG_BUILDING) if rep % 2 == 0 else int(flags[cand]))  # place / sell alternately
        t0 = time.perf_counter()
        ctx.waypoints_update_dev(wp)
        ctx.synchronize()
        res.append(time.perf_counter() - t0)
    inc = ctx.big_stats()
    ctx.debug_set(ctx.DEBUG_INCREMENTAL, 0)
    full = []
    for rep in range(60):
        ctx.grid_set_flags(cand, (int(flags[cand]) | synth.FLAG_BUILDING) if rep % 2 == 0 else int(flags[cand]))
        t0 = time.perf_counter()
        ctx.waypoints_update_dev(wp)
        ctx.synchronize()
        full.append(time.perf_counter() - t0)
    ctx.debug_set(ctx.DEBUG_INCREMENTAL, 1)
    out["C1"] = {"workload": "C1: default generated map (seed k=0), scripted tower placement: trial + flag delta + Waypoints::update, 101x101, 4 layers", "metric": "us per placement", "value": med * 1e6, "unit": "us",
                 "higher_is_better": False, "tile_relaxations_per_s": T * NWP / med, "e2e": {"value": med * 1e6, "unit": "us", "d2h_bytes": T * NWP * 8 + 4 * 4096 * NWP, "api": "mtd_placing_blocks + mtd_grid_set_flags + mtd_waypoints_update (host outputs), wall clock"},
                 "resident_us": float(np.median(res[20:])) * 1e6, "resident_from_scratch_us": float(np.median(full[10:])) * 1e6, "incremental_updates": inc["rounds"], "incremental_trials": inc["visits"],
                 "candidate_tile": cand, "roofline": None, "roofline_note": "latency-bound by design (10 201 tiles in one SM's shared memory): reported as latency, SURVEY.md 8d"}
    if cpu:
        out["C1"]["cpu_baseline"] = cpu_c1(elev, flags, wp, cand)

    nevents, nmon = args.c3_events, args.c3_monsters
    ctx.grid_upload(W, H, elev, flags, 50, 50)
    ctx.waypoints_update_dev(wp)
    mon = game_events.initial_monsters(nmon, path, wp, ctx.waypoints_find, speed_range=(0.03, 0.06))
    game = game_events.GpuGame(ctx, W, H, elev, flags, wp)
    ctx.monsters_upload(mon)
    l0 = ctx.launches
    ctx.synchronize()
    t0 = time.perf_counter()
    stats = game_events.run_events(game, nevents)
    _, alive = ctx.monsters_tick(nevents + 1)
    wall = time.perf_counter() - t0
    peak, src = hbm_peak()
    ach = 28.0 * nmon * nevents / wall / 1e9
    out["C3"] = {"workload": f"C3: {nevents} random tower place/remove events on the default map, {nmon} monsters steering along the flow field, one tick per event", "metric": "monster ticks/s", "value": nmon * nevents / wall, "unit": "monster ticks/s",
                 "events_per_s": nevents / wall, "wall_s": wall, "rebuilds": stats["placed"] + stats["cleared"], "trials": stats["placed"] + stats["rejected_trial"], "stats": stats, "alive": int(alive), "gpu_launches": ctx.launches - l0,
                 "roofline": {"bound": "hbm", "kernel": "k_monsters_tick", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None, "peak_source": src,
                              "note": "28 B per monster per tick (SURVEY.md 8d) over the WALL time of the whole script incl. rebuilds and trials (host-driven event loop)"},
                 "e2e": {"value": nmon * nevents / wall, "unit": "monster ticks/s", "note": "the script itself is the host-facing path: every event is a synchronous C-ABI call sequence"}}
    if cpu:
        out["C3"]["cpu_baseline"] = cpu_c3(elev, flags, wp, path)
    return out


def leg_c3_sharded(ctx, dev, rank, world, args):
    """C3 at N > 1: every rank holds a replica of the 101x101 map and 1/N of the monsters; per trial one
    all-reduce of the verdict, per accepted event one broadcast of the 326 KB of layers from rank 0"""
    import numpy as np
    import torch
    import torch.distributed as dist

    from mazetd_b200 import game_events, sharded_game

    s1, s2 = KAT_SEED
    g = ctx.mapgen_batch(s1, s2, 0, 1, cap=4096)
    elev, flags, wp = g["elev"][0], g["flags"][0], g["wp_tiles"][0]
    sp = int(g["agg"][0][2])
    path = g["fullpath"][0][sp][: int(g["fullpath_len"][0][sp])]
    ctx.grid_upload(W, H, elev, flags, 50, 50)
    ctx.waypoints_update_dev(wp)
    nevents, nmon = min(args.c3_events, 2000), args.c3_monsters
    mon = game_events.initial_monsters(nmon, path, wp, ctx.waypoints_find, speed_range=(0.03, 0.06))
    mine = sharded_game.shard_monsters(mon, rank, world)
    out = {}
    for mode in ("broadcast", "replicate"):
        local = sharded_game.GpuShardBackend(ctx, W, H, elev, flags, wp, dev)
        ctx.monsters_upload(mine)
        game = sharded_game.ShardedGame(local, rank, world, mode=mode, device=dev)
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        stats = game_events.run_events(game, nevents)
        _, alive = ctx.monsters_tick(nevents + 1)
        tot = torch.tensor([int(alive)], dtype=torch.int64, device=dev)
        dist.all_reduce(tot)
        wall = time.perf_counter() - t0
        t = torch.tensor([wall], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall = float(t.item())
        out[mode] = {"wall_s": wall, "value": nmon * nevents / wall, "unit": "monster ticks/s", "events_per_s": nevents / wall, "collectives_per_rank": game.collectives, "alive_total": int(tot.item()),
                     "stats": stats}
    return {"workload": f"C3 sharded: {nevents} place/remove events on the default map (replica per GPU), {nmon} monsters split by index over {world} GPUs", "n_gpus": world,
            "partition": "monsters by index; trial verdict all-reduce(MAX) per trial; layers broadcast from rank 0 per accepted event (mode broadcast) or rebuilt on every rank (mode replicate)",
            "value": out["broadcast"]["value"], "unit": "monster ticks/s", "modes": out,
            "note": "latency-bound on the per-event rebuild and trial (one SM each); sharding the monsters shortens only the tick"}


def cpu_c1(elev, flags, wp, cand):
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    from mazetd_b200 import synth

    use_ref = oracle_lib.Ref.available()
    if not use_ref:
        oracle_lib.build_oracle()
    eng = oracle_lib.Ref() if use_ref else oracle_lib.Oracle()
    cores = host_cores()
    f2 = flags.copy()
    f2[cand] |= synth.FLAG_BUILDING
    if use_ref:
        eng.L.ref_set_task_threads(min(4, cores))
    h = eng.map_create(W, H, elev, f2, wp, 50, 50) if use_ref else None
    lat = []
    for _ in range(60):
        t0 = time.perf_counter()
        if use_ref:
            eng.directions_update(W, H, elev, f2, int(wp[0]), 50, 50)
            eng.L.ref_waypoints_update(h)
        else:
            eng.directions_update(W, H, elev, f2, wp[:1], 50, 50)
            eng.waypoints_update(W, H, elev, f2, wp, 50, 50)
        lat.append(time.perf_counter() - t0)
    if h:
        eng.L.ref_map_destroy(h)
    med = float(np.median(lat[5:]))
    return {"value": med * 1e6, "unit": "us", "cores": min(4, cores) if use_ref else 1, "kind": "reference" if use_ref else "port", "cpu": cpu_model(), "sample": "median of 55 placements (trial Dijkstra + Waypoints::update on the live objects)"}


def cpu_c3(elev, flags, wp, path, nevents=300, nmon=100000):
    """bounded sample: the same script, fewer events and monsters, single host thread (the reference's ECS visitor and its
    placement code run on the control thread)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    from mazetd_b200 import game_events

    oracle_lib.build_oracle()
    orc = oracle_lib.Oracle()
    w = orc.waypoints_update(W, H, elev, flags, wp, 50, 50)
    view = orc.waypoints_view(W, H, elev, flags, wp, w["tiles"], w["dist"], 50, 50)
    import numpy as np

    find = lambda pos, bits: tuple(np.array(x, np.uint32) for x in zip(*[orc.find(view, int(p), int(b)) for p, b in zip(pos, bits)]))  # noqa: E731
    mon = game_events.initial_monsters(2000, path, wp, find, speed_range=(0.03, 0.06))
    reps = nmon // 2000
    mon = {k: np.tile(v, reps) for k, v in mon.items()}
    og = oracle_lib.OracleGame(orc, W, H, elev, flags, wp, mon)
    t0 = time.perf_counter()
    stats = game_events.run_events(og, nevents)
    dt = time.perf_counter() - t0
    n = len(mon["tile"])
    return {"value": n * nevents / dt, "unit": "monster ticks/s", "events_per_s": nevents / dt, "cores": 1, "kind": "port", "cpu": cpu_model(),
            "sample": f"{nevents} events x {n} monsters ({dt:.1f} s): oracle restatement of monsters.cpp:55-99 + Waypoints::update / trial per event, one thread; rebuilds {stats['placed'] + stats['cleared']}"}


def workload_config(args, maps):
    return {
        "workload": f"C2: batch of {maps} procedurally generated {W}x{H} maps per GPU (seeds (s1+k, s2), k = rank*{maps}..), map gen + {NWP} distance/direction layers + full paths + aggregates",
        "maps_per_gpu": maps,
        "grid": [W, H],
        "layers": NWP,
        "seed": list(KAT_SEED),
        "fullpath_cap": FULLPATH_CAP,
        "l2_policy": "per-step outputs (1.5 GB) exceed the 126 MB L2; no input is re-read across steps",
    }


def run_ours(args):
    import numpy as np
    import torch

    from mazetd_b200 import capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libmazetd_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = None
    if world > 1:
        import torch.distributed as dist
        from mazetd_b200 import numa as numa_mod

        # before any pinned allocation: this rank's host buffers go to the NUMA node its GPU hangs off
        numa = numa_mod.bind_to_gpu_node(local)
        dist.init_process_group("nccl", device_id=dev)
    # a real (non-default) stream: the library treats a NULL handle as "create your own", and torch's events must
    # be recorded on the stream the kernels are launched on
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ctx = capi.Context(local, stream=stream.cuda_stream)
    if args.gw:
        ctx.debug_set(ctx.DEBUG_GROUP_WARPS, args.gw)  # tuning probe only; the default is the shipped configuration
    if args.walker:
        ctx.debug_set(ctx.DEBUG_WALKER, args.walker)
    maps = args.maps
    first = rank * maps

    # ---- device-resident outputs (everything Grid + Waypoints hold) ----
    d = dict(
        elev=torch.empty((maps, T), dtype=torch.int16, device=dev),
        flags=torch.empty((maps, T), dtype=torch.int16, device=dev),
        wp_tiles=torch.empty((maps, NWP), dtype=torch.int32, device=dev),
        tiles=torch.empty((maps, NWP, T), dtype=torch.int32, device=dev),
        dist=torch.empty((maps, NWP, T), dtype=torch.int32, device=dev),
        fullpath=torch.empty((maps, NWP, FULLPATH_CAP), dtype=torch.int32, device=dev),
        fullpath_len=torch.empty((maps, NWP), dtype=torch.int32, device=dev),
        full_distance=torch.empty((maps, NWP), dtype=torch.int32, device=dev),
        agg=torch.empty((maps, 3), dtype=torch.int32, device=dev),
    )
    ptrs = {k: v.data_ptr() for k, v in d.items()}

    def step():
        ctx.mapgen_batch_dev(KAT_SEED[0], KAT_SEED[1], first, maps, ptrs, W, H, 50, 50, 45, NWP, FULLPATH_CAP)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    ctx.synchronize()  # raises on any deferred kernel status

    # ---- timed region: exactly K steps, CUDA events on the launching stream ----
    sampler = ClockSampler(local)
    ctx.profile_enable(True)
    ctx.profile_read(capi.Context.K_WAYPOINTS_SMALL)
    ctx.profile_read(capi.Context.K_MAPGEN_GRID)
    launches0 = ctx.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.start()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms_total = e0.elapsed_time(e1)
    launches = ctx.launches - launches0
    wp_ms, wp_n = ctx.profile_read(capi.Context.K_WAYPOINTS_SMALL)
    mg_ms, mg_n = ctx.profile_read(capi.Context.K_MAPGEN_GRID)
    ctx.profile_enable(False)
    ctx.synchronize()
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    relax_per_step = maps * T * NWP * world
    value = relax_per_step / (ms_step * 1e-3)

    # ---- roofline of the dominant kernel (k_waypoints_small): algorithmic bytes per DESIGN.md ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (of fallback)"
    bytes_per_map = T * (4 + 8 * NWP)  # elev+flags once (read, or written when the generator is fused in), dist+tiles written per layer
    wp_avg_ms = wp_ms / max(wp_n, 1)
    achieved = maps * bytes_per_map / (wp_avg_ms * 1e-3) / 1e9 if wp_avg_ms > 0 else 0.0
    roofline = {
        "bound": "hbm",
        "kernel": "k_waypoints_small" + (" (generator fused in)" if mg_n == 0 else ""),
        "achieved": achieved,
        "peak": peak,
        "unit": "GB/s",
        "frac": achieved / peak,
        "traffic": None,
        "peak_source": peak_src,
        "algorithmic_bytes_per_launch": maps * bytes_per_map,
        "avg_launch_ms": wp_avg_ms,
        "launches_timed": wp_n,
        "share_of_step": wp_ms / (ms_total if ms_total > 0 else 1.0),
        "other_kernels": {"k_mapgen_grid_avg_ms": mg_ms / mg_n} if mg_n else {},
    }
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get("k_waypoints_small_dram_bytes_per_launch")
        except Exception:
            pass

    # ---- e2e: the host-buffer C-ABI call a drop-in user makes; pinned host outputs, D2H inside the timed region ----
    e2e = None
    if not args.no_e2e:
        hp = dict(
            elev=torch.empty((maps, T), dtype=torch.int16, pin_memory=True),
            flags=torch.empty((maps, T), dtype=torch.int16, pin_memory=True),
            wp_tiles=torch.empty((maps, NWP), dtype=torch.int32, pin_memory=True),
            tiles=torch.empty((maps, NWP, T), dtype=torch.int32, pin_memory=True),
            dist=torch.empty((maps, NWP, T), dtype=torch.int32, pin_memory=True),
            fullpath=torch.empty((maps, NWP, FULLPATH_CAP), dtype=torch.int32, pin_memory=True),
            fullpath_len=torch.empty((maps, NWP), dtype=torch.int32, pin_memory=True),
            full_distance=torch.empty((maps, NWP), dtype=torch.int32, pin_memory=True),
            agg=torch.empty((maps, 3), dtype=torch.int32, pin_memory=True),
        )
        hptrs = {k: v.data_ptr() for k, v in hp.items()}
        d2h = sum(v.numel() * v.element_size() for v in hp.values())
        p = capi.MapgenParams(W, H, 50, 50, 45, NWP, FULLPATH_CAP)

        def e2e_step():
            rc = ctx.L.mtd_mapgen_batch(ctx.h, C.byref(p), KAT_SEED[0], KAT_SEED[1], first, maps, *[C.c_void_p(hptrs[k]) for k in ("elev", "flags", "wp_tiles", "tiles", "dist", "fullpath", "fullpath_len", "full_distance", "agg")])
            ctx._ck(rc)

        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        # spot-check: the host copy of the aggregates must equal the device-resident run's
        assert np.array_equal(hp["agg"].numpy(), d["agg"].cpu().numpy()), "e2e host outputs differ from the device-resident run"
        e2e = {
            "value": relax_per_step / dt,
            "unit": "tile relaxations/s",
            "map_rebuilds_per_s": maps * world / dt,
            "ms_per_step": dt * 1e3,
            "steps": e2e_steps,
            "h2d_bytes_per_step": 64,  # seeds + mtd_mapgen_params; the generator has no bulk input
            "d2h_bytes_per_step": d2h,
            "api": "mtd_mapgen_batch (include/mazetd_b200.h), pinned host outputs, synchronous",
            "host_numa": numa,  # N > 1: rank 0's process was bound to its GPU's NUMA node before the pinned buffers were allocated
        }

    # ---- the same batch in STRONG form: 4096 maps in total, a contiguous block of map indices per rank (BASELINE.json
    #      configs[1]: "sharded per GPU at 2/4/8"); the weak form above keeps 4096 maps on every GPU ----
    strong = None
    if world > 1:
        from mazetd_b200 import shard

        s_first, s_last = shard.map_range(rank, world, maps)
        s_n = s_last - s_first

        def step_strong():
            ctx.mapgen_batch_dev(KAT_SEED[0], KAT_SEED[1], s_first, s_n, ptrs, W, H, 50, 50, 45, NWP, FULLPATH_CAP)

        for _ in range(3):
            step_strong()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for _ in range(args.steps):
            step_strong()
        f1.record(stream)
        barrier()
        ts = torch.tensor([f0.elapsed_time(f1)], dtype=torch.float64, device=dev)
        dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        ms_strong = float(ts.item()) / args.steps
        strong = {"scaling": "strong", "maps_total": maps, "maps_per_gpu": s_n, "ms_per_step": ms_strong, "value": maps * T * NWP / (ms_strong * 1e-3), "unit": "tile relaxations/s",
                  "map_rebuilds_per_s": maps / (ms_strong * 1e-3)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = time_cpu_sample()

    # ---- the other BASELINE.json configurations, same run (C2 above stays the headline `value`) ----
    del d
    torch.cuda.empty_cache()
    configs = {}
    line_so_far = {}

    def bail_out():
        # a leg that hangs (a rank died inside a collective) must not take the headline measurement with it
        if rank == 0 and line_so_far:
            line_so_far["configs"] = dict(configs, error="the remaining legs were cut off after the time limit")
            print(json.dumps(line_so_far), flush=True)
        os._exit(0 if line_so_far or rank != 0 else 1)

    watchdog = threading.Timer(args.legs_timeout, bail_out)
    watchdog.daemon = True
    if rank == 0:
        line_so_far.update({
            "metric": "tile relaxations/sec (map gen + Waypoints::update on batched 101x101 maps)", "value": value, "unit": "tile relaxations/s",
            "map_rebuilds_per_s": maps * world / (ms_step * 1e-3), "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic", "config": workload_config(args, maps),
            "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "strong": strong,
        })
    watchdog.start()
    if not args.no_configs:
        for name, fn in (("C5", lambda: leg_c5(ctx, dev, stream, rank, world, args)), ("C4", lambda: leg_c4(ctx, dev, stream, rank, world, args))):
            try:
                configs[name] = fn()
            except Exception as e:  # noqa: BLE001  (a failing leg must not take the headline line with it)
                configs[name] = {"error": f"{type(e).__name__}: {e}"[:400]}
            torch.cuda.empty_cache()
        if world == 1:
            try:
                configs.update(leg_c1_c3(ctx, args, cpu=not args.no_cpu))
            except Exception as e:  # noqa: BLE001
                configs["C1_C3"] = {"error": f"{type(e).__name__}: {e}"[:400]}
        else:
            try:
                configs["C3"] = leg_c3_sharded(ctx, dev, rank, world, args)
            except Exception as e:  # noqa: BLE001
                configs["C3"] = {"error": f"{type(e).__name__}: {e}"[:400]}

    watchdog.cancel()
    if rank == 0:
        line = {
            "metric": "tile relaxations/sec (map gen + Waypoints::update on batched 101x101 maps)",
            "value": value,
            "unit": "tile relaxations/s",
            "map_rebuilds_per_s": maps * world / (ms_step * 1e-3),
            "n_gpus": world,
            "steps": args.steps,
            "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step,
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": None,
            "dtype": "u32",
            "data": "synthetic",
            "config": workload_config(args, maps),
            "roofline": roofline,
            "cpu_baseline": cpu,
            "clocks": clocks,
            "e2e": e2e,
            "gpu_launches": launches,
            "strong": strong,
            "configs": configs,
        }
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--maps", type=int, default=4096, help="maps per GPU per step")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1/C3/C4/C5 legs (C2 headline only)")
    ap.add_argument("--legs-timeout", type=float, default=900.0, help="seconds the C1/C3/C4/C5 legs may take before the line is printed without them")
    ap.add_argument("--c4-size", type=int, default=16384)
    ap.add_argument("--c3-events", type=int, default=10000)
    ap.add_argument("--c3-monsters", type=int, default=1000000)
    ap.add_argument("--workload", default="C2", choices=["C2", "C1", "C5"], help="reference arm only: CPU figures for the non-headline configurations")
    ap.add_argument("--size", type=int, default=4096, help="grid side of --workload C5")
    ap.add_argument("--gw", type=int, default=0, help="tuning probe: warps per layer in the one-CTA kernel (1, 2, 4)")
    ap.add_argument("--walker", type=int, default=0, help="tuning probe: 1 = workers walk the paths, 2 = walker warp")
    args = ap.parse_args()
    if args.impl == "reference" and args.workload != "C2":
        if int(os.environ.get("RANK", "0")) == 0:
            run_reference_other(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        world = int(os.environ.get("WORLD_SIZE", "1"))
        if world == 1 and args.gpus > 1:
            # convenience: re-exec under torchrun
            import subprocess

            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
            raise SystemExit(subprocess.call(cmd))
        run_ours(args)


if __name__ == "__main__":
    main()
