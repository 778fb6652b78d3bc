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
    if world > 1:
        import torch.distributed as dist

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
    bytes_per_map = T * (4 + 8 * NWP)  # read elev+flags once, write dist+tiles per layer
    wp_avg_ms = wp_ms / max(wp_n, 1)
    achieved = maps * bytes_per_map / (wp_avg_ms * 1e-3) / 1e9 if wp_avg_ms > 0 else 0.0
    roofline = {
        "bound": "hbm",
        "kernel": "k_waypoints_small",
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
        "other_kernels": {"k_mapgen_grid_avg_ms": mg_ms / max(mg_n, 1)},
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
        }

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = time_cpu_sample()

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
