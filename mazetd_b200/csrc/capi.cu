// capi.cu -- C ABI of libmazetd_b200.so (include/mazetd_b200.h): context, device buffers, kernel launches.
// Single translation unit: all kernels are header-only templates included below.
//
// Build (see mazetd_b200/build.py):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -shared -Xcompiler -fPIC
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <climits>
#include <map>
#include <mutex>
#include <new>
#include <utility>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/mazetd_b200.h"
#include "big_grid.cuh"
#include "common.cuh"
#include "monsters.cuh"
#include "noise.cuh"
#include "small_maps.cuh"
#include "sssp_warp.cuh"
#include "tile_async.cuh"
#include "towers.cuh"

using namespace mtd;

namespace
{
	constexpr size_t MAX_SMEM = 227 * 1024 - 2560; // dynamic budget: the opt-in limit is 227 KB per CTA including static __shared__ (k_waypoints_small: 2.2 KB)
	constexpr int BIG_TS = 32;   // tile side of the large-grid path = warp width (tile_warp.cuh)
	constexpr int BIG_WARPS = 4; // warps per CTA, each working on its own (layer, tile) item

	template<class T>
	struct DevBuf
	{
		T *p = nullptr;
		size_t cap = 0; // elements
		cudaError_t reserve(size_t n)
		{
			if (n <= cap)
				return cudaSuccess;
			if (p)
				cudaFree(p);
			p = nullptr;
			cap = 0;
			cudaError_t e = cudaMalloc((void **)&p, n * sizeof(T));
			if (e == cudaSuccess)
				cap = n;
			return e;
		}
		void release()
		{
			if (p)
				cudaFree(p);
			p = nullptr;
			cap = 0;
		}
	};

	struct MapgenStage
	{
		DevBuf<int16_t> elev;
		DevBuf<uint16_t> flags;
		DevBuf<uint32_t> wp, tiles, dist, fp, fplen, fdist, agg;
		void release()
		{
			elev.release();
			flags.release();
			wp.release();
			tiles.release();
			dist.release();
			fp.release();
			fplen.release();
			fdist.release();
			agg.release();
		}
	};

	struct HostFlags
	{
		int status;            // OR of ST_* bits written by kernels (copied back on demand)
		uint32_t next_count;   // (unused)
		uint32_t round_out[2]; // written by k_tile_rounds: items queued for the next round, index of the next round
		int trial_blocked;
		uint32_t alive;
		unsigned long long damage;
		uint32_t spawn[3];
		uint32_t big_items;     // (layer, tile) visits of the last tiled relaxation (statistics)
		uint32_t stage_blocked; // pinned staging word of the trial grid's extra blocked tile
		uint32_t stripe_improved; // halo cells lowered by the last mtd_stripe_exchange_dev
		uint32_t async_ctl[4];    // ACTL_* words of the last k_tile_async run
	};
}

// one flag word (and its blocking byte on the tiled path) set from kernel arguments: mtd_grid_set_flags
__global__ void k_poke_flags(uint16_t *flag, uint16_t value, uint8_t *blocked, uint8_t b)
{
	*flag = value;
	if (blocked)
		*blocked = b;
}

struct mtd_ctx
{
	int device = 0;
	int sms = 148;
	cudaStream_t stream = nullptr;
	bool own_stream = false;
	char err[512] = { 0 };
	uint64_t launches = 0;
	std::mutex mu;

	// grid
	int W = 0, H = 0, ox = 0, oy = 0;
	size_t T = 0;
	bool have_grid = false;
	bool small = false;
	DevBuf<int16_t> elev;
	DevBuf<uint16_t> flags;
	// big path derived data
	DevBuf<uint32_t> wpack;
	DevBuf<uint8_t> blocked;
	DevBuf<uint32_t> list0, list1, list2, stamp, tilekey, counters; // counters: count[3], cursor[3], total_active, improved
	DevBuf<uint32_t> sources;
	int tilesX = 0, tilesY = 0;
	uint32_t last_rounds = 0;
	DevBuf<unsigned long long> a_inbox;
	DevBuf<uint32_t> a_tail, a_ctamin, a_ctl, a_fly;
	int async_workers = ASYNC_WORKERS; // worker warps per SM in use
	int async_late_looks = 1;          // a visit re-reads its apron once after converging
	int async_cont = 0;                // 1: revisits continue on the frame of the worker that visited the tile last (measured: no gain, r03x)
	int async_nb_rule = 0, async_depth = 2; // measured (profiles/r02g_*): the neighbour rule halves the visits but not the time; two mailboxes: 14.5 -> 13.0 ms on C4
	uint32_t async_nb_margin = 0; // dataflow scheduler of k_tile_async: inbox rings, ring tails, per-CTA smallest keys, control words
	bool big_async = true;         // large grids: k_tile_async (no global rounds); false = k_tile_rounds (A/B and multi-stripe contexts)
	uint32_t async_delta = 65536;  // priority window of the dataflow scheduler (label units); swept 8k..256k on C4 / C5 (profiles/r02c_*, r02g_*)
	uint32_t last_visits = 0;
	// incremental re-pathing of the one-CTA path: the resident layers are valid for the grid as it was before `inc_dirty` blocking
	// flips (inc_dirty == 1: the tile is inc_tile); h_blocking mirrors the blocking bits of a grid uploaded from the host
	bool inc_enabled = true, inc_layers_ok = false;
	int inc_dirty = 0, inc_n = 0;
	uint32_t inc_tile = MTD_M, inc_wp[MTD_MAX_WAYPOINTS] = { 0 };
	std::vector<uint8_t> h_blocking;
	std::vector<int16_t> h_elev; // host mirror of the elevations of a small grid uploaded from the host: a re-upload of the same grid is a flag delta
	uint64_t inc_updates = 0, inc_trials = 0;
	bool last_async = false, last_tma = false;
	bool big_tma = true;           // dataflow scheduler: stage frames with bulk tensor copies (TMA) when the width allows it
	uint32_t tile_delta = 98304; // visit window of the tile-level priority (label units); swept 8k..inf on 4096^2 and 16384^2 (weights 255..570): flat optimum 64k..256k
	// row stripe of a larger grid (multi-GPU, one rank per stripe): rows [sy0, sy1) of a W x sH grid are this context's grid
	bool stripe = false;
	int sy0 = 0, sy1 = 0, sH = 0, stripe_round = 0, stripe_nsrc = 0, stripe_budget = 0;
	uint32_t stripe_pending = 0;
	uint32_t *stripe_dist = nullptr;                 // caller-owned [sy1 - sy0][W]
	DevBuf<uint32_t> halo_up, halo_down, up_wpack;   // local copies of the neighbour stripes' boundary rows; wpack of row sy0-1
	// peer-coupled stripes: one block per rank that the neighbour ranks write into (layout: peer_layout())
	DevBuf<uint32_t> peer_blk;
	uint32_t *peers[PEER_MAX_WORLD] = { nullptr }; // every rank's block (mine included), mapped here
	uint32_t *peer_above = nullptr, *peer_below = nullptr, *peer_root = nullptr; // = peers[rank - 1], peers[rank + 1], peers[0]
	bool peer_ipc = false;      // mapped with cudaIpcOpenMemHandle (to be closed)
	bool peer_linked = false, peer_on = false;
	int peer_rank = 0, peer_world = 1, peer_ctas = 0;
	DevBuf<uint8_t> up_blocked;

	// staging of mtd_mapgen_batch (host outputs): two buffer sets, compute overlaps the drain of the previous chunk
	MapgenStage mg[2];
	cudaStream_t copy_stream = nullptr;
	cudaEvent_t mg_done[2] = { nullptr, nullptr }, mg_copied[2] = { nullptr, nullptr };

	// resident layers
	int nlayers = 0;
	DevBuf<uint32_t> ltiles, ldist, lwp, dtab;
	DevBuf<uint32_t> fp_buf, fp_len, fp_dist, agg;
	uint32_t fp_cap = 0;
	DevBuf<uint32_t> trial_dist;
	DevBuf<uint32_t> map_counter;
	int debug_qseg_cap = 0;
	int debug_gw = SMALL_GROUP_WARPS;
	bool fused_mapgen = false; // mtd_mapgen_batch with waypoints: generator inside k_waypoints_small (measured slower, r02s / r02t)
	int gen_warps = 8;        // ... warps that compute the next map's elevations during the relaxation
	int debug_walker = 0; // 0 auto, 1 never, 2 always
	DevBuf<uint32_t> scratch_tiles, scratch_dist;
	DevBuf<int16_t> scratch_elev_in, scratch_elev;
	DevBuf<uint16_t> scratch_flags;

	// monsters
	int nmon = 0;
	DevBuf<uint32_t> m_tile, m_tileStart, m_tileEnd, m_timeStart, m_timeEnd, m_bits, m_tta, m_damage;
	DevBuf<uint8_t> m_dot, m_alive;
	bool m_has_dot = false, m_has_damage = false, m_has_dots = false, m_has_resist = false, m_has_money = false;
	DevBuf<uint32_t> m_dot_duration, m_dot_damage, m_money; // damage-over-time state: [5][n], [5][n], [n]
	DevBuf<long long> m_life;
	DevBuf<uint8_t> m_resist;
	DevBuf<float> m_pos;
	DevBuf<uint8_t> m_class;
	bool m_has_class = false;
	// towers (attack pass) and the uniform-grid monster index they query
	int ntowers = 0;
	DevBuf<uint32_t> t_tile, t_damage, t_overTime, t_period, t_manaCost, t_delay, t_mana, t_hits;
	DevBuf<float> t_range, t_splash;
	DevBuf<uint8_t> t_invalid, t_element, t_bonus, t_targeting;
	DevBuf<unsigned long long> t_random;
	DevBuf<int> t_target;
	DevBuf<uint32_t> sp_count, sp_start, sp_binof, sp_rank, sp_sorted;
	DevBuf<float2> sp_xz;

	// scratch
	DevBuf<uint32_t> scratch_u32a, scratch_u32b, scratch_u32c, scratch_u32d;
	DevBuf<float> scratch_f;
	DevBuf<int> d_status;
	DevBuf<unsigned long long> d_damage;
	DevBuf<uint32_t> d_small; // alive count, spawn out ...
	DevBuf<int> d_trial;
	HostFlags *hflags = nullptr; // pinned
	unsigned char *pin_stage = nullptr; // pinned staging for small device->host results that go to pageable memory
	size_t pin_stage_bytes = 0;

	// timing probe
	bool profiling = false;
	struct ProfRegion
	{
		int kind;
		cudaEvent_t a, b;
	};
	std::vector<ProfRegion> prof_open;
	std::vector<cudaEvent_t> prof_pool;
};

namespace
{
	int fail(mtd_ctx *c, int code, const char *fmt, ...)
	{
		if (c)
		{
			va_list ap;
			va_start(ap, fmt);
			vsnprintf(c->err, sizeof(c->err), fmt, ap);
			va_end(ap);
		}
		return code;
	}

#define CU(call)                                                                                          \
	do                                                                                                    \
	{                                                                                                     \
		cudaError_t e__ = (call);                                                                         \
		if (e__ != cudaSuccess)                                                                           \
			(void)cudaGetLastError();                                                                     \
		if (e__ != cudaSuccess)                                                                           \
			return fail(ctx, e__ == cudaErrorMemoryAllocation ? MTD_ERR_NOMEM : MTD_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__)); \
	} while (0)

#define LAUNCHED()                                                  \
	do                                                              \
	{                                                               \
		ctx->launches++;                                            \
		CU(cudaGetLastError());                                     \
	} while (0)

	// RAII bracket of one kernel (or a loop of launches) with CUDA events when profiling is on
	struct ProfScope
	{
		mtd_ctx *ctx;
		cudaEvent_t b = nullptr;
		ProfScope(mtd_ctx *c, int kind) : ctx(c)
		{
			if (!c->profiling)
				return;
			auto take = [&]() -> cudaEvent_t
			{
				cudaEvent_t e = nullptr;
				if (!c->prof_pool.empty())
				{
					e = c->prof_pool.back();
					c->prof_pool.pop_back();
				}
				else
					cudaEventCreate(&e);
				return e;
			};
			cudaEvent_t a = take();
			b = take();
			cudaEventRecord(a, c->stream);
			c->prof_open.push_back({ kind, a, b });
		}
		~ProfScope()
		{
			if (b)
				cudaEventRecord(b, ctx->stream);
		}
	};

	struct DeviceGuard
	{
		int prev = -1;
		explicit DeviceGuard(int dev)
		{
			cudaGetDevice(&prev);
			if (prev != dev)
				cudaSetDevice(dev);
			else
				prev = -1;
		}
		~DeviceGuard()
		{
			if (prev >= 0)
				cudaSetDevice(prev);
		}
	};

	int status_to_error(mtd_ctx *ctx, int st)
	{
		if (st & ST_GUARD)
			return fail(ctx, MTD_ERR_OVERFLOW, "a distance reached the uint32 guard band (the reference's uint32 arithmetic would wrap)");
		if (st & ST_NO_VALID)
			return fail(ctx, MTD_ERR_STATE, "a generated map has no valid tile (the reference would never return from findSpawnPoint)");
		if (st & ST_PATH_TRUNC)
			return fail(ctx, MTD_ERR_CAPACITY, "a full path is longer than fullpath_cap (lengths are exact, tiles were truncated)");
		if (st & ST_PEER_TIMEOUT)
			return fail(ctx, MTD_ERR_STATE, "peer-coupled stripes: another rank did not reach a barrier in time (20 s to enter the run, 4 s per round: not launched, or died)");
		if (st & ST_ASYNC_TIMEOUT)
			return fail(ctx, MTD_ERR_CUDA, "the dataflow relaxation hit its time limit");
		if (st & ST_MONSTER_STUCK)
			return fail(ctx, MTD_ERR_STATE, "a monster stands on a tile outside the grid or cannot reach its next waypoint (the reference asserts on this, sources/game/monsters.cpp:80): its state was left untouched");
		return MTD_OK;
	}

	int sync_and_status(mtd_ctx *ctx)
	{
		CU(cudaMemcpyAsync(&ctx->hflags->status, ctx->d_status.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
		CU(cudaStreamSynchronize(ctx->stream));
		const int st = ctx->hflags->status;
		if (st)
		{
			CU(cudaMemsetAsync(ctx->d_status.p, 0, sizeof(int), ctx->stream));
			return status_to_error(ctx, st);
		}
		return MTD_OK;
	}

	// Results of one call going to the caller's host memory. A copy from the device into PAGEABLE memory is staged by the
	// driver copy by copy and blocks the calling thread each time (six of them cost more than the kernel on the game-sized
	// map); small result sets therefore travel into one pinned arena with asynchronous copies, one synchronisation, and are
	// moved to the caller's buffers by memcpy. Page-locked destinations and large results are copied straight.
	struct D2HItem
	{
		void *dst;
		const void *src;
		size_t bytes;
	};

	int d2h_results(mtd_ctx *ctx, const D2HItem *items, int n)
	{
		size_t total = 0, biggest = 0;
		int ib = -1;
		for (int i = 0; i < n; i++)
			if (items[i].dst && items[i].bytes)
			{
				total += (items[i].bytes + 15) & ~(size_t)15;
				if (items[i].bytes > biggest)
				{
					biggest = items[i].bytes;
					ib = i;
				}
			}
		bool stage = ib >= 0 && total <= ((size_t)4 << 20);
		if (stage)
		{
			cudaPointerAttributes at;
			if (cudaPointerGetAttributes(&at, items[ib].dst) == cudaSuccess)
				stage = at.type == cudaMemoryTypeUnregistered;
			else
				cudaGetLastError();
		}
		if (stage && ctx->pin_stage_bytes < total)
		{
			if (ctx->pin_stage)
				cudaFreeHost(ctx->pin_stage);
			ctx->pin_stage = nullptr;
			ctx->pin_stage_bytes = 0;
			const size_t want = total < ((size_t)1 << 20) ? ((size_t)1 << 20) : ((size_t)4 << 20);
			if (cudaHostAlloc((void **)&ctx->pin_stage, want, cudaHostAllocDefault) == cudaSuccess)
				ctx->pin_stage_bytes = want;
			else
			{
				cudaGetLastError();
				stage = false;
			}
		}
		size_t off = 0;
		for (int i = 0; i < n; i++)
			if (items[i].dst && items[i].bytes)
			{
				CU(cudaMemcpyAsync(stage ? (void *)(ctx->pin_stage + off) : items[i].dst, items[i].src, items[i].bytes, cudaMemcpyDeviceToHost, ctx->stream));
				off += (items[i].bytes + 15) & ~(size_t)15;
			}
		const int rc = sync_and_status(ctx);
		if (rc || !stage)
			return rc;
		off = 0;
		for (int i = 0; i < n; i++)
			if (items[i].dst && items[i].bytes)
			{
				memcpy(items[i].dst, ctx->pin_stage + off, items[i].bytes);
				off += (items[i].bytes + 15) & ~(size_t)15;
			}
		return MTD_OK;
	}

	// resident layers per pass that fit shared memory next to a usable queue (0 = the map does not fit the one-CTA path)
	int small_resident(int W, int H, int n)
	{
		if (W > 256 || H > 256 || n < 1)
			return 0;
		if ((size_t)(W + 1) * (H + 2) >= 65536) // queue entries are 16-bit indices into the padded frame
			return 0;
		for (int R = n < SMALL_MAX_RESIDENT ? n : SMALL_MAX_RESIDENT; R >= 1; R--)
		{
			// sized for the widest group (most queue segments), so every GW instantiation fits
			const SmallLayout L = small_layout(W, H, R, (uint32_t)MAX_SMEM, 0, 4);
			if (L.qseg >= 64 && L.total <= MAX_SMEM)
				return R;
		}
		return 0;
	}

	bool small_fits(int W, int H, int n)
	{
		return small_resident(W, H, n) > 0;
	}

	int grid_blocks(mtd_ctx *ctx, size_t n, int threads, int per_sm = 8)
	{
		size_t b = (n + threads - 1) / threads;
		const size_t cap = (size_t)ctx->sms * per_sm;
		return (int)(b < cap ? (b ? b : 1) : cap);
	}


	// cudaFuncSetAttribute / occupancy are per DEVICE: one entry per (device, kernel), guarded by a process-wide mutex
	// (contexts of several devices may live in one process and are locked independently)
	std::mutex g_cfg_mu;
	std::map<std::pair<int, const void *>, int> g_cfg;

	// opts `func` in to `smem` bytes of dynamic shared memory on the context's device (once) and returns the resident
	// CTAs per SM for `threads` threads, or -1 with the error recorded
	int kernel_config(mtd_ctx *ctx, const void *func, int threads, size_t smem)
	{
		std::lock_guard<std::mutex> lk(g_cfg_mu);
		const auto key = std::make_pair(ctx->device, func);
		auto it = g_cfg.find(key);
		if (it != g_cfg.end())
			return it->second;
		cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		int per_sm = 0;
		if (e == cudaSuccess)
			e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, func, threads, smem);
		if (e != cudaSuccess)
		{
			(void)cudaGetLastError();
			fail(ctx, MTD_ERR_CUDA, "kernel configuration failed: %s", cudaGetErrorString(e));
			return -1;
		}
		g_cfg[key] = per_sm;
		return per_sm;
	}

	// k_big_finalize_tiled moves rows as 16-byte vectors: the width and every base it touches must allow that
	bool finalize_vec_ok(mtd_ctx *ctx, const void *d_dist, const void *d_tiles)
	{
		return ctx->W % 4 == 0 && ctx->W >= 4 && (((uintptr_t)d_dist | (uintptr_t)d_tiles) & 15u) == 0;
	}

	// next-hop pass over rows [0, Hl) of `nlayers` layers, labels and weights staged through shared memory (needs finalize_vec_ok)
	void launch_finalize_tiled(mtd_ctx *ctx, int W, int Hl, int y0, int H, uint32_t *d_dist, size_t stride, int nsrc, int nlayers, const uint32_t *up_row, const uint32_t *up_wpack, const uint32_t *down_row, uint32_t *d_tiles)
	{
		const int ntx = (W + FIN_TW - 1) / FIN_TW, nty = (Hl + FIN_TH - 1) / FIN_TH;
		k_big_finalize_tiled<<<dim3((unsigned)(ntx * nty), (unsigned)nlayers), 256, 0, ctx->stream>>>(W, Hl, y0, H, d_dist, stride, ctx->wpack.p, ctx->blocked.p, ctx->sources.p, nsrc, up_row, up_wpack, down_row, d_tiles, stride, ntx);
	}

	// ---- small path launcher -------------------------------------------------------------------------
	template<int GW>
	int launch_waypoints_small_gw(mtd_ctx *ctx, WaypointsSmallParams P)
	{
		const int T = P.W * P.H;
		P.R = small_resident(P.W, P.H, P.n);
		if (P.R < 1)
			return fail(ctx, MTD_ERR_CAPACITY, "map of %d tiles with %d layers does not fit shared memory", T, P.n);
		P.smem_budget = (uint32_t)MAX_SMEM;
		P.qseg_cap = ctx->debug_qseg_cap;
		const SmallLayout L = small_layout(P.W, P.H, P.R, P.smem_budget, P.qseg_cap, GW);
		if (kernel_config(ctx, (const void *)k_waypoints_small<GW>, SMALL_BLOCK, MAX_SMEM) < 0)
			return MTD_ERR_CUDA;
		CU(ctx->map_counter.reserve(1));
		CU(cudaMemsetAsync(ctx->map_counter.p, 0, 4, ctx->stream));
		P.map_counter = ctx->map_counter.p;
		const int per_sm = L.total * 2 + 2048 <= MAX_SMEM ? 2 : 1;
		const int blocks = P.nmaps < ctx->sms * per_sm ? P.nmaps : ctx->sms * per_sm;
		// paths: a CTA that owns several maps hands the walks to its walker warp, which chases the layers in global memory
		const bool want_walks = P.fullpath_len && P.nsrc <= 1;
		P.walker = want_walks && ((P.nmaps > blocks && ctx->debug_walker != 1) || ctx->debug_walker == 2) ? 1 : 0;
		if (want_walks && (P.walker || P.n > P.R) && (!P.tiles || !P.dist))
		{
			// the caller did not ask for the layers: keep them in library-owned scratch
			const size_t need = (size_t)P.nmaps * P.n * T;
			if (!P.tiles)
			{
				CU(ctx->scratch_tiles.reserve(need));
				P.tiles = ctx->scratch_tiles.p;
			}
			if (!P.dist)
			{
				CU(ctx->scratch_dist.reserve(need));
				P.dist = ctx->scratch_dist.p;
			}
		}
		ProfScope prof(ctx, MTD_K_WAYPOINTS_SMALL);
		k_waypoints_small<GW><<<blocks, SMALL_BLOCK, L.total, ctx->stream>>>(P);
		LAUNCHED();
		return MTD_OK;
	}

	int launch_waypoints_small(mtd_ctx *ctx, const WaypointsSmallParams &P)
	{
		switch (ctx->debug_gw)
		{
		case 1:
			return launch_waypoints_small_gw<1>(ctx, P);
		case 2:
			return launch_waypoints_small_gw<2>(ctx, P);
		default:
			return launch_waypoints_small_gw<4>(ctx, P);
		}
	}

	// ---- big path -----------------------------------------------------------------------------------
	int big_prepare(mtd_ctx *ctx)
	{
		// derived data for the tiled path: packed edge weights + blocking mask + work-lists
		CU(ctx->wpack.reserve(ctx->T));
		CU(ctx->blocked.reserve(ctx->T));
		ctx->tilesX = (ctx->W + BIG_TS - 1) / BIG_TS;
		ctx->tilesY = (ctx->H + BIG_TS - 1) / BIG_TS;
		CU(ctx->counters.reserve(16));
		ProfScope prof(ctx, MTD_K_BIG_PREPARE);
		k_edge_weights<<<grid_blocks(ctx, ctx->T, 256), 256, 0, ctx->stream>>>(ctx->W, ctx->H, ctx->elev.p, nullptr, ctx->flags.p, ctx->wpack.p, ctx->blocked.p);
		LAUNCHED();
		return MTD_OK;
	}

	// Runs relaxation rounds from `round` until the work-list is empty or `budget` rounds are done (budget <= 0: no limit).
	// One cooperative launch of the persistent kernel; rounds are separated by grid-wide barriers inside it.
	int run_rounds(mtd_ctx *ctx, const BigParams &P, int &round, int budget, uint32_t &pending)
	{
		using WF = WarpFrame<BIG_TS>;
		const size_t smem = sizeof(WF) * BIG_WARPS;
		const int ctas_per_sm = kernel_config(ctx, (const void *)k_tile_rounds<BIG_TS, BIG_WARPS>, BIG_WARPS * 32, smem);
		if (ctas_per_sm < 0)
			return MTD_ERR_CUDA;
		if (ctas_per_sm < 1)
			return fail(ctx, MTD_ERR_CUDA, "the tile relaxation kernel does not fit an SM");
		const int limit = 4 * (ctx->tilesX + ctx->tilesY) * BIG_TS + 100000; // far above any wavefront depth: a safety net, not a tuning knob
		int max_rounds = budget > 0 ? budget : limit;
		uint32_t *d_out = nullptr;
		CU(cudaHostGetDevicePointer((void **)&d_out, ctx->hflags->round_out, 0));
		BigParams Pk = P;
		int round0 = round;
		void *args[] = { (void *)&Pk, (void *)&round0, (void *)&max_rounds, (void *)&d_out };
		int ctas = ctx->sms * ctas_per_sm;
		if (Pk.soft_grid)
			ctas = ctx->peer_ctas < ctas ? ctx->peer_ctas : ctas;
		{
			ProfScope prof(ctx, MTD_K_BIG_RELAX);
			if (Pk.soft_grid) // several stripes share this GPU: plain launches of small grids that are resident together
				k_tile_rounds<BIG_TS, BIG_WARPS><<<dim3((unsigned)ctas), dim3(BIG_WARPS * 32), smem, ctx->stream>>>(Pk, round0, max_rounds, d_out);
			else
				CU(cudaLaunchCooperativeKernel((const void *)k_tile_rounds<BIG_TS, BIG_WARPS>, dim3((unsigned)ctas), dim3(BIG_WARPS * 32), args, smem, ctx->stream));
			LAUNCHED();
		}
		if (P.ctl0)
			return MTD_OK; // peer-coupled: the ranks stop together inside the kernel; the caller waits in mtd_stripe_peer_finish
		CU(cudaStreamSynchronize(ctx->stream));
		pending = ctx->hflags->round_out[0];
		round = (int)ctx->hflags->round_out[1];
		if (budget <= 0 && pending != 0)
			return fail(ctx, MTD_ERR_CUDA, "tile relaxation did not converge after %d rounds", round);
		return MTD_OK;
	}


	// The dataflow form of run_rounds (tile_async.cuh): ONE launch, no global rounds. Returns MTD_ERR_CAPACITY (without a
	// message) when a CTA would own more items than its scheduler holds: the caller then uses the round-based kernel.
	bool async_fits(mtd_ctx *ctx, size_t nitems)
	{
		return (nitems + (size_t)ctx->sms - 1) / (size_t)ctx->sms + 64 <= (size_t)ASYNC_PCAP;
	}

	// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (no link-time dependency on libcuda)
	typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
									  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
	EncodeTiledFn encode_tiled_fn()
	{
		static EncodeTiledFn fn = nullptr;
		static std::once_flag once;
		std::call_once(once, []()
					   {
						   void *p = nullptr;
						   cudaDriverEntryPointQueryResult q;
						   if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
							   fn = (EncodeTiledFn)p;
						   else
							   (void)cudaGetLastError();
					   });
		return fn;
	}

	// a tensor of uint32 [nz][ny][nx] (row pitch nx, plane pitch `plane` elements) read in boxes of 40 x 34 x 1 words
	bool frame_tensor_map(CUtensorMap *tm, const void *base, int nx, int ny, int nz, size_t plane)
	{
		EncodeTiledFn enc = encode_tiled_fn();
		if (!enc || nx % 4 != 0 || ((uintptr_t)base & 15u) != 0 || (plane % 4) != 0)
			return false;
		const cuuint64_t gdim[3] = { (cuuint64_t)nx, (cuuint64_t)ny, (cuuint64_t)nz };
		const cuuint64_t gstride[2] = { (cuuint64_t)nx * 4u, (cuuint64_t)plane * 4u };
		const cuuint32_t box[3] = { (cuuint32_t)(BIG_TS + 8), (cuuint32_t)(BIG_TS + 2), 1u };
		const cuuint32_t estr[3] = { 1u, 1u, 1u };
		const int rank = nz > 0 ? 3 : 2;
		return enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, (cuuint32_t)rank, const_cast<void *>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
				   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
	}

	// peer = true: peer-coupled row stripes (P.ctl0 set). The kernel is only queued: the ranks enter, exchange border offers
	// and detect the end of the run on the device; the caller waits in mtd_stripe_peer_finish.
	int run_async(mtd_ctx *ctx, BigParams P, size_t nitems, bool peer = false)
	{
		const size_t smem = async_smem_bytes<BIG_TS>();
		const int threads = (ASYNC_WORKERS + 1) * 32;
		const int per_sm = kernel_config(ctx, (const void *)k_tile_async<BIG_TS>, threads, smem);
		if (per_sm < 0)
			return MTD_ERR_CUDA;
		if (per_sm < 1)
			return fail(ctx, MTD_ERR_CUDA, "the dataflow relaxation kernel does not fit an SM");
		int ctas = ctx->sms;
		if (peer && P.soft_grid)
		{
			// several stripes share this GPU (tests): every stripe's kernel must be resident at once, one CTA per SM
			const int share = ctx->sms / (ctx->peer_world > 0 ? ctx->peer_world : 1);
			ctas = ctx->peer_ctas < share ? ctx->peer_ctas : share;
			if (ctas < 1)
				return fail(ctx, MTD_ERR_CAPACITY, "more stripes than SMs on one GPU");
		}
		uint32_t cap = 64;
		while ((size_t)cap < (nitems + ctas - 1) / ctas + 64)
			cap <<= 1;
		CU(ctx->a_inbox.reserve((size_t)ctas * cap));
		CU(ctx->a_tail.reserve((size_t)ctas * 32));
		CU(ctx->a_ctamin.reserve((size_t)ctas));
		CU(ctx->a_ctl.reserve(ACTL_WORDS));
		CU(ctx->a_fly.reserve(nitems));
		CU(cudaMemsetAsync(ctx->a_fly.p, 0xFF, nitems * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->a_inbox.p, 0, (size_t)ctas * cap * 8, ctx->stream));
		CU(cudaMemsetAsync(ctx->a_tail.p, 0, (size_t)ctas * 32 * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->a_ctamin.p, 0xFF, (size_t)ctas * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->a_ctl.p, 0, ACTL_WORDS * 4, ctx->stream));
		AsyncParams A;
		memset(&A, 0, sizeof(A));
		A.inbox = ctx->a_inbox.p;
		A.tail = ctx->a_tail.p;
		A.ctamin = ctx->a_ctamin.p;
		A.ctl = ctx->a_ctl.p;
		A.cap_mask = cap - 1u;
		A.ctas = ctas;
		A.fly = ctx->a_fly.p;
		A.nb_rule = ctx->async_nb_rule;
		A.nb_margin = ctx->async_nb_margin;
		A.depth = ctx->async_depth;
		A.workers = ctx->async_workers;
		A.late_looks = ctx->async_late_looks;
		A.cont = ctx->async_cont;
		A.time_limit_ns = peer ? 8000000000ull : 20000000000ull; // an update that takes this long is reported, not waited for
		P.tile_delta = ctx->async_delta;
		// frames staged by TMA: one tensor map over the labels of all layers, one over the packed weights
		alignas(64) CUtensorMap tm_dist, tm_w;
		memset(&tm_dist, 0, sizeof(tm_dist));
		memset(&tm_w, 0, sizeof(tm_w));
		A.use_tma = !peer && ctx->big_tma && P.aligned16 && frame_tensor_map(&tm_dist, P.dist, P.W, P.y1 - P.y0, P.nlayers, P.layer_stride) && frame_tensor_map(&tm_w, P.wpack, P.W, P.y1 - P.y0, 0, 0) ? 1 : 0;
		ctx->last_tma = A.use_tma != 0;
		if (peer) // stripe_prepare seeded the round-based work-lists, which also posts keys: the dataflow kernel starts from "nothing pending"
			CU(cudaMemsetAsync(ctx->tilekey.p, 0xFF, nitems * 4, ctx->stream));
		k_async_seed<BIG_TS><<<(P.nlayers * P.nsrc + 127) / 128, 128, 0, ctx->stream>>>(P, A);
		LAUNCHED();
		void *args[] = { (void *)&P, (void *)&A, (void *)&tm_dist, (void *)&tm_w };
		{
			ProfScope prof(ctx, MTD_K_BIG_RELAX);
			if (peer && P.soft_grid)
				k_tile_async<BIG_TS><<<dim3((unsigned)ctas), dim3((unsigned)threads), smem, ctx->stream>>>(P, A, tm_dist, tm_w);
			else
				CU(cudaLaunchCooperativeKernel((const void *)k_tile_async<BIG_TS>, dim3((unsigned)ctas), dim3((unsigned)threads), args, smem, ctx->stream));
			LAUNCHED();
		}
		if (peer)
		{
			ctx->hflags->round_out[1] = 0u; // no rounds in this mode
			CU(cudaMemcpyAsync(ctx->hflags->async_ctl, ctx->a_ctl.p, 4 * 4, cudaMemcpyDeviceToHost, ctx->stream));
			ctx->last_async = true;
			return MTD_OK;
		}
		CU(cudaMemcpyAsync(ctx->hflags->async_ctl, ctx->a_ctl.p, 4 * 4, cudaMemcpyDeviceToHost, ctx->stream));
		CU(cudaStreamSynchronize(ctx->stream));
		if (ctx->hflags->async_ctl[ACTL_ABORT] != 0 || ctx->hflags->async_ctl[ACTL_OUTSTANDING] != 0)
			return fail(ctx, MTD_ERR_CUDA, "dataflow relaxation stopped with %u items outstanding (time limit)", ctx->hflags->async_ctl[ACTL_OUTSTANDING]);
		ctx->last_visits = ctx->hflags->async_ctl[ACTL_VISITS];
#ifdef MTD_TIMING
		{
			unsigned long long t[16], z[16] = { 0 };
			cudaMemcpyFromSymbol(t, g_async_timing, sizeof(t));
			cudaMemcpyToSymbol(g_async_timing, z, sizeof(z));
			const double v = (double)ctx->last_visits;
			double tot = 0;
			for (int q = 0; q < 7; q++)
				tot += (double)t[q];
			printf("ASYNCTIMING visits=%.0f | worker cycles per visit: wait=%.0f consume=%.0f stage=%.0f seed=%.0f engine=%.0f writeback=%.0f offers+finish=%.0f | idle share %.2f | steps/visit=%.1f cells/step=%.1f cycles/step=%.0f\n", v, t[0] / v, t[1] / v, t[2] / v,
				   t[3] / v, t[4] / v, t[5] / v, t[6] / v, tot > 0 ? t[0] / tot : 0.0, t[7] / v, t[7] ? (double)t[8] / t[7] : 0.0, t[7] ? (double)t[4] / t[7] : 0.0);
			unsigned long long x[4], zx[4] = { 0 };
			cudaMemcpyFromSymbol(x, g_async_sched_extra, sizeof(x));
			cudaMemcpyToSymbol(g_async_sched_extra, zx, sizeof(zx));
			{
				uint32_t dm = 0, zero = 0;
				cudaMemcpyFromSymbol(&dm, g_async_depth_max, 4);
				cudaMemcpyToSymbol(g_async_depth_max, &zero, 4);
				cudaMemset(nullptr, 0, 0);
				void *dp = nullptr;
				cudaGetSymbolAddress(&dp, g_async_depth);
				cudaMemset(dp, 0, sizeof(uint32_t) << 20);
				uint32_t sm = 0;
				cudaMemcpyFromSymbol(&sm, g_async_sdepth_max, 4);
				cudaMemcpyToSymbol(g_async_sdepth_max, &zero, 4);
				cudaGetSymbolAddress(&dp, g_async_sdepth);
				cudaMemset(dp, 0, sizeof(uint32_t) << 20);
				cudaGetSymbolAddress(&dp, g_async_sdepth_off);
				cudaMemset(dp, 0, sizeof(uint32_t) << 20);
				printf("ASYNCCHAIN longest chain of dependent visits = %u (offers only), %u (counting the previous visit of the same tile as a predecessor)\n", dm, sm);
			}
			{
				unsigned long long ph[8], zp[8] = { 0 };
				cudaMemcpyFromSymbol(ph, g_async_sched_phase, sizeof(ph));
				cudaMemcpyToSymbol(g_async_sched_phase, zp, sizeof(zp));
				const double it = t[9] ? (double)t[9] : 1.0;
				printf("ASYNCSCHEDPHASE cycles per scheduler iteration: loads + key refresh %.0f | drain + mailboxes + ctamin %.0f | hand-out loop %.0f | consume + ring %.0f | loop tail (stop test, sleep) %.0f\n", ph[0] / it, ph[2] / it, ph[3] / it, ph[4] / it, ph[5] / it);
			}
			printf("ASYNCSCHED2 per iteration: handed out %.2f | pending left: runnable but no free mailbox here %.2f, tile in flight %.2f, beyond the window %.2f\n", t[9] ? (double)x[1] / t[9] : 0.0, t[9] ? (double)t[14] / t[9] : 0.0,
				   t[9] ? (double)x[0] / t[9] : 0.0, t[9] ? (double)t[15] / t[9] : 0.0);
			printf("ASYNCSCHED iterations=%llu (%.2f us each at %d schedulers over the kernel) | per iteration: idle workers seen %.2f, left idle %.2f (held back by the window %.2f), pending entries %.1f\n", t[9], 0.0, ctas, t[9] ? (double)t[10] / t[9] : 0.0,
				   t[9] ? (double)t[11] / t[9] : 0.0, t[9] ? (double)t[12] / t[9] : 0.0, t[9] ? (double)t[13] / t[9] : 0.0);
			fflush(stdout);
		}
#endif
		return MTD_OK;
	}

	// nlayers independent layers, nsrc sources each (h_sources = [nlayers][nsrc], host), relaxed together: a work item is
	// one (layer, tile) pair. d_dist / d_tiles = [nlayers][T] (d_tiles may be null).
	int big_layers(mtd_ctx *ctx, const uint32_t *h_sources, int nlayers, int nsrc, uint32_t extra_blocked, uint32_t *d_tiles, uint32_t *d_dist)
	{
		const size_t nt = (size_t)ctx->tilesX * ctx->tilesY;
		const size_t nitems = nt * (size_t)nlayers;
		if (nitems >= 0xFFFFFFFFull)
			return fail(ctx, MTD_ERR_CAPACITY, "too many (layer, tile) work items");
		CU(ctx->list0.reserve(nitems));
		CU(ctx->list1.reserve(nitems));
		CU(ctx->list2.reserve(nitems));
		CU(ctx->stamp.reserve(nitems));
		CU(ctx->tilekey.reserve(nitems));
		CU(ctx->sources.reserve((size_t)nlayers * nsrc));
		CU(cudaMemcpyAsync(ctx->sources.p, h_sources, (size_t)nlayers * nsrc * 4, cudaMemcpyHostToDevice, ctx->stream));
		CU(cudaMemsetAsync(ctx->stamp.p, 0, nitems * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->tilekey.p, 0xFF, nitems * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->counters.p, 0, 16 * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->counters.p + 8, 0xFF, 3 * 4, ctx->stream)); // minkey[3] = m
		uint8_t prev_blocked = 0;
		// the trial grid of placingBuildingBlocksMonsters: one more blocked tile for the duration of this call, taken back on
		// EVERY way out (an error must not leave the trial building in the mask)
		struct RestoreBlocked
		{
			mtd_ctx *ctx;
			uint32_t tile;
			uint8_t prev;
			bool armed;
			~RestoreBlocked()
			{
				if (!armed)
					return;
				ctx->hflags->stage_blocked = prev;
				cudaMemcpyAsync(ctx->blocked.p + tile, &ctx->hflags->stage_blocked, 1, cudaMemcpyHostToDevice, ctx->stream);
				cudaStreamSynchronize(ctx->stream); // the staging word is reused by the next call
			}
		} restore{ ctx, extra_blocked, 0, false };
		if (extra_blocked != MTD_M)
		{
			CU(cudaMemcpyAsync(&prev_blocked, ctx->blocked.p + extra_blocked, 1, cudaMemcpyDeviceToHost, ctx->stream));
			CU(cudaStreamSynchronize(ctx->stream));
			ctx->hflags->stage_blocked = 1u;
			CU(cudaMemcpyAsync(ctx->blocked.p + extra_blocked, &ctx->hflags->stage_blocked, 1, cudaMemcpyHostToDevice, ctx->stream));
			restore.prev = prev_blocked;
			restore.armed = true;
		}
		{
			ProfScope prof(ctx, MTD_K_BIG_PREPARE);
			if (finalize_vec_ok(ctx, d_dist, nullptr))
				k_big_init4<<<grid_blocks(ctx, ctx->T / 4, 256), 256, 0, ctx->stream>>>(d_dist, ctx->T, nlayers, ctx->T / 4, ctx->blocked.p);
			else
				k_big_init<<<grid_blocks(ctx, ctx->T, 256), 256, 0, ctx->stream>>>(d_dist, ctx->T, nlayers, ctx->T, ctx->blocked.p);
			LAUNCHED();
		}

		BigParams P;
		memset(&P, 0, sizeof(P));
		P.W = ctx->W;
		P.H = ctx->H;
		P.y0 = 0;
		P.y1 = ctx->H;
		P.tilesX = ctx->tilesX;
		P.tilesY = ctx->tilesY;
		P.nlayers = nlayers;
		P.nsrc = nsrc;
		P.dist = d_dist;
		P.layer_stride = ctx->T;
		P.wpack = ctx->wpack.p;
		P.blocked = ctx->blocked.p;
		P.sources = ctx->sources.p;
		P.list[0] = ctx->list0.p;
		P.list[1] = ctx->list1.p;
		P.list[2] = ctx->list2.p;
		P.count = ctx->counters.p;
		P.cursor = ctx->counters.p + 3;
		P.total_active = ctx->counters.p + 6;
		P.stamp = ctx->stamp.p;
		P.key = ctx->tilekey.p;
		P.minkey = ctx->counters.p + 8;
		P.tile_delta = ctx->tile_delta;
		P.status = ctx->d_status.p;
		P.aligned16 = (ctx->W % 4 == 0) && ((uintptr_t)d_dist % 16 == 0) && ((uintptr_t)ctx->wpack.p % 16 == 0) ? 1 : 0;

		int round = 0;
		uint32_t pending = 0;
		if (ctx->big_async && async_fits(ctx, nitems))
		{
			const int rc = run_async(ctx, P, nitems);
			if (rc)
				return rc;
			ctx->last_async = true;
		}
		else
		{
			k_big_seed<BIG_TS><<<(nlayers * nsrc + 127) / 128, 128, 0, ctx->stream>>>(P);
			LAUNCHED();
			const int rc = run_rounds(ctx, P, round, 0, pending);
			if (rc)
				return rc;
			ctx->last_async = false;
		}
		ctx->last_rounds = (uint32_t)round;
		CU(cudaMemcpyAsync(&ctx->hflags->big_items, P.total_active, 4, cudaMemcpyDeviceToHost, ctx->stream));
#ifdef MTD_TIMING
		{
			unsigned long long t[16], z[16] = { 0 };
			cudaMemcpyFromSymbol(t, g_big_timing, sizeof(t));
			cudaMemcpyToSymbol(g_big_timing, z, sizeof(z));
			cudaStreamSynchronize(ctx->stream);
			const double v = (double)ctx->hflags->big_items;
			printf("BIGTIMING rounds=%d visits=%.0f | per visit cycles: stage=%.0f seed=%.0f engine=%.0f writeback=%.0f activate=%.0f | buckets/visit=%.1f cells/bucket=%.1f | longest visit=%llu cycles, visits over 60k cycles=%llu (steps per such visit %.1f, stage cycles %.0f), over 100k=%llu | list entries dropped as stale=%llu, deferred by the priority window=%llu\n", round, v, t[0] / v, t[1] / v, t[2] / v, t[3] / v, t[4] / v, t[5] / v, t[5] ? (double)t[6] / t[5] : 0.0, t[7], t[8], t[8] ? (double)t[10] / t[8] : 0.0, t[8] ? (double)t[11] / t[8] : 0.0, t[9], t[12], t[13]);
		}
#endif
		{
			ProfScope prof(ctx, MTD_K_BIG_DIRECTIONS);
			if (finalize_vec_ok(ctx, d_dist, d_tiles))
			{
				launch_finalize_tiled(ctx, ctx->W, ctx->H, 0, ctx->H, d_dist, ctx->T, nsrc, nlayers, nullptr, nullptr, nullptr, d_tiles);
			}
			else
			{
				dim3 grid((unsigned)grid_blocks(ctx, ctx->T, 256, nlayers > 4 ? 2 : 8), (unsigned)nlayers);
				k_big_finalize<<<grid, 256, 0, ctx->stream>>>(ctx->W, ctx->H, 0, ctx->H, d_dist, ctx->T, ctx->wpack.p, ctx->blocked.p, ctx->sources.p, nsrc, nullptr, nullptr, nullptr, d_tiles, ctx->T);
			}
			LAUNCHED();
		}
		return MTD_OK; // `restore` takes the trial building out of the mask
	}

	// ---- row stripe of a larger grid (one rank per stripe) -------------------------------
	// words of a peer block: control, the two halo rows the neighbours write, the two key arrays they post offers in
	struct PeerLayout
	{
		size_t halo_up, halo_down, act_up, act_down, words;
	};
	PeerLayout peer_layout(int W)
	{
		const size_t Wp = ((size_t)W + 3) & ~(size_t)3, Tp = (((size_t)W + BIG_TS - 1) / BIG_TS + 3) & ~(size_t)3;
		PeerLayout L;
		L.halo_up = PEER_CTL_WORDS;
		L.halo_down = L.halo_up + Wp;
		L.act_up = L.halo_down + Wp;
		L.act_down = L.act_up + Tp;
		L.words = L.act_down + Tp;
		return L;
	}

	BigParams stripe_params(mtd_ctx *ctx)
	{
		BigParams P;
		memset(&P, 0, sizeof(P));
		P.W = ctx->W;
		P.H = ctx->sH;
		P.y0 = ctx->sy0;
		P.y1 = ctx->sy1;
		P.tilesX = ctx->tilesX;
		P.tilesY = ctx->tilesY;
		P.nlayers = 1;
		P.nsrc = ctx->stripe_nsrc;
		P.dist = ctx->stripe_dist;
		P.layer_stride = ctx->T;
		P.wpack = ctx->wpack.p;
		P.blocked = ctx->blocked.p;
		P.sources = ctx->sources.p;
		P.list[0] = ctx->list0.p;
		P.list[1] = ctx->list1.p;
		P.list[2] = ctx->list2.p;
		P.count = ctx->counters.p;
		P.cursor = ctx->counters.p + 3;
		P.total_active = ctx->counters.p + 6;
		P.stamp = ctx->stamp.p;
		P.key = ctx->tilekey.p;
		P.minkey = ctx->counters.p + 8;
		P.tile_delta = ctx->tile_delta;
		P.status = ctx->d_status.p;
		P.aligned16 = (ctx->W % 4 == 0) && ((uintptr_t)ctx->stripe_dist % 16 == 0) && ((uintptr_t)ctx->wpack.p % 16 == 0) && ((uintptr_t)ctx->halo_up.p % 16 == 0) && ((uintptr_t)ctx->halo_down.p % 16 == 0) && ((uintptr_t)ctx->up_wpack.p % 16 == 0) ? 1 : 0;
		P.up_row = ctx->sy0 > 0 ? ctx->halo_up.p : nullptr;
		P.up_wpack = ctx->sy0 > 0 ? ctx->up_wpack.p : nullptr;
		P.down_row = ctx->sy1 < ctx->sH ? ctx->halo_down.p : nullptr;
		if (ctx->peer_on)
		{
			const PeerLayout L = peer_layout(ctx->W);
			uint32_t *mine = ctx->peer_blk.p;
			P.up_row = ctx->sy0 > 0 ? mine + L.halo_up : nullptr;
			P.down_row = ctx->sy1 < ctx->sH ? mine + L.halo_down : nullptr;
			P.act_up = ctx->peer_above ? mine + L.act_up : nullptr;
			P.act_down = ctx->peer_below ? mine + L.act_down : nullptr;
			// my first row is the "down" halo of the rank above, my last row the "up" halo of the rank below
			P.peer_up_halo = ctx->peer_above ? ctx->peer_above + L.halo_down : nullptr;
			P.peer_up_act = ctx->peer_above ? ctx->peer_above + L.act_down : nullptr;
			P.peer_down_halo = ctx->peer_below ? ctx->peer_below + L.halo_up : nullptr;
			P.peer_down_act = ctx->peer_below ? ctx->peer_below + L.act_up : nullptr;
			P.ctl0 = ctx->peer_root;
			P.ctl = mine;
			for (int r = 0; r < PEER_MAX_WORLD; r++)
				P.peers[r] = ctx->peers[r];
			P.world = ctx->peer_world;
			P.rank = ctx->peer_rank;
			P.soft_grid = ctx->peer_ctas > 0 ? 1 : 0;
		}
		return P;
	}

	// relax the stripe until its work-list is empty (the halo copies are constant meanwhile)
	int stripe_run(mtd_ctx *ctx, const BigParams &P)
	{
		// with a round budget the stripe stops early and reports what is still queued: the caller swaps halo rows more often,
		// so labels that will be overwritten from the other side of the cut are not propagated all the way first
		const int rc = run_rounds(ctx, P, ctx->stripe_round, ctx->stripe_budget, ctx->stripe_pending);
		if (rc)
			return rc;
		ctx->last_rounds = (uint32_t)ctx->stripe_round;
		return MTD_OK;
	}

	int directions_dev(mtd_ctx *ctx, const uint32_t *sources, int nsrc, uint32_t extra_blocked, uint32_t *d_tiles, uint32_t *d_dist)
	{
		if (!ctx->have_grid || ctx->stripe)
			return fail(ctx, MTD_ERR_STATE, ctx->stripe ? "context holds a row stripe: use the mtd_stripe_* calls" : "no grid uploaded");
		if (!sources || nsrc < 1)
			return fail(ctx, MTD_ERR_ARG, "sources must hold at least one tile");
		for (int i = 0; i < nsrc; i++)
			if (sources[i] >= ctx->T)
				return fail(ctx, MTD_ERR_ARG, "source tile %u outside the grid", sources[i]);
		if (extra_blocked != MTD_M && extra_blocked >= ctx->T)
			return fail(ctx, MTD_ERR_ARG, "extra_blocked tile outside the grid");
		if (!d_dist)
		{
			CU(ctx->trial_dist.reserve(ctx->T));
			d_dist = ctx->trial_dist.p;
		}
		if (ctx->small && nsrc <= MTD_MAXSRC_SMALL)
		{
			CU(ctx->lwp.reserve(64));
			CU(cudaMemcpyAsync(ctx->lwp.p + 32, sources, (size_t)nsrc * 4, cudaMemcpyHostToDevice, ctx->stream));
			WaypointsSmallParams P;
			memset(&P, 0, sizeof(P));
			P.W = ctx->W;
			P.H = ctx->H;
			P.n = 1;
			P.nmaps = 1;
			P.nsrc = nsrc;
			P.elev = ctx->elev.p;
			P.flags = ctx->flags.p;
			P.wp = ctx->lwp.p + 32;
			P.extra_blocked = extra_blocked;
			P.tiles = d_tiles;
			P.dist = d_dist;
			P.status = ctx->d_status.p;
			P.inc_tile = MTD_M;
			// the trial of placingBuildingBlocksMonsters from waypoint 0: layer 0 of the resident layers, with the candidate
			// blocked, is a repair of that layer (buildingsPlacement.cpp:18-42 runs a fresh Dijkstra for it)
			if (ctx->inc_enabled && ctx->inc_layers_ok && ctx->inc_dirty == 0 && nsrc == 1 && extra_blocked != MTD_M && ctx->inc_n >= 1 && sources[0] == ctx->inc_wp[0] && !ctx->h_blocking.empty() &&
				!ctx->h_blocking[extra_blocked] && d_dist != ctx->ldist.p)
			{
				P.prev_dist = ctx->ldist.p;
				P.inc_tile = extra_blocked;
				ctx->inc_trials++;
			}
			return launch_waypoints_small(ctx, P);
		}
		return big_layers(ctx, sources, 1, nsrc, extra_blocked, d_tiles, d_dist);
	}

	LayersView layers_view(mtd_ctx *ctx)
	{
		LayersView L;
		L.W = ctx->W;
		L.H = ctx->H;
		L.ox = ctx->ox;
		L.oy = ctx->oy;
		L.n = ctx->nlayers;
		L.elev = ctx->elev.p;
		L.tiles = ctx->ltiles.p;
		L.dist = ctx->ldist.p;
		L.wp = ctx->lwp.p;
		L.Dtab = ctx->dtab.p;
		return L;
	}

	MonstersView monsters_view(mtd_ctx *ctx)
	{
		MonstersView M;
		M.n = ctx->nmon;
		M.tile = ctx->m_tile.p;
		M.tileStart = ctx->m_tileStart.p;
		M.tileEnd = ctx->m_tileEnd.p;
		M.timeStart = ctx->m_timeStart.p;
		M.timeEnd = ctx->m_timeEnd.p;
		M.bits = ctx->m_bits.p;
		M.timeToArrive = ctx->m_tta.p;
		M.dotflags = ctx->m_has_dot ? ctx->m_dot.p : nullptr;
		M.damage = ctx->m_has_damage ? ctx->m_damage.p : nullptr;
		M.alive = ctx->m_alive.p;
		M.status = ctx->d_status.p;
		return M;
	}

	int waypoints_dev(mtd_ctx *ctx, const uint32_t *wp_tiles, int n, uint32_t fullpath_cap)
	{
		if (!ctx->have_grid || ctx->stripe)
			return fail(ctx, MTD_ERR_STATE, "no grid uploaded (or the context holds a row stripe)");
		if (n < 0 || n > MTD_MAX_WAYPOINTS)
			return fail(ctx, MTD_ERR_ARG, "waypoint count %d outside [0, %d]", n, MTD_MAX_WAYPOINTS);
		for (int i = 0; i < n; i++)
			if (wp_tiles[i] >= ctx->T)
				return fail(ctx, MTD_ERR_ARG, "waypoint tile %u outside the grid", wp_tiles[i]); // a rejected call leaves the resident layers as they were
		ctx->nlayers = 0; // from here on the resident layers are being rewritten: they count again once the launches went through
		if (n == 0)
			return MTD_OK; // paths.cpp:113-114
		CU(ctx->ltiles.reserve((size_t)n * ctx->T));
		CU(ctx->ldist.reserve((size_t)n * ctx->T));
		CU(ctx->lwp.reserve(64));
		CU(ctx->dtab.reserve(FIND_MAXN * FIND_MAXN));
		CU(ctx->fp_len.reserve(MTD_MAX_WAYPOINTS));
		CU(ctx->fp_dist.reserve(MTD_MAX_WAYPOINTS));
		CU(ctx->agg.reserve(4));
		ctx->fp_cap = fullpath_cap;
		if (fullpath_cap)
			CU(ctx->fp_buf.reserve((size_t)n * fullpath_cap));
		CU(cudaMemcpyAsync(ctx->lwp.p, wp_tiles, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
		CU(cudaMemsetAsync(ctx->fp_len.p, 0, MTD_MAX_WAYPOINTS * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->fp_dist.p, 0, MTD_MAX_WAYPOINTS * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->agg.p, 0, 16, ctx->stream));
		const bool walks = n <= FIND_MAXN;
		const bool small_path = ctx->small && small_fits(ctx->W, ctx->H, n);
		if (small_path)
		{
			WaypointsSmallParams P;
			memset(&P, 0, sizeof(P));
			P.W = ctx->W;
			P.H = ctx->H;
			P.n = n;
			P.nmaps = 1;
			P.nsrc = 1;
			P.elev = ctx->elev.p;
			P.flags = ctx->flags.p;
			P.wp = ctx->lwp.p;
			P.extra_blocked = MTD_M;
			P.tiles = ctx->ltiles.p;
			P.dist = ctx->ldist.p;
			P.fullpath = fullpath_cap ? ctx->fp_buf.p : nullptr;
			P.cap = fullpath_cap;
			P.fullpath_len = walks ? ctx->fp_len.p : nullptr;
			P.full_distance = ctx->fp_dist.p;
			P.agg = ctx->agg.p;
			P.status = ctx->d_status.p;
			// one blocking flip since the layers were computed for these waypoints: repair them instead of starting over
			// (same outputs: distances are the unique fixpoint, next hops / walks / aggregates are recomputed from them)
			bool same_wp = ctx->inc_layers_ok && ctx->inc_n == n && n <= small_resident(ctx->W, ctx->H, n);
			for (int i = 0; same_wp && i < n; i++)
				same_wp = ctx->inc_wp[i] == wp_tiles[i];
			if (ctx->inc_enabled && same_wp && ctx->inc_dirty == 1 && !ctx->h_blocking.empty())
			{
				P.prev_dist = ctx->ldist.p;
				P.inc_tile = ctx->inc_tile;
				ctx->inc_updates++;
			}
			else
				P.inc_tile = MTD_M;
			ctx->inc_layers_ok = false;
			int rc = launch_waypoints_small(ctx, P);
			if (rc)
				return rc;
			ctx->inc_layers_ok = !ctx->h_blocking.empty();
			ctx->inc_dirty = 0;
			ctx->inc_n = n;
			for (int i = 0; i < n; i++)
				ctx->inc_wp[i] = wp_tiles[i];
		}
		else
		{
			// large grid: all layers in one tiled relaxation (a work item is a (layer, tile) pair)
			int rc = big_layers(ctx, wp_tiles, n, 1, MTD_M, ctx->ltiles.p, ctx->ldist.p);
			if (rc)
				return rc;
		}
		ctx->nlayers = n;
		if (walks)
		{
			k_build_dtab<<<1, 64, 0, ctx->stream>>>(layers_view(ctx), ctx->dtab.p);
			LAUNCHED();
			if (!small_path)
			{
				// large grid: the walks run on the layers in global memory
				k_walk_global<<<1, 32, 0, ctx->stream>>>(layers_view(ctx), fullpath_cap ? ctx->fp_buf.p : nullptr, fullpath_cap, ctx->fp_len.p, ctx->fp_dist.p, ctx->agg.p, ctx->d_status.p);
				LAUNCHED();
			}
		}
		return MTD_OK;
	}
}

// =====================================================================================================
// lifecycle
// =====================================================================================================
extern "C" int mtd_version(void)
{
	return 100;
}

extern "C" int mtd_create(mtd_ctx **out, int device, void *stream)
{
	if (!out)
		return MTD_ERR_ARG;
	*out = nullptr;
	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || device < 0 || device >= ndev)
		return MTD_ERR_CUDA; // no CPU fallback by design
	mtd_ctx *ctx = new (std::nothrow) mtd_ctx();
	if (!ctx)
		return MTD_ERR_NOMEM;
	ctx->device = device;
	if (cudaSetDevice(device) != cudaSuccess)
	{
		delete ctx;
		return MTD_ERR_CUDA;
	}
	cudaDeviceGetAttribute(&ctx->sms, cudaDevAttrMultiProcessorCount, device);
	if (stream)
		ctx->stream = (cudaStream_t)stream;
	else
	{
		if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess)
		{
			delete ctx;
			return MTD_ERR_CUDA;
		}
		ctx->own_stream = true;
	}
	if (cudaHostAlloc((void **)&ctx->hflags, sizeof(HostFlags), cudaHostAllocMapped) != cudaSuccess || ctx->d_status.reserve(1) != cudaSuccess || ctx->d_damage.reserve(1) != cudaSuccess || ctx->d_small.reserve(8) != cudaSuccess || ctx->d_trial.reserve(1) != cudaSuccess)
	{
		mtd_destroy(ctx);
		return MTD_ERR_CUDA;
	}
	memset(ctx->hflags, 0, sizeof(HostFlags));
	cudaMemsetAsync(ctx->d_status.p, 0, sizeof(int), ctx->stream);
	*out = ctx;
	return MTD_OK;
}

namespace
{
	void peer_unlink(mtd_ctx *ctx);
}

extern "C" void mtd_destroy(mtd_ctx *ctx)
{
	if (!ctx)
		return;
	DeviceGuard g(ctx->device);
	cudaStreamSynchronize(ctx->stream);
	ctx->elev.release();
	ctx->flags.release();
	ctx->wpack.release();
	ctx->blocked.release();
	ctx->list0.release();
	ctx->list1.release();
	ctx->list2.release();
	ctx->tilekey.release();
	peer_unlink(ctx);
	ctx->peer_blk.release();
	ctx->halo_up.release();
	ctx->halo_down.release();
	ctx->up_wpack.release();
	ctx->up_blocked.release();
	ctx->scratch_elev_in.release();
	ctx->scratch_elev.release();
	ctx->scratch_flags.release();
	for (int b = 0; b < 2; b++)
	{
		ctx->mg[b].release();
		if (ctx->mg_done[b])
			cudaEventDestroy(ctx->mg_done[b]);
		if (ctx->mg_copied[b])
			cudaEventDestroy(ctx->mg_copied[b]);
	}
	if (ctx->copy_stream)
		cudaStreamDestroy(ctx->copy_stream);
	ctx->stamp.release();
	ctx->counters.release();
	ctx->sources.release();
	ctx->ltiles.release();
	ctx->ldist.release();
	ctx->lwp.release();
	ctx->dtab.release();
	ctx->fp_buf.release();
	ctx->fp_len.release();
	ctx->fp_dist.release();
	ctx->agg.release();
	ctx->trial_dist.release();
	ctx->map_counter.release();
	ctx->m_tile.release();
	ctx->m_tileStart.release();
	ctx->m_tileEnd.release();
	ctx->m_timeStart.release();
	ctx->m_timeEnd.release();
	ctx->m_bits.release();
	ctx->m_tta.release();
	ctx->m_damage.release();
	ctx->m_dot.release();
	ctx->m_alive.release();
	ctx->m_pos.release();
	ctx->scratch_u32a.release();
	ctx->scratch_u32b.release();
	ctx->scratch_u32c.release();
	ctx->scratch_u32d.release();
	ctx->scratch_f.release();
	ctx->d_status.release();
	ctx->d_damage.release();
	ctx->m_dot_duration.release();
	ctx->m_dot_damage.release();
	ctx->m_money.release();
	ctx->m_life.release();
	ctx->m_resist.release();
	ctx->d_small.release();
	ctx->d_trial.release();
	if (ctx->hflags)
		cudaFreeHost(ctx->hflags);
	if (ctx->pin_stage)
		cudaFreeHost(ctx->pin_stage);
	if (ctx->own_stream && ctx->stream)
		cudaStreamDestroy(ctx->stream);
	delete ctx;
}

extern "C" const char *mtd_last_error(const mtd_ctx *ctx)
{
	return ctx ? ctx->err : "null context";
}

extern "C" int mtd_synchronize(mtd_ctx *ctx)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return sync_and_status(ctx);
}

extern "C" uint64_t mtd_launch_count(const mtd_ctx *ctx)
{
	return ctx ? ctx->launches : 0;
}

extern "C" int mtd_debug_set(mtd_ctx *ctx, int key, int64_t value)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	switch (key)
	{
	case MTD_DEBUG_QSEG_CAP:
		ctx->debug_qseg_cap = (int)value;
		return MTD_OK;
	case MTD_DEBUG_GROUP_WARPS:
		if (value != 1 && value != 2 && value != 4)
			return fail(ctx, MTD_ERR_ARG, "group warps must be 1, 2 or 4");
		ctx->debug_gw = (int)value;
		return MTD_OK;
	case MTD_DEBUG_WALKER:
		ctx->debug_walker = (int)value;
		return MTD_OK;
	case MTD_DEBUG_TILE_DELTA:
		if (value < 1 || value > 0xFFFFFFFFll)
			return fail(ctx, MTD_ERR_ARG, "tile delta must be in 1 .. 2^32-1");
		ctx->tile_delta = (uint32_t)value;
		return MTD_OK;
	case MTD_DEBUG_BIG_ASYNC:
		ctx->big_async = value != 0;
		return MTD_OK;
	case MTD_DEBUG_ASYNC_NB_RULE:
		ctx->async_nb_rule = value >= 0 ? 1 : 0; // value = margin (label units); negative = rule off
		ctx->async_nb_margin = value >= 0 ? (uint32_t)value : 0u;
		return MTD_OK;
	case MTD_DEBUG_ASYNC_CONT:
		ctx->async_cont = value != 0;
		return MTD_OK;
	case MTD_DEBUG_ASYNC_LATE_LOOKS:
		ctx->async_late_looks = value < 0 ? 0 : value > 4 ? 4 : (int)value;
		return MTD_OK;
	case MTD_DEBUG_ASYNC_WORKERS:
		ctx->async_workers = value < 1 ? 1 : value > ASYNC_WORKERS ? ASYNC_WORKERS : (int)value;
		return MTD_OK;
	case MTD_DEBUG_ASYNC_DEPTH:
		ctx->async_depth = value >= 2 ? 2 : 1;
		return MTD_OK;
	case MTD_DEBUG_INCREMENTAL:
		ctx->inc_enabled = value != 0;
		return MTD_OK;
	case MTD_DEBUG_FUSED_MAPGEN:
		ctx->fused_mapgen = value != 0;
		return MTD_OK;
	case MTD_DEBUG_GEN_WARPS:
		ctx->gen_warps = value < 0 ? 0 : value > 16 ? 16 : (int)value;
		return MTD_OK;
	case MTD_DEBUG_BIG_TMA:
		ctx->big_tma = value != 0;
		return MTD_OK;
	case MTD_DEBUG_ASYNC_DELTA:
		if (value < 1 || value > 0xFFFFFFFFll)
			return fail(ctx, MTD_ERR_ARG, "async delta must be in 1 .. 2^32-1");
		ctx->async_delta = (uint32_t)value;
		return MTD_OK;
	default:
		return fail(ctx, MTD_ERR_ARG, "unknown debug key %d", key);
	}
}

extern "C" int mtd_big_stats(mtd_ctx *ctx, uint32_t *out4)
{
	if (!ctx || !out4)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	out4[0] = ctx->last_rounds;
	out4[1] = ctx->last_async ? ctx->last_visits : ctx->hflags->big_items;
	out4[2] = ctx->last_async ? ctx->hflags->async_ctl[ACTL_IDLE_POLLS] : 0u;
	out4[3] = ctx->last_async ? (ctx->last_tma ? 3u : 1u) : 0u;
	if (ctx->small)
	{
		// one-CTA path: how many updates / trials were incremental repairs since the context was created
		out4[0] = (uint32_t)ctx->inc_updates;
		out4[1] = (uint32_t)ctx->inc_trials;
	}
	return MTD_OK;
}

extern "C" int mtd_profile_enable(mtd_ctx *ctx, int on)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	ctx->profiling = on != 0;
	return MTD_OK;
}

extern "C" int mtd_profile_read(mtd_ctx *ctx, int kind, double *total_ms, uint64_t *regions)
{
	if (!ctx || kind < 0 || kind >= MTD_K_COUNT)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	CU(cudaStreamSynchronize(ctx->stream));
	double sum = 0;
	uint64_t n = 0;
	std::vector<mtd_ctx::ProfRegion> keep;
	for (const auto &r : ctx->prof_open)
	{
		if (r.kind != kind)
		{
			keep.push_back(r);
			continue;
		}
		float ms = 0;
		if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess)
		{
			sum += ms;
			n++;
		}
		else
			(void)cudaGetLastError();
		ctx->prof_pool.push_back(r.a);
		ctx->prof_pool.push_back(r.b);
	}
	ctx->prof_open.swap(keep);
	if (total_ms)
		*total_ms = sum;
	if (regions)
		*regions = n;
	return MTD_OK;
}

// =====================================================================================================
// grid
// =====================================================================================================
static int grid_common(mtd_ctx *ctx, int32_t W, int32_t H, int32_t ox, int32_t oy, const int16_t *elev, const uint16_t *flags, cudaMemcpyKind kind)
{
	if (W < 1 || H < 1 || !elev || !flags)
		return fail(ctx, MTD_ERR_ARG, "grid needs W, H >= 1 and non-null elev/flags");
	if ((uint64_t)W * (uint64_t)H >= 0xFFFFFFFFull)
		return fail(ctx, MTD_ERR_ARG, "grid too large for uint32 tile indices");
	// The reference mutates Grid::flags in place and calls update() (buildingsPlacement.cpp:128-129), so the drop-in
	// adapter re-uploads the whole grid before every update. When it is the grid this context already holds (same shape,
	// same elevations) only the flags travel, the resident layers stay, and the blocking bits that flipped are counted:
	// exactly one flip lets the next update repair the layers instead of recomputing them.
	if (kind == cudaMemcpyHostToDevice && ctx->have_grid && !ctx->stripe && ctx->small && W == ctx->W && H == ctx->H && ox == ctx->ox && oy == ctx->oy && ctx->h_elev.size() == (size_t)W * H &&
		ctx->h_blocking.size() == ctx->h_elev.size() && memcmp(ctx->h_elev.data(), elev, ctx->h_elev.size() * 2) == 0)
	{
		for (size_t i = 0; i < ctx->T; i++)
		{
			const uint8_t nb = (flags[i] & F_BLOCKING) ? 1 : 0;
			if (nb != ctx->h_blocking[i])
			{
				ctx->h_blocking[i] = nb;
				ctx->inc_dirty++;
				ctx->inc_tile = (uint32_t)i;
			}
		}
		CU(cudaMemcpyAsync(ctx->flags.p, flags, ctx->T * 2, kind, ctx->stream));
		CU(cudaStreamSynchronize(ctx->stream));
		return MTD_OK;
	}
	ctx->have_grid = false;
	ctx->stripe = false;
	ctx->nlayers = 0;
	ctx->W = W;
	ctx->H = H;
	ctx->ox = ox;
	ctx->oy = oy;
	ctx->T = (size_t)W * H;
	CU(ctx->elev.reserve(ctx->T));
	CU(ctx->flags.reserve(ctx->T));
	CU(cudaMemcpyAsync(ctx->elev.p, elev, ctx->T * 2, kind, ctx->stream));
	CU(cudaMemcpyAsync(ctx->flags.p, flags, ctx->T * 2, kind, ctx->stream));
	ctx->small = small_fits(W, H, 1);
	if (!ctx->small)
	{
		int rc = big_prepare(ctx);
		if (rc)
			return rc;
	}
	if (kind == cudaMemcpyHostToDevice)
		CU(cudaStreamSynchronize(ctx->stream)); // caller may reuse its host arrays (pageable memory is staged, but be explicit)
	ctx->inc_layers_ok = false;
	ctx->inc_dirty = 0;
	ctx->h_blocking.clear();
	ctx->h_elev.clear();
	if (kind == cudaMemcpyHostToDevice && ctx->small)
	{
		ctx->h_blocking.resize(ctx->T);
		for (size_t i = 0; i < ctx->T; i++)
			ctx->h_blocking[i] = (flags[i] & F_BLOCKING) ? 1 : 0;
		ctx->h_elev.assign(elev, elev + ctx->T);
	}
	ctx->have_grid = true;
	return MTD_OK;
}

extern "C" int mtd_grid_upload(mtd_ctx *ctx, int32_t W, int32_t H, int32_t ox, int32_t oy, const int16_t *elev, const uint16_t *flags)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return grid_common(ctx, W, H, ox, oy, elev, flags, cudaMemcpyHostToDevice);
}

extern "C" int mtd_grid_upload_dev(mtd_ctx *ctx, int32_t W, int32_t H, int32_t ox, int32_t oy, const int16_t *d_elev, const uint16_t *d_flags)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return grid_common(ctx, W, H, ox, oy, d_elev, d_flags, cudaMemcpyDeviceToDevice);
}

// ---- row stripes ---------------------------------------------------------------------------------------
static int stripe_common(mtd_ctx *ctx, int32_t W, int32_t H, int32_t y0, int32_t y1, const int16_t *elev_rows, const uint16_t *flags_rows, cudaMemcpyKind kind)
{
	if (W < 1 || H < 1 || y0 < 0 || y1 > H || y0 >= y1 || !elev_rows || !flags_rows)
		return fail(ctx, MTD_ERR_ARG, "stripe needs 0 <= y0 < y1 <= H and non-null rows");
	if ((uint64_t)W * (uint64_t)H >= 0xFFFFFFFFull)
		return fail(ctx, MTD_ERR_ARG, "grid too large for uint32 tile indices");
	const int Hl = y1 - y0;
	const int above = y0 > 0 ? 1 : 0, below = y1 < H ? 1 : 0;
	ctx->have_grid = false;
	ctx->nlayers = 0;
	ctx->W = W;
	ctx->H = Hl;
	ctx->ox = ctx->oy = 0;
	ctx->T = (size_t)W * Hl;
	ctx->sy0 = y0;
	ctx->sy1 = y1;
	ctx->sH = H;
	ctx->stripe = true;
	ctx->small = false;
	ctx->stripe_dist = nullptr;
	// elev and flags keep the halo rows too: [above + Hl + below][W]
	const size_t rows = (size_t)(above + Hl + below);
	CU(ctx->elev.reserve(rows * W));
	CU(ctx->flags.reserve(rows * W));
	CU(cudaMemcpyAsync(ctx->elev.p, elev_rows, rows * W * 2, kind, ctx->stream));
	CU(cudaMemcpyAsync(ctx->flags.p, flags_rows, rows * W * 2, kind, ctx->stream));
	CU(ctx->wpack.reserve(ctx->T));
	CU(ctx->blocked.reserve(ctx->T));
	CU(ctx->halo_up.reserve(W));
	CU(ctx->halo_down.reserve(W));
	CU(ctx->up_wpack.reserve(W));
	CU(ctx->up_blocked.reserve(W));
	CU(ctx->counters.reserve(16));
	ctx->tilesX = (W + BIG_TS - 1) / BIG_TS;
	ctx->tilesY = (Hl + BIG_TS - 1) / BIG_TS;
	const int16_t *own_e = ctx->elev.p + (size_t)above * W;
	const uint16_t *own_f = ctx->flags.p + (size_t)above * W;
	{
		ProfScope prof(ctx, MTD_K_BIG_PREPARE);
		k_edge_weights<<<grid_blocks(ctx, ctx->T, 256), 256, 0, ctx->stream>>>(W, Hl, own_e, below ? own_e + (size_t)Hl * W : nullptr, own_f, ctx->wpack.p, ctx->blocked.p);
		LAUNCHED();
		if (above) // weights of the edges (y0-1) -> y0, stored on the upper cell like everywhere else
		{
			k_edge_weights<<<grid_blocks(ctx, W, 256), 256, 0, ctx->stream>>>(W, 1, ctx->elev.p, own_e, ctx->flags.p, ctx->up_wpack.p, ctx->up_blocked.p);
			LAUNCHED();
		}
	}
	if (kind == cudaMemcpyHostToDevice)
		CU(cudaStreamSynchronize(ctx->stream));
	ctx->have_grid = true;
	return MTD_OK;
}

extern "C" int mtd_stripe_upload(mtd_ctx *ctx, int32_t W, int32_t H, int32_t y0, int32_t y1, const int16_t *elev_rows, const uint16_t *flags_rows)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return stripe_common(ctx, W, H, y0, y1, elev_rows, flags_rows, cudaMemcpyHostToDevice);
}

extern "C" int mtd_stripe_upload_dev(mtd_ctx *ctx, int32_t W, int32_t H, int32_t y0, int32_t y1, const int16_t *d_elev_rows, const uint16_t *d_flags_rows)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return stripe_common(ctx, W, H, y0, y1, d_elev_rows, d_flags_rows, cudaMemcpyDeviceToDevice);
}

namespace
{
	int stripe_prepare(mtd_ctx *ctx, const uint32_t *sources, int32_t nsrc, uint32_t *d_dist, bool peer);

	// work-lists and per-tile words of a stripe (allocations may wait for the device: peer-coupled stripes make them when
	// they connect, not between the launches of ranks that already wait for each other)
	int stripe_reserve(mtd_ctx *ctx, int nsrc)
	{
		const size_t nt = (size_t)ctx->tilesX * ctx->tilesY;
		CU(ctx->list0.reserve(nt));
		CU(ctx->list1.reserve(nt));
		CU(ctx->list2.reserve(nt));
		CU(ctx->stamp.reserve(nt));
		CU(ctx->tilekey.reserve(nt));
		CU(ctx->sources.reserve(nsrc));
		return MTD_OK;
	}
}

extern "C" int mtd_stripe_begin_dev(mtd_ctx *ctx, const uint32_t *sources, int32_t nsrc, uint32_t *d_dist)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	int rc = stripe_prepare(ctx, sources, nsrc, d_dist, false);
	if (rc)
		return rc;
	return stripe_run(ctx, stripe_params(ctx));
}

namespace
{
// distances := m / parked, sources seeded, border tiles under halo sources queued; everything on the context's stream
int stripe_prepare(mtd_ctx *ctx, const uint32_t *sources, int32_t nsrc, uint32_t *d_dist, bool peer)
{
	if (!ctx->have_grid || !ctx->stripe)
		return fail(ctx, MTD_ERR_STATE, "no stripe uploaded");
	if (!sources || nsrc < 1 || !d_dist)
		return fail(ctx, MTD_ERR_ARG, "stripe_begin needs sources and a distance buffer");
	for (int i = 0; i < nsrc; i++)
		if ((uint64_t)sources[i] >= (uint64_t)ctx->W * (uint64_t)ctx->sH)
			return fail(ctx, MTD_ERR_ARG, "source tile %u outside the grid", sources[i]);
	const size_t nt = (size_t)ctx->tilesX * ctx->tilesY;
	int rc = stripe_reserve(ctx, nsrc);
	if (rc)
		return rc;
	CU(cudaMemcpyAsync(ctx->sources.p, sources, (size_t)nsrc * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemsetAsync(ctx->stamp.p, 0, nt * 4, ctx->stream));
	CU(cudaMemsetAsync(ctx->tilekey.p, 0xFF, nt * 4, ctx->stream));
	CU(cudaMemsetAsync(ctx->counters.p, 0, 16 * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->counters.p + 8, 0xFF, 3 * 4, ctx->stream)); // minkey[3] = m
	if (!peer)
	{
		CU(cudaMemsetAsync(ctx->halo_up.p, 0xFF, (size_t)ctx->W * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->halo_down.p, 0xFF, (size_t)ctx->W * 4, ctx->stream));
	}
	ctx->peer_on = peer; // peer-coupled: the halo rows live in the peer block and the neighbours fill them
	ctx->stripe_dist = d_dist;
	ctx->stripe_nsrc = nsrc;
	ctx->stripe_round = 0;
	ctx->stripe_pending = 0;
	{
		ProfScope prof(ctx, MTD_K_BIG_PREPARE);
		if (finalize_vec_ok(ctx, d_dist, nullptr))
			k_big_init4<<<grid_blocks(ctx, ctx->T / 4, 256), 256, 0, ctx->stream>>>(d_dist, ctx->T, 1, ctx->T / 4, ctx->blocked.p);
		else
			k_big_init<<<grid_blocks(ctx, ctx->T, 256), 256, 0, ctx->stream>>>(d_dist, ctx->T, 1, ctx->T, ctx->blocked.p);
		LAUNCHED();
	}
	const BigParams P = stripe_params(ctx);
	k_big_seed<BIG_TS><<<(nsrc + 127) / 128, 128, 0, ctx->stream>>>(P);
	LAUNCHED();
	k_stripe_seed_halo<BIG_TS><<<(nsrc + 127) / 128, 128, 0, ctx->stream>>>(P);
	LAUNCHED();
	return MTD_OK;
}
}

extern "C" int mtd_stripe_exchange_dev(mtd_ctx *ctx, const uint32_t *d_row_above, const uint32_t *d_row_below, uint32_t *improved_out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->have_grid || !ctx->stripe || !ctx->stripe_dist)
		return fail(ctx, MTD_ERR_STATE, "stripe_exchange before stripe_begin");
	const BigParams P = stripe_params(ctx);
	uint32_t *d_improved = ctx->counters.p + 7;
	CU(cudaMemsetAsync(d_improved, 0, 4, ctx->stream));
	const int blocks = (ctx->W + 255) / 256;
	if (d_row_above && ctx->sy0 > 0)
	{
		k_stripe_merge<BIG_TS><<<blocks, 256, 0, ctx->stream>>>(P, ctx->halo_up.p, d_row_above, 0, ctx->stripe_round, d_improved);
		LAUNCHED();
	}
	if (d_row_below && ctx->sy1 < ctx->sH)
	{
		k_stripe_merge<BIG_TS><<<blocks, 256, 0, ctx->stream>>>(P, ctx->halo_down.p, d_row_below, ctx->tilesY - 1, ctx->stripe_round, d_improved);
		LAUNCHED();
	}
	CU(cudaMemcpyAsync(&ctx->hflags->stripe_improved, d_improved, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	const uint32_t improved = ctx->hflags->stripe_improved;
	if (improved_out)
		*improved_out = improved;
	if (improved == 0 && ctx->stripe_pending == 0)
		return MTD_OK;
	return stripe_run(ctx, P);
}

extern "C" int mtd_stripe_round_budget(mtd_ctx *ctx, int32_t rounds)
{
	if (!ctx || rounds < 0)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	ctx->stripe_budget = rounds;
	return MTD_OK;
}

extern "C" int mtd_stripe_pending(mtd_ctx *ctx, uint32_t *items_out)
{
	if (!ctx || !items_out)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	*items_out = ctx->stripe ? ctx->stripe_pending : 0u;
	return MTD_OK;
}

extern "C" int mtd_stripe_finalize_dev(mtd_ctx *ctx, uint32_t *d_tiles)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->have_grid || !ctx->stripe || !ctx->stripe_dist)
		return fail(ctx, MTD_ERR_STATE, "stripe_finalize before stripe_begin");
	const BigParams P = stripe_params(ctx);
	{
		ProfScope prof(ctx, MTD_K_BIG_DIRECTIONS);
		if (finalize_vec_ok(ctx, ctx->stripe_dist, d_tiles))
			launch_finalize_tiled(ctx, ctx->W, ctx->H, ctx->sy0, ctx->sH, ctx->stripe_dist, ctx->T, ctx->stripe_nsrc, 1, P.up_row, P.up_wpack, P.down_row, d_tiles);
		else
			k_big_finalize<<<dim3((unsigned)grid_blocks(ctx, ctx->T, 256, 8), 1u), 256, 0, ctx->stream>>>(ctx->W, ctx->H, ctx->sy0, ctx->sH, ctx->stripe_dist, ctx->T, ctx->wpack.p, ctx->blocked.p, ctx->sources.p, ctx->stripe_nsrc, P.up_row, P.up_wpack, P.down_row, d_tiles, ctx->T);
		LAUNCHED();
	}
	ctx->stripe_dist = nullptr;
	return MTD_OK;
}

// ---- peer-coupled stripes -----------------------------------------------------------------------------
namespace
{
	int peer_block_create(mtd_ctx *ctx)
	{
		if (!ctx->have_grid || !ctx->stripe)
			return fail(ctx, MTD_ERR_STATE, "no stripe uploaded");
		const PeerLayout L = peer_layout(ctx->W);
		if (ctx->peer_blk.cap >= L.words)
		{
			// (Re-)connecting: the barrier words restart from zero and the halo / key arrays from "nothing offered", so a rank
			// that timed out in an earlier run cannot leave the ranks at different epochs (every rank resets its OWN block here;
			// the caller's host barrier after the connect makes sure nobody launches before all blocks are reset)
			CU(cudaMemsetAsync(ctx->peer_blk.p, 0xFF, L.words * 4, ctx->stream));
			CU(cudaMemsetAsync(ctx->peer_blk.p, 0, PEER_CTL_WORDS * 4, ctx->stream));
			CU(cudaStreamSynchronize(ctx->stream));
			return MTD_OK;
		}
		if (ctx->peer_linked)
			return fail(ctx, MTD_ERR_STATE, "the stripe grew while its peer block is mapped by other ranks: destroy the contexts and connect again");
		CU(ctx->peer_blk.reserve(L.words));
		// halo rows and key arrays: m (nothing offered); control words: 0, smallest keys m
		CU(cudaMemsetAsync(ctx->peer_blk.p, 0xFF, L.words * 4, ctx->stream));
		CU(cudaMemsetAsync(ctx->peer_blk.p, 0, PEER_CTL_WORDS * 4, ctx->stream));
		CU(cudaStreamSynchronize(ctx->stream));
		return MTD_OK;
	}

	void peer_unlink(mtd_ctx *ctx)
	{
		for (int r = 0; r < PEER_MAX_WORLD; r++)
		{
			if (ctx->peer_ipc && ctx->peers[r] && ctx->peers[r] != ctx->peer_blk.p)
				cudaIpcCloseMemHandle(ctx->peers[r]);
			ctx->peers[r] = nullptr;
		}
		ctx->peer_above = ctx->peer_below = ctx->peer_root = nullptr;
		ctx->peer_ipc = ctx->peer_linked = ctx->peer_on = false;
	}

	int peer_linked_up(mtd_ctx *ctx, int rank, int world, int ctas)
	{
		ctx->peer_above = rank > 0 ? ctx->peers[rank - 1] : nullptr;
		ctx->peer_below = rank < world - 1 ? ctx->peers[rank + 1] : nullptr;
		ctx->peer_root = ctx->peers[0];
		int rc = stripe_reserve(ctx, 4096);
		if (rc)
			return rc;
		ctx->peer_rank = rank;
		ctx->peer_world = world;
		ctx->peer_ctas = ctas;
		ctx->peer_linked = true;
		return MTD_OK;
	}
}

extern "C" int mtd_stripe_peer_export(mtd_ctx *ctx, void *handle64)
{
	if (!ctx || !handle64)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the ABI passes IPC handles as 64 opaque bytes");
	int rc = peer_block_create(ctx);
	if (rc)
		return rc;
	cudaIpcMemHandle_t h;
	CU(cudaIpcGetMemHandle(&h, ctx->peer_blk.p));
	memcpy(handle64, &h, 64);
	return MTD_OK;
}

extern "C" int mtd_stripe_peer_connect(mtd_ctx *ctx, int32_t rank, int32_t world, const void *handles)
{
	if (!ctx || !handles || world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	int rc = peer_block_create(ctx);
	if (rc)
		return rc;
	peer_unlink(ctx);
	ctx->peer_ipc = true;
	for (int r = 0; r < world; r++)
	{
		if (r == rank)
		{
			ctx->peers[r] = ctx->peer_blk.p;
			continue;
		}
		cudaIpcMemHandle_t h;
		memcpy(&h, (const uint8_t *)handles + (size_t)r * 64, 64);
		void *p = nullptr;
		CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
		ctx->peers[r] = (uint32_t *)p;
	}
	return peer_linked_up(ctx, rank, world, 0);
}

extern "C" int mtd_stripe_peer_connect_local(mtd_ctx *ctx, int32_t rank, int32_t world, mtd_ctx *const *all, int32_t ctas)
{
	if (!ctx || !all || world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world || ctas < 1 || all[rank] != ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	int rc = peer_block_create(ctx);
	if (rc)
		return rc;
	peer_unlink(ctx);
	for (int r = 0; r < world; r++)
	{
		if (!all[r] || !all[r]->peer_blk.p || all[r]->device != ctx->device)
			return fail(ctx, MTD_ERR_STATE, "a linked context has no peer block yet (call mtd_stripe_peer_export on every context first) or lives on another device");
		ctx->peers[r] = all[r]->peer_blk.p;
	}
	return peer_linked_up(ctx, rank, world, ctas);
}

extern "C" int mtd_stripe_peer_launch_dev(mtd_ctx *ctx, const uint32_t *sources, int32_t nsrc, uint32_t *d_dist, uint32_t *d_tiles)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->peer_linked)
		return fail(ctx, MTD_ERR_STATE, "stripe_peer_launch before stripe_peer_connect");
	int rc = stripe_prepare(ctx, sources, nsrc, d_dist, true);
	if (rc)
		return rc;
	const BigParams P = stripe_params(ctx);
	const size_t nitems = (size_t)ctx->tilesX * ctx->tilesY;
	if (ctx->big_async && async_fits(ctx, nitems) && (!P.soft_grid || ctx->sms / ctx->peer_world >= 1))
		rc = run_async(ctx, P, nitems, true); // dataflow kernel: border offers and termination go rank to rank, no rounds
	else
	{
		int round = 0;
		uint32_t pending = 0;
		rc = run_rounds(ctx, P, round, 0, pending); // returns once the kernel is queued: the ranks stop together on the device
		ctx->last_async = false;
	}
	if (rc)
		return rc;
	{
		ProfScope prof(ctx, MTD_K_BIG_DIRECTIONS);
		if (finalize_vec_ok(ctx, d_dist, d_tiles))
			launch_finalize_tiled(ctx, ctx->W, ctx->H, ctx->sy0, ctx->sH, d_dist, ctx->T, nsrc, 1, P.up_row, P.up_wpack, P.down_row, d_tiles);
		else
			k_big_finalize<<<dim3((unsigned)grid_blocks(ctx, ctx->T, 256, 8), 1u), 256, 0, ctx->stream>>>(ctx->W, ctx->H, ctx->sy0, ctx->sH, d_dist, ctx->T, ctx->wpack.p, ctx->blocked.p, ctx->sources.p, nsrc, P.up_row, P.up_wpack, P.down_row, d_tiles, ctx->T);
		LAUNCHED();
	}
	ctx->stripe_dist = nullptr;
	ctx->peer_on = false;
	return MTD_OK;
}

extern "C" int mtd_stripe_peer_finish(mtd_ctx *ctx, uint32_t *rounds_out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	const int rc = sync_and_status(ctx);
	ctx->last_rounds = ctx->hflags->round_out[1];
	if (ctx->last_async)
		ctx->last_visits = ctx->hflags->async_ctl[ACTL_VISITS];
	if (rounds_out)
		*rounds_out = ctx->last_rounds;
	if (rc == MTD_ERR_STATE && ctx->peer_root && ctx->peer_blk.p)
	{
		uint32_t mine[PEER_CTL_WORDS] = { 0 };
		cudaMemcpy(mine, ctx->peer_blk.p, sizeof(mine), cudaMemcpyDeviceToHost);
		uint32_t behind = 0;
		for (int r = 0; r < ctx->peer_world; r++)
			behind += (int32_t)(mine[PEER_SLOT_A + r] - mine[PEER_SLOT_A + ctx->peer_rank]) < 0 ? 1u : 0u;
		return fail(ctx, MTD_ERR_STATE, "peer-coupled stripes: another rank did not reach a barrier in time (20 s to enter the run, 4 s per round: not launched, or died); rank %d of %d, phases passed %u, ranks behind %u, rounds %u", ctx->peer_rank, ctx->peer_world, mine[PEER_EPOCH], behind, ctx->last_rounds);
	}
	return rc;
}

extern "C" int mtd_grid_set_flags(mtd_ctx *ctx, uint32_t tile, uint16_t flags)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->have_grid || ctx->stripe)
		return fail(ctx, MTD_ERR_STATE, "no grid uploaded (or the context holds a row stripe)");
	if (tile >= ctx->T)
		return fail(ctx, MTD_ERR_ARG, "tile outside the grid");
	if (!ctx->h_blocking.empty())
	{
		const uint8_t nb = (flags & F_BLOCKING) ? 1 : 0;
		if (nb != ctx->h_blocking[tile])
		{
			ctx->h_blocking[tile] = nb;
			ctx->inc_dirty++;
			ctx->inc_tile = tile;
		}
	}
	else
		ctx->inc_layers_ok = false; // no host mirror of the blocking bits: the next update is a full one
	// the value travels as a kernel argument: no staging memory whose reuse would need a synchronisation, so a scripted
	// game (flag flip, rebuild, tick, ...) stays asynchronous up to the call that returns something
	k_poke_flags<<<1, 1, 0, ctx->stream>>>(ctx->flags.p + tile, flags, ctx->small ? nullptr : ctx->blocked.p + tile, (uint8_t)((flags & F_BLOCKING) ? 1 : 0));
	LAUNCHED();
	return MTD_OK;
}

extern "C" int mtd_neighbor_distances(mtd_ctx *ctx, const uint32_t *a, const uint32_t *b, int32_t n, uint32_t *out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->have_grid || ctx->stripe)
		return fail(ctx, MTD_ERR_STATE, "no grid uploaded (or the context holds a row stripe)");
	if (n < 0 || (n && (!a || !b || !out)))
		return fail(ctx, MTD_ERR_ARG, "bad arguments");
	if (n == 0)
		return MTD_OK;
	for (int i = 0; i < n; i++)
		if (a[i] >= ctx->T || b[i] >= ctx->T)
			return fail(ctx, MTD_ERR_ARG, "tile outside the grid");
	CU(ctx->scratch_u32a.reserve(n));
	CU(ctx->scratch_u32b.reserve(n));
	CU(ctx->scratch_u32c.reserve(n));
	CU(cudaMemcpyAsync(ctx->scratch_u32a.p, a, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->scratch_u32b.p, b, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
	k_neighbor_distances<<<(n + 255) / 256, 256, 0, ctx->stream>>>(ctx->W, ctx->ox, ctx->oy, ctx->elev.p, ctx->scratch_u32a.p, ctx->scratch_u32b.p, n, ctx->scratch_u32c.p);
	LAUNCHED();
	CU(cudaMemcpyAsync(out, ctx->scratch_u32c.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return MTD_OK;
}

extern "C" int mtd_edge_weights(mtd_ctx *ctx, uint32_t *right, uint32_t *down)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->have_grid || ctx->stripe)
		return fail(ctx, MTD_ERR_STATE, "no grid uploaded (or the context holds a row stripe)");
	if (!right || !down)
		return fail(ctx, MTD_ERR_ARG, "null output");
	CU(ctx->scratch_u32a.reserve(ctx->T));
	CU(ctx->scratch_u32b.reserve(ctx->T));
	k_unpack_weights<<<grid_blocks(ctx, ctx->T, 256), 256, 0, ctx->stream>>>(ctx->W, ctx->H, ctx->elev.p, ctx->scratch_u32a.p, ctx->scratch_u32b.p);
	LAUNCHED();
	CU(cudaMemcpyAsync(right, ctx->scratch_u32a.p, ctx->T * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaMemcpyAsync(down, ctx->scratch_u32b.p, ctx->T * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return MTD_OK;
}

// =====================================================================================================
// Directions::update
// =====================================================================================================
extern "C" int mtd_directions_update_dev(mtd_ctx *ctx, const uint32_t *sources, int32_t nsrc, uint32_t extra_blocked, uint32_t *d_tiles, uint32_t *d_dist)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return directions_dev(ctx, sources, nsrc, extra_blocked, d_tiles, d_dist);
}

extern "C" int mtd_directions_layers_dev(mtd_ctx *ctx, const uint32_t *targets, int32_t n, uint32_t *d_tiles, uint32_t *d_dist)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->have_grid || ctx->stripe)
		return fail(ctx, MTD_ERR_STATE, "no grid uploaded (or the context holds a row stripe)");
	if (!targets || n < 1 || !d_dist)
		return fail(ctx, MTD_ERR_ARG, "layers need targets and a distance buffer");
	for (int i = 0; i < n; i++)
		if (targets[i] >= ctx->T)
			return fail(ctx, MTD_ERR_ARG, "target tile %u outside the grid", targets[i]);
	if (!ctx->small)
		return big_layers(ctx, targets, n, 1, MTD_M, d_tiles, d_dist); // all layers in flight together
	for (int i = 0; i < n; i++)
	{
		int rc = directions_dev(ctx, targets + i, 1, MTD_M, d_tiles ? d_tiles + (size_t)i * ctx->T : nullptr, d_dist + (size_t)i * ctx->T);
		if (rc)
			return rc;
	}
	return MTD_OK;
}

extern "C" int mtd_directions_update(mtd_ctx *ctx, const uint32_t *sources, int32_t nsrc, uint32_t extra_blocked, uint32_t *tiles_out, uint32_t *dist_out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->have_grid || ctx->stripe)
		return fail(ctx, MTD_ERR_STATE, "no grid uploaded (or the context holds a row stripe)");
	CU(ctx->scratch_u32a.reserve(ctx->T));
	CU(ctx->scratch_u32b.reserve(ctx->T));
	int rc = directions_dev(ctx, sources, nsrc, extra_blocked, tiles_out ? ctx->scratch_u32a.p : nullptr, ctx->scratch_u32b.p);
	if (rc)
		return rc;
	const D2HItem out[2] = { { tiles_out, ctx->scratch_u32a.p, (size_t)ctx->T * 4 }, { dist_out, ctx->scratch_u32b.p, (size_t)ctx->T * 4 } };
	return d2h_results(ctx, out, 2);
}

// =====================================================================================================
// Waypoints::update / find
// =====================================================================================================
extern "C" int mtd_waypoints_update_dev(mtd_ctx *ctx, const uint32_t *wp_tiles, int32_t n)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (n > 0 && !wp_tiles)
		return fail(ctx, MTD_ERR_ARG, "null waypoint list");
	// device-resident variant: the path buffers are library-owned; size them for long detours (8 x the half perimeter
	// per visited waypoint), never below the 4096 entries the game-sized maps use
	size_t cap = (size_t)8 * (size_t)(ctx->W + ctx->H) * (size_t)(n > 0 ? n : 1);
	if (cap < 4096)
		cap = 4096;
	if (cap > ctx->T)
		cap = ctx->T;
	if (ctx->fp_cap > cap)
		cap = ctx->fp_cap;
	return waypoints_dev(ctx, wp_tiles, n, (uint32_t)cap);
}

extern "C" int mtd_waypoints_update(mtd_ctx *ctx, const uint32_t *wp_tiles, int32_t n, uint32_t *tiles_out, uint32_t *dist_out, uint32_t *fullpath_buf, uint32_t fullpath_cap, uint32_t *fullpath_len, uint32_t *full_distance, uint32_t *agg3)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (n > 0 && !wp_tiles)
		return fail(ctx, MTD_ERR_ARG, "null waypoint list");
	int rc = waypoints_dev(ctx, wp_tiles, n, fullpath_buf ? fullpath_cap : 0u);
	if (rc || n == 0)
		return rc;
	const D2HItem out[6] = {
		{ tiles_out, ctx->ltiles.p, (size_t)n * ctx->T * 4 },
		{ dist_out, ctx->ldist.p, (size_t)n * ctx->T * 4 },
		{ fullpath_cap ? fullpath_buf : nullptr, ctx->fp_buf.p, (size_t)n * fullpath_cap * 4 },
		{ fullpath_len, ctx->fp_len.p, (size_t)n * 4 },
		{ full_distance, ctx->fp_dist.p, (size_t)n * 4 },
		{ agg3, ctx->agg.p, 12 },
	};
	return d2h_results(ctx, out, 6);
}

extern "C" int mtd_waypoints_layers_dev(mtd_ctx *ctx, uint32_t **d_tiles, uint32_t **d_dist)
{
	if (!ctx)
		return MTD_ERR_ARG;
	if (ctx->nlayers <= 0)
		return fail(ctx, MTD_ERR_STATE, "no resident layers");
	if (d_tiles)
		*d_tiles = ctx->ltiles.p;
	if (d_dist)
		*d_dist = ctx->ldist.p;
	return MTD_OK;
}

extern "C" int mtd_waypoints_find(mtd_ctx *ctx, const uint32_t *pos, const uint32_t *bits, int32_t nq, uint32_t *tile_out, uint32_t *dist_out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (ctx->nlayers <= 0 || ctx->nlayers > FIND_MAXN)
		return fail(ctx, MTD_ERR_STATE, "find needs 1..%d resident layers", FIND_MAXN);
	if (nq < 0 || (nq && (!pos || !bits || !tile_out || !dist_out)))
		return fail(ctx, MTD_ERR_ARG, "bad arguments");
	if (nq == 0)
		return MTD_OK;
	for (int i = 0; i < nq; i++)
		if (pos[i] >= ctx->T)
			return fail(ctx, MTD_ERR_ARG, "position outside the grid");
	CU(ctx->scratch_u32a.reserve(nq));
	CU(ctx->scratch_u32b.reserve(nq));
	CU(ctx->scratch_u32c.reserve(nq));
	CU(ctx->scratch_u32d.reserve(nq));
	CU(cudaMemcpyAsync(ctx->scratch_u32a.p, pos, (size_t)nq * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->scratch_u32b.p, bits, (size_t)nq * 4, cudaMemcpyHostToDevice, ctx->stream));
	k_find_batch<<<(nq + 127) / 128, 128, 0, ctx->stream>>>(layers_view(ctx), ctx->scratch_u32a.p, ctx->scratch_u32b.p, nq, ctx->scratch_u32c.p, ctx->scratch_u32d.p);
	LAUNCHED();
	CU(cudaMemcpyAsync(tile_out, ctx->scratch_u32c.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaMemcpyAsync(dist_out, ctx->scratch_u32d.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return MTD_OK;
}

extern "C" int mtd_placing_blocks(mtd_ctx *ctx, uint32_t candidate, int32_t *blocked_out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!blocked_out)
		return fail(ctx, MTD_ERR_ARG, "null output");
	if (ctx->nlayers <= 0)
		return fail(ctx, MTD_ERR_STATE, "placing_blocks needs resident waypoints (call mtd_waypoints_update first)");
	if (candidate >= ctx->T)
		return fail(ctx, MTD_ERR_ARG, "candidate outside the grid");
	// the trial source is waypoint 0 (buildingsPlacement.cpp:27); its tile is mirrored on the host in hflags->spawn[2]
	uint32_t src;
	if (ctx->inc_layers_ok && ctx->inc_n >= 1)
		src = ctx->inc_wp[0]; // host mirror of the resident waypoints
	else
	{
		CU(cudaMemcpyAsync(&ctx->hflags->spawn[2], ctx->lwp.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
		CU(cudaStreamSynchronize(ctx->stream));
		src = ctx->hflags->spawn[2];
	}
	CU(ctx->trial_dist.reserve(ctx->T));
	int rc = directions_dev(ctx, &src, 1, candidate, nullptr, ctx->trial_dist.p);
	if (rc)
		return rc;
	CU(cudaMemsetAsync(ctx->d_trial.p, 0, sizeof(int), ctx->stream));
	const int nthreads = ctx->nmon > ctx->nlayers ? ctx->nmon : ctx->nlayers;
	k_trial_check<<<(nthreads + 255) / 256, 256, 0, ctx->stream>>>(ctx->trial_dist.p, ctx->lwp.p, ctx->nlayers, ctx->m_tile.p, ctx->m_alive.p, ctx->nmon, ctx->d_trial.p);
	LAUNCHED();
	CU(cudaMemcpyAsync(&ctx->hflags->trial_blocked, ctx->d_trial.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
	rc = sync_and_status(ctx);
	*blocked_out = ctx->hflags->trial_blocked;
	return rc;
}

// =====================================================================================================
// procedural generator
// =====================================================================================================
static int mapgen_dev(mtd_ctx *ctx, const mtd_mapgen_params *p, uint64_t s1, uint64_t s2, int32_t first, int32_t nmaps, int16_t *d_elev, uint16_t *d_flags, uint32_t *d_wp, uint32_t *d_tiles, uint32_t *d_dist, uint32_t *d_fullpath, uint32_t *d_fplen, uint32_t *d_fdist, uint32_t *d_agg, const int16_t *d_elev_in = nullptr)
{
	if (!p || p->W < 1 || p->H < 1 || p->nwp < 0 || p->nwp > MTD_MAX_WAYPOINTS || nmaps < 0)
		return fail(ctx, MTD_ERR_ARG, "bad mapgen parameters");
	if (nmaps == 0)
		return MTD_OK;
	if (!d_elev || !d_flags || !d_wp)
		return fail(ctx, MTD_ERR_ARG, "elev, flags and wp_tiles outputs are required");
	const int T = p->W * p->H;
	if (p->W > 256 || p->H > 256 || T >= 65536 || mapgen_grid_smem(T) > MAX_SMEM || !small_fits(p->W, p->H, p->nwp > 0 ? p->nwp : 1))
		return fail(ctx, MTD_ERR_CAPACITY, "procedural maps must fit one CTA's shared memory (reference size is 101x101)");
	if (kernel_config(ctx, (const void *)k_mapgen_grid, 512, MAX_SMEM) < 0)
		return MTD_ERR_CUDA;
	MapgenGridParams G;
	memset(&G, 0, sizeof(G));
	G.W = p->W;
	G.H = p->H;
	G.ox = p->ox;
	G.oy = p->oy;
	G.radius = p->radius;
	G.nwp = p->nwp;
	G.s1 = s1;
	G.s2 = s2;
	G.first = first;
	G.nmaps = nmaps;
	G.elev = d_elev;
	G.flags = d_flags;
	G.wp_tiles = d_wp;
	G.status = ctx->d_status.p;
	G.elev_in = d_elev_in;
	// maps WITH waypoints: the generator runs inside k_waypoints_small (the next map's elevations are computed by the warps
	// that have no layer to relax); maps only, the caller's elevations, or shapes the fused mode cannot hold: k_mapgen_grid
	const int Rres = p->nwp > 0 ? small_resident(p->W, p->H, p->nwp) : 0;
	const bool fused = ctx->fused_mapgen && p->nwp > 0 && !d_elev_in && p->nwp <= Rres && small_gen_fits(p->W, p->H, Rres);
	if (!fused)
	{
		const size_t gsm = mapgen_grid_smem(T);
		const int per_sm = (int)(MAX_SMEM / (gsm + 1024)) > 0 ? (int)(MAX_SMEM / (gsm + 1024)) : 1;
		const int gblocks = nmaps < ctx->sms * per_sm ? nmaps : ctx->sms * per_sm;
		ProfScope prof(ctx, MTD_K_MAPGEN_GRID);
		k_mapgen_grid<<<gblocks, 512, gsm, ctx->stream>>>(G);
		LAUNCHED();
	}
	if (p->nwp == 0)
		return MTD_OK;

	WaypointsSmallParams P;
	memset(&P, 0, sizeof(P));
	P.W = p->W;
	P.H = p->H;
	P.n = p->nwp;
	P.nmaps = nmaps;
	P.nsrc = 1;
	P.elev = d_elev;
	P.flags = d_flags;
	P.wp = d_wp;
	P.extra_blocked = MTD_M;
	P.tiles = d_tiles;
	P.dist = d_dist;
	P.fullpath = d_fullpath;
	P.cap = d_fullpath ? p->fullpath_cap : 0u;
	P.fullpath_len = p->nwp <= FIND_MAXN ? d_fplen : nullptr;
	P.full_distance = d_fdist;
	P.agg = d_agg;
	P.status = ctx->d_status.p;
	if (fused)
	{
		P.gen = 1;
		P.gen_warps = ctx->gen_warps;
		P.gen_ox = p->ox;
		P.gen_oy = p->oy;
		P.gen_radius = p->radius;
		P.gen_first = first;
		P.gen_s1 = s1;
		P.gen_s2 = s2;
	}
	if (P.fullpath_len && (!d_fdist || !d_agg))
		return fail(ctx, MTD_ERR_ARG, "fullpath_len, full_distance and agg outputs go together");
	if (p->nwp > SMALL_MAX_RESIDENT && P.fullpath_len && !d_dist)
		return fail(ctx, MTD_ERR_ARG, "more than %d waypoints need the dist output for the path walk", SMALL_MAX_RESIDENT);
	return launch_waypoints_small(ctx, P);
}

extern "C" int mtd_mapgen_batch_dev(mtd_ctx *ctx, const mtd_mapgen_params *p, uint64_t s1, uint64_t s2, int32_t first, int32_t nmaps, int16_t *d_elev, uint16_t *d_flags, uint32_t *d_wp_tiles, uint32_t *d_tiles, uint32_t *d_dist, uint32_t *d_fullpath, uint32_t *d_fullpath_len, uint32_t *d_full_distance, uint32_t *d_agg)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return mapgen_dev(ctx, p, s1, s2, first, nmaps, d_elev, d_flags, d_wp_tiles, d_tiles, d_dist, d_fullpath, d_fullpath_len, d_full_distance, d_agg);
}

extern "C" int mtd_mapgen_batch(mtd_ctx *ctx, const mtd_mapgen_params *p, uint64_t s1, uint64_t s2, int32_t first, int32_t nmaps, int16_t *elev, uint16_t *flags, uint32_t *wp_tiles, uint32_t *tiles, uint32_t *dist, uint32_t *fullpath, uint32_t *fullpath_len, uint32_t *full_distance, uint32_t *agg)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!p || nmaps < 0)
		return fail(ctx, MTD_ERR_ARG, "bad mapgen parameters");
	const size_t T = (size_t)p->W * p->H;
	const int n = p->nwp;
	const size_t cap = fullpath ? p->fullpath_cap : 0;
	// Chunked so the device staging stays bounded (layers are 8 bytes per tile per waypoint), and double buffered: the
	// kernels of chunk i + 1 run on the context's stream while chunk i drains to the host on a copy stream. The staging
	// buffers live in the context (no cudaMalloc / cudaFree per call).
	const int chunk = ctx->sms * 4;
	if (!ctx->copy_stream)
	{
		CU(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
		for (int b = 0; b < 2; b++)
		{
			CU(cudaEventCreateWithFlags(&ctx->mg_done[b], cudaEventDisableTiming));
			CU(cudaEventCreateWithFlags(&ctx->mg_copied[b], cudaEventDisableTiming));
		}
	}
	for (int b = 0; b < 2; b++)
	{
		MapgenStage &S = ctx->mg[b];
		CU(S.elev.reserve(chunk * T));
		CU(S.flags.reserve(chunk * T));
		CU(S.wp.reserve((size_t)chunk * (n ? n : 1)));
		if (n)
		{
			if (tiles)
				CU(S.tiles.reserve((size_t)chunk * n * T));
			if (dist || n > SMALL_MAX_RESIDENT)
				CU(S.dist.reserve((size_t)chunk * n * T));
			if (cap)
				CU(S.fp.reserve((size_t)chunk * n * cap));
			CU(S.fplen.reserve((size_t)chunk * n));
			CU(S.fdist.reserve((size_t)chunk * n));
			CU(S.agg.reserve((size_t)chunk * 3));
		}
	}
	int rc = MTD_OK;
	int i = 0;
	cudaStream_t cs = ctx->copy_stream;
	for (int done = 0; done < nmaps && rc == MTD_OK; done += chunk, i++)
	{
		const int m = nmaps - done < chunk ? nmaps - done : chunk;
		const int b = i & 1;
		MapgenStage &S = ctx->mg[b];
		if (i >= 2)
			CU(cudaStreamWaitEvent(ctx->stream, ctx->mg_copied[b], 0)); // the buffer has drained
		rc = mapgen_dev(ctx, p, s1, s2, first + done, m, S.elev.p, S.flags.p, S.wp.p, tiles ? S.tiles.p : nullptr, (dist || n > SMALL_MAX_RESIDENT) ? S.dist.p : nullptr, cap ? S.fp.p : nullptr, S.fplen.p, S.fdist.p, S.agg.p);
		if (rc)
			break;
		CU(cudaEventRecord(ctx->mg_done[b], ctx->stream));
		CU(cudaStreamWaitEvent(cs, ctx->mg_done[b], 0));
		if (elev)
			CU(cudaMemcpyAsync(elev + (size_t)done * T, S.elev.p, (size_t)m * T * 2, cudaMemcpyDeviceToHost, cs));
		if (flags)
			CU(cudaMemcpyAsync(flags + (size_t)done * T, S.flags.p, (size_t)m * T * 2, cudaMemcpyDeviceToHost, cs));
		if (wp_tiles && n)
			CU(cudaMemcpyAsync(wp_tiles + (size_t)done * n, S.wp.p, (size_t)m * n * 4, cudaMemcpyDeviceToHost, cs));
		if (tiles && n)
			CU(cudaMemcpyAsync(tiles + (size_t)done * n * T, S.tiles.p, (size_t)m * n * T * 4, cudaMemcpyDeviceToHost, cs));
		if (dist && n)
			CU(cudaMemcpyAsync(dist + (size_t)done * n * T, S.dist.p, (size_t)m * n * T * 4, cudaMemcpyDeviceToHost, cs));
		if (fullpath && cap && n)
			CU(cudaMemcpyAsync(fullpath + (size_t)done * n * cap, S.fp.p, (size_t)m * n * cap * 4, cudaMemcpyDeviceToHost, cs));
		if (fullpath_len && n)
			CU(cudaMemcpyAsync(fullpath_len + (size_t)done * n, S.fplen.p, (size_t)m * n * 4, cudaMemcpyDeviceToHost, cs));
		if (full_distance && n)
			CU(cudaMemcpyAsync(full_distance + (size_t)done * n, S.fdist.p, (size_t)m * n * 4, cudaMemcpyDeviceToHost, cs));
		if (agg && n)
			CU(cudaMemcpyAsync(agg + (size_t)done * 3, S.agg.p, (size_t)m * 3 * 4, cudaMemcpyDeviceToHost, cs));
		CU(cudaEventRecord(ctx->mg_copied[b], cs));
	}
	const cudaError_t ce = cudaStreamSynchronize(cs);
	const int rs = sync_and_status(ctx);
	if (rc == MTD_OK && ce != cudaSuccess)
		return fail(ctx, MTD_ERR_CUDA, "device to host copy failed: %s", cudaGetErrorString(ce));
	return rc ? rc : rs;
}

__global__ void k_elevation_samples(uint64_t s1, uint64_t s2, const float *xy, int n, float *out)
{
	__shared__ float2 grad[128];
	__shared__ noise::ElevationParams ep;
	noise::fill_gradients(grad);
	if (threadIdx.x == 0)
	{
		Rng g;
		g.s0 = s1;
		g.s1 = s2;
		const float r1 = g.chance();
		const float r2 = g.chance();
		ep.off_x = __fmul_rn(r1, 1000.0f);
		ep.off_y = __fmul_rn(r2, 1000.0f);
		ep.seed = (int)(uint32_t)g.next();
		ep.bounding = noise::fractal_bounding(4, 0.55f);
	}
	__syncthreads();
	const noise::ElevationParams lp = ep;
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
		out[i] = noise::elevation(grad, lp, xy[i * 2 + 0], xy[i * 2 + 1]);
}

// Terrain SDF over the marching-cubes lattice: Maker::sdf (sources/map/generate.cpp:85-96) at the lattice positions of
// MarchingCubesCreateConfig::position (externals/cage/sources/libcore/mesh/marchingCubes.cpp:247-257), stored x fastest,
// then y, then z (marchingCubes.cpp:149-165). elevation() depends on (x, z) only, so one thread owns a lattice column:
// one noise evaluation serves the whole column (the reference evaluates the noise at every lattice point).
__global__ void k_sdf_lattice(uint64_t s1, uint64_t s2, int W, int H, int ox, int oy, const uint16_t *__restrict__ flags, float ax, float ay, float az, float bx, float by, float bz, int rx, int ry, int rz, float *__restrict__ out)
{
	__shared__ float2 grad[128];
	__shared__ noise::ElevationParams ep;
	noise::fill_gradients(grad);
	if (threadIdx.x == 0)
	{
		Rng g;
		g.s0 = s1;
		g.s1 = s2;
		const float r1 = g.chance();
		const float r2 = g.chance();
		ep.off_x = __fmul_rn(r1, 1000.0f);
		ep.off_y = __fmul_rn(r2, 1000.0f);
		ep.seed = (int)(uint32_t)g.next();
		ep.bounding = noise::fractal_bounding(4, 0.55f);
	}
	__syncthreads();
	const noise::ElevationParams lp = ep;
	const float dx = __fsub_rn((float)rx, 5.0f), dy = __fsub_rn((float)ry, 5.0f), dz = __fsub_rn((float)rz, 5.0f);
	const float sx = __fsub_rn(bx, ax), sy = __fsub_rn(by, ay), sz = __fsub_rn(bz, az);
	const int columns = rx * rz;
	for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < columns; c += gridDim.x * blockDim.x)
	{
		const int z = c / rx, x = c - z * rx;
		const float px = __fadd_rn(__fmul_rn(sx, __fdiv_rn(__fsub_rn((float)x, 2.0f), dx)), ax);
		const float pz = __fadd_rn(__fmul_rn(sz, __fdiv_rn(__fsub_rn((float)z, 2.0f), dz)), az);
		const float e = noise::elevation(grad, lp, px, pz);
		// Grid::index(Vec3) -> index(Vec2(x, z)) -> Vec2i(pos + 200.5) - 200 (sources/map/grid.cpp:108-116)
		const int tx = __float2int_rz(__fadd_rn(px, 200.5f)) - 200 + ox;
		const int tz = __float2int_rz(__fadd_rn(pz, 200.5f)) - 200 + oy;
		bool cliff = false; // invalid land tiles get a wall below the water line (generate.cpp:89-94)
		if (tx >= 0 && tx < W && tz >= 0 && tz < H)
		{
			const uint16_t f = flags[tz * W + tx];
			cliff = (f & F_INVALID) && !(f & F_WATER);
		}
		for (int y = 0; y < ry; y++)
		{
			const float py = __fadd_rn(__fmul_rn(sy, __fdiv_rn(__fsub_rn((float)y, 2.0f), dy)), ay);
			float d = __fsub_rn(e, py);
			if (cliff)
			{
				float t = __fdiv_rn(__fsub_rn(py, -7.5f), __fsub_rn(-6.5f, -7.5f)); // find(y, -7.5, -6.5), sources/map/generate.h:37-40
				t = t > 0.0f ? t : 0.0f; // saturate = clamp(t, 0, 1) = min(max(t, 0), 1) with cage's a < b ? a : b selects
				t = t < 1.0f ? t : 1.0f;
				d = __fadd_rn(d, t);
			}
			__stcs(out + ((size_t)z * ry + y) * rx + x, d);
		}
	}
}

extern "C" int mtd_sdf_lattice(mtd_ctx *ctx, uint64_t s1, uint64_t s2, int32_t W, int32_t H, int32_t ox, int32_t oy, const uint16_t *flags, const float *box_a, const float *box_b, const int32_t *res, float *out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (W < 1 || H < 1 || !flags || !box_a || !box_b || !res || !out)
		return fail(ctx, MTD_ERR_ARG, "sdf_lattice needs flags, a box, a resolution and an output");
	if (res[0] < 6 || res[1] < 6 || res[2] < 6) // marchingCubes.cpp:249-251
		return fail(ctx, MTD_ERR_ARG, "lattice resolution must exceed 5 in every dimension");
	const size_t T = (size_t)W * H, n = (size_t)res[0] * res[1] * res[2];
	CU(ctx->scratch_flags.reserve(T));
	CU(ctx->scratch_f.reserve(n));
	CU(cudaMemcpyAsync(ctx->scratch_flags.p, flags, T * 2, cudaMemcpyHostToDevice, ctx->stream));
	const int columns = res[0] * res[2];
	k_sdf_lattice<<<(columns + 127) / 128, 128, 0, ctx->stream>>>(s1, s2, W, H, ox, oy, ctx->scratch_flags.p, box_a[0], box_a[1], box_a[2], box_b[0], box_b[1], box_b[2], res[0], res[1], res[2], ctx->scratch_f.p);
	LAUNCHED();
	CU(cudaMemcpyAsync(out, ctx->scratch_f.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	return sync_and_status(ctx);
}

extern "C" int mtd_grid_from_elevations(mtd_ctx *ctx, const mtd_mapgen_params *p, int32_t nmaps, const int16_t *elev, uint16_t *flags_out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!p || nmaps < 0 || (nmaps && (!elev || !flags_out)))
		return fail(ctx, MTD_ERR_ARG, "grid_from_elevations needs elevations in and flags out");
	if (nmaps == 0)
		return MTD_OK;
	const size_t T = (size_t)p->W * p->H;
	mtd_mapgen_params q = *p;
	q.nwp = 0; // spawn picking belongs to newWaypoints (sources/map/paths.cpp:185-196), not newGrid
	const int chunk = 1024;
	CU(ctx->scratch_elev_in.reserve((size_t)chunk * T));
	CU(ctx->scratch_elev.reserve((size_t)chunk * T));
	CU(ctx->scratch_flags.reserve((size_t)chunk * T));
	CU(ctx->scratch_u32a.reserve(64));
	for (int first = 0; first < nmaps; first += chunk)
	{
		const int cnt = nmaps - first < chunk ? nmaps - first : chunk;
		CU(cudaMemcpyAsync(ctx->scratch_elev_in.p, elev + (size_t)first * T, (size_t)cnt * T * 2, cudaMemcpyHostToDevice, ctx->stream));
		int rc = mapgen_dev(ctx, &q, 0, 0, 0, cnt, ctx->scratch_elev.p, ctx->scratch_flags.p, ctx->scratch_u32a.p, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, ctx->scratch_elev_in.p);
		if (rc)
			return rc;
		CU(cudaMemcpyAsync(flags_out + (size_t)first * T, ctx->scratch_flags.p, (size_t)cnt * T * 2, cudaMemcpyDeviceToHost, ctx->stream));
		rc = sync_and_status(ctx);
		if (rc)
			return rc;
	}
	return MTD_OK;
}

extern "C" int mtd_elevation_samples(mtd_ctx *ctx, uint64_t s1, uint64_t s2, const float *xy, int32_t n, float *out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (n < 0 || (n && (!xy || !out)))
		return fail(ctx, MTD_ERR_ARG, "bad arguments");
	if (n == 0)
		return MTD_OK;
	CU(ctx->scratch_f.reserve((size_t)n * 3));
	CU(cudaMemcpyAsync(ctx->scratch_f.p, xy, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
	k_elevation_samples<<<grid_blocks(ctx, n, 256), 256, 0, ctx->stream>>>(s1, s2, ctx->scratch_f.p, n, ctx->scratch_f.p + (size_t)n * 2);
	LAUNCHED();
	CU(cudaMemcpyAsync(out, ctx->scratch_f.p + (size_t)n * 2, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return MTD_OK;
}

// =====================================================================================================
// monsters
// =====================================================================================================
extern "C" int mtd_monsters_upload(mtd_ctx *ctx, int32_t n, const uint32_t *tile, const uint32_t *tileStart, const uint32_t *tileEnd, const uint32_t *timeStart, const uint32_t *timeEnd, const uint32_t *bits, const uint32_t *timeToArrive, const uint8_t *dotflags, const uint32_t *damage, const uint8_t *alive)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (n < 0 || (n && (!tile || !tileStart || !tileEnd || !timeStart || !timeEnd || !bits || !timeToArrive)))
		return fail(ctx, MTD_ERR_ARG, "bad monster arrays");
	ctx->nmon = n;
	if (n == 0)
		return MTD_OK;
	const size_t N = (size_t)n;
	CU(ctx->m_tile.reserve(N));
	CU(ctx->m_tileStart.reserve(N));
	CU(ctx->m_tileEnd.reserve(N));
	CU(ctx->m_timeStart.reserve(N));
	CU(ctx->m_timeEnd.reserve(N));
	CU(ctx->m_bits.reserve(N));
	CU(ctx->m_tta.reserve(N));
	CU(ctx->m_damage.reserve(N));
	CU(ctx->m_dot.reserve(N));
	CU(ctx->m_alive.reserve(N));
	CU(cudaMemcpyAsync(ctx->m_tile.p, tile, N * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->m_tileStart.p, tileStart, N * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->m_tileEnd.p, tileEnd, N * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->m_timeStart.p, timeStart, N * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->m_timeEnd.p, timeEnd, N * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->m_bits.p, bits, N * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->m_tta.p, timeToArrive, N * 4, cudaMemcpyHostToDevice, ctx->stream));
	ctx->m_has_dot = dotflags != nullptr;
	ctx->m_has_dots = false; // a new population: its damage-over-time state has to be uploaded again
	ctx->m_has_damage = damage != nullptr;
	if (dotflags)
		CU(cudaMemcpyAsync(ctx->m_dot.p, dotflags, N, cudaMemcpyHostToDevice, ctx->stream));
	if (damage)
		CU(cudaMemcpyAsync(ctx->m_damage.p, damage, N * 4, cudaMemcpyHostToDevice, ctx->stream));
	if (alive)
		CU(cudaMemcpyAsync(ctx->m_alive.p, alive, N, cudaMemcpyHostToDevice, ctx->stream));
	else
		CU(cudaMemsetAsync(ctx->m_alive.p, 1, N, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return MTD_OK;
}

extern "C" int mtd_monsters_download(mtd_ctx *ctx, uint32_t *tile, uint32_t *tileStart, uint32_t *tileEnd, uint32_t *timeStart, uint32_t *timeEnd, uint32_t *bits, uint8_t *alive)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	const size_t N = (size_t)ctx->nmon;
	if (N == 0)
		return MTD_OK;
	if (tile)
		CU(cudaMemcpyAsync(tile, ctx->m_tile.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (tileStart)
		CU(cudaMemcpyAsync(tileStart, ctx->m_tileStart.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (tileEnd)
		CU(cudaMemcpyAsync(tileEnd, ctx->m_tileEnd.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (timeStart)
		CU(cudaMemcpyAsync(timeStart, ctx->m_timeStart.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (timeEnd)
		CU(cudaMemcpyAsync(timeEnd, ctx->m_timeEnd.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (bits)
		CU(cudaMemcpyAsync(bits, ctx->m_bits.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (alive)
		CU(cudaMemcpyAsync(alive, ctx->m_alive.p, N, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return MTD_OK;
}

static int monsters_tick_dev(mtd_ctx *ctx, uint32_t gameTime)
{
	if (ctx->nlayers <= 0 || ctx->nlayers > FIND_MAXN)
		return fail(ctx, MTD_ERR_STATE, "monster tick needs 1..%d resident waypoint layers", FIND_MAXN);
	CU(cudaMemsetAsync(ctx->d_damage.p, 0, 8, ctx->stream));
	CU(cudaMemsetAsync(ctx->d_small.p, 0, 4, ctx->stream));
	if (ctx->nmon == 0)
		return MTD_OK;
	ProfScope prof(ctx, MTD_K_MONSTERS_TICK);
	k_monsters_tick<<<(ctx->nmon + 255) / 256, 256, 0, ctx->stream>>>(layers_view(ctx), monsters_view(ctx), gameTime, ctx->d_damage.p, ctx->d_small.p);
	LAUNCHED();
	return MTD_OK;
}

extern "C" int mtd_monsters_tick_dev(mtd_ctx *ctx, uint32_t gameTime)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return monsters_tick_dev(ctx, gameTime);
}

extern "C" int mtd_monsters_tick(mtd_ctx *ctx, uint32_t gameTime, int64_t *damage_out, uint32_t *alive_out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	int rc = monsters_tick_dev(ctx, gameTime);
	if (rc)
		return rc;
	CU(cudaMemcpyAsync(&ctx->hflags->damage, ctx->d_damage.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaMemcpyAsync(&ctx->hflags->alive, ctx->d_small.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
	rc = sync_and_status(ctx); // ST_MONSTER_STUCK and friends
	if (damage_out)
		*damage_out = (int64_t)ctx->hflags->damage;
	if (alive_out)
		*alive_out = ctx->hflags->alive;
	return rc;
}

// ---- damage-over-time tick ----
static DotsView dots_view(mtd_ctx *ctx)
{
	DotsView D;
	D.n = ctx->nmon;
	D.duration = ctx->m_dot_duration.p;
	D.damage = ctx->m_dot_damage.p;
	D.resist = ctx->m_has_resist ? ctx->m_resist.p : nullptr;
	D.life = ctx->m_life.p;
	D.money = ctx->m_has_money ? ctx->m_money.p : nullptr;
	D.alive = ctx->m_alive.p;
	D.dotflags = ctx->m_dot.p;
	return D;
}

extern "C" int mtd_monsters_dots_upload(mtd_ctx *ctx, const uint32_t *duration, const uint32_t *damage, const uint8_t *resist, const int64_t *life, const uint32_t *money)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	const size_t n = (size_t)ctx->nmon;
	if (n == 0)
		return fail(ctx, MTD_ERR_STATE, "upload the monsters first (mtd_monsters_upload)");
	if (!duration || !damage || !life)
		return fail(ctx, MTD_ERR_ARG, "dots need duration, damage and life arrays");
	CU(ctx->m_dot_duration.reserve(5 * n));
	CU(ctx->m_dot_damage.reserve(5 * n));
	CU(ctx->m_life.reserve(n));
	CU(ctx->m_resist.reserve(n));
	CU(ctx->m_money.reserve(n));
	CU(ctx->m_dot.reserve(n));
	CU(cudaMemcpyAsync(ctx->m_dot_duration.p, duration, 5 * n * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->m_dot_damage.p, damage, 5 * n * 4, cudaMemcpyHostToDevice, ctx->stream));
	CU(cudaMemcpyAsync(ctx->m_life.p, life, n * 8, cudaMemcpyHostToDevice, ctx->stream));
	ctx->m_has_resist = resist != nullptr;
	ctx->m_has_money = money != nullptr;
	if (resist)
		CU(cudaMemcpyAsync(ctx->m_resist.p, resist, n, cudaMemcpyHostToDevice, ctx->stream));
	if (money)
		CU(cudaMemcpyAsync(ctx->m_money.p, money, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	ctx->m_has_dots = true;
	ctx->m_has_dot = true; // from now on the steering tick reads the fire / water bits derived from the dots
	k_dots_flags<<<(ctx->nmon + 255) / 256, 256, 0, ctx->stream>>>(dots_view(ctx));
	LAUNCHED();
	CU(cudaStreamSynchronize(ctx->stream));
	return MTD_OK;
}

static int dots_tick_dev(mtd_ctx *ctx)
{
	if (!ctx->m_has_dots || ctx->nmon == 0)
		return fail(ctx, MTD_ERR_STATE, "no dots uploaded (mtd_monsters_dots_upload)");
	CU(cudaMemsetAsync(ctx->d_damage.p, 0, 8, ctx->stream));
	CU(cudaMemsetAsync(ctx->d_small.p + 4, 0, 4, ctx->stream));
	k_monsters_dots<<<(ctx->nmon + 255) / 256, 256, 0, ctx->stream>>>(dots_view(ctx), ctx->d_damage.p, ctx->d_small.p + 4);
	LAUNCHED();
	return MTD_OK;
}

extern "C" int mtd_monsters_dots_tick_dev(mtd_ctx *ctx)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return dots_tick_dev(ctx);
}

extern "C" int mtd_monsters_dots_tick(mtd_ctx *ctx, int64_t *money_out, uint32_t *killed_out)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	int rc = dots_tick_dev(ctx);
	if (rc)
		return rc;
	CU(cudaMemcpyAsync(&ctx->hflags->damage, ctx->d_damage.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaMemcpyAsync(&ctx->hflags->alive, ctx->d_small.p + 4, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	if (money_out)
		*money_out = (int64_t)ctx->hflags->damage;
	if (killed_out)
		*killed_out = ctx->hflags->alive;
	return MTD_OK;
}

extern "C" int mtd_monsters_dots_download(mtd_ctx *ctx, uint32_t *duration, uint32_t *damage, int64_t *life, uint8_t *alive)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->m_has_dots)
		return fail(ctx, MTD_ERR_STATE, "no dots uploaded (mtd_monsters_dots_upload)");
	const size_t n = (size_t)ctx->nmon;
	if (duration)
		CU(cudaMemcpyAsync(duration, ctx->m_dot_duration.p, 5 * n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (damage)
		CU(cudaMemcpyAsync(damage, ctx->m_dot_damage.p, 5 * n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (life)
		CU(cudaMemcpyAsync(life, ctx->m_life.p, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
	if (alive)
		CU(cudaMemcpyAsync(alive, ctx->m_alive.p, n, cudaMemcpyDeviceToHost, ctx->stream));
	return sync_and_status(ctx);
}

extern "C" int mtd_monsters_positions_dev(mtd_ctx *ctx, uint32_t gameTime, float *d_out3n)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (!ctx->have_grid || ctx->stripe)
		return fail(ctx, MTD_ERR_STATE, "no grid uploaded (or the context holds a row stripe)");
	if (ctx->nmon == 0)
		return MTD_OK;
	if (!d_out3n)
		return fail(ctx, MTD_ERR_ARG, "null output");
	k_monsters_positions<<<(ctx->nmon + 255) / 256, 256, 0, ctx->stream>>>(layers_view(ctx), monsters_view(ctx), gameTime, d_out3n);
	LAUNCHED();
	return MTD_OK;
}

extern "C" int mtd_monsters_positions(mtd_ctx *ctx, uint32_t gameTime, float *out3n)
{
	if (!ctx)
		return MTD_ERR_ARG;
	{
		std::lock_guard<std::mutex> lk(ctx->mu);
		DeviceGuard g(ctx->device);
		if (ctx->nmon == 0)
			return MTD_OK;
		if (!out3n)
			return fail(ctx, MTD_ERR_ARG, "null output");
		CU(ctx->m_pos.reserve((size_t)ctx->nmon * 3));
	}
	int rc = mtd_monsters_positions_dev(ctx, gameTime, ctx->m_pos.p);
	if (rc)
		return rc;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	CU(cudaMemcpyAsync(out3n, ctx->m_pos.p, (size_t)ctx->nmon * 12, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return MTD_OK;
}

extern "C" int mtd_monster_spawn(mtd_ctx *ctx, uint32_t gameTime, float speed, uint32_t *out3)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (ctx->nlayers <= 0 || ctx->nlayers > FIND_MAXN)
		return fail(ctx, MTD_ERR_STATE, "spawn needs 1..%d resident waypoint layers", FIND_MAXN);
	if (!out3)
		return fail(ctx, MTD_ERR_ARG, "null output");
	// minDistanceSpawner is agg[2] of the last update
	CU(cudaMemcpyAsync(&ctx->hflags->spawn[0], ctx->agg.p + 2, 4, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	const uint32_t spawner = ctx->hflags->spawn[0];
	k_monster_spawn<<<1, 1, 0, ctx->stream>>>(layers_view(ctx), spawner, gameTime, speed, ctx->d_small.p + 4);
	LAUNCHED();
	CU(cudaMemcpyAsync(ctx->hflags->spawn, ctx->d_small.p + 4, 12, cudaMemcpyDeviceToHost, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	memcpy(out3, ctx->hflags->spawn, 12);
	return MTD_OK;
}

// =====================================================================================================
// towers: monster spatial index + attack pass (sources/game/spatial.cpp:24-34, sources/game/attacks.cpp:132-160)
// =====================================================================================================
extern "C" int mtd_monsters_class_upload(mtd_ctx *ctx, const uint8_t *monsterClass)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (ctx->nmon <= 0)
		return fail(ctx, MTD_ERR_STATE, "upload the monsters first (mtd_monsters_upload)");
	ctx->m_has_class = monsterClass != nullptr;
	if (monsterClass)
	{
		CU(ctx->m_class.reserve((size_t)ctx->nmon));
		CU(cudaMemcpyAsync(ctx->m_class.p, monsterClass, (size_t)ctx->nmon, cudaMemcpyHostToDevice, ctx->stream));
		CU(cudaStreamSynchronize(ctx->stream));
	}
	return MTD_OK;
}

extern "C" int mtd_towers_upload(mtd_ctx *ctx, int32_t n, const uint32_t *tile, const uint32_t *damage, const uint32_t *overTime, const uint32_t *firingPeriod, const float *firingRange, const float *splashRadius,
								 const uint32_t *manaCost, const uint8_t *invalidClasses, const uint8_t *element, const uint8_t *bonus, const uint8_t *targeting, const uint32_t *firingDelay, const uint32_t *mana)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (n < 0 || (n && (!tile || !damage || !overTime || !firingPeriod || !firingRange || !splashRadius || !manaCost || !invalidClasses || !element || !bonus || !targeting || !firingDelay || !mana)))
		return fail(ctx, MTD_ERR_ARG, "bad tower arrays");
	if (!ctx->have_grid || ctx->stripe)
		return fail(ctx, MTD_ERR_STATE, "no grid uploaded (or the context holds a row stripe)");
	for (int i = 0; i < n; i++)
		if (tile[i] >= ctx->T || element[i] >= 5 || bonus[i] > 6 || targeting[i] > 6)
			return fail(ctx, MTD_ERR_ARG, "tower %d: tile outside the grid or element / bonus / targeting out of range", i);
	ctx->ntowers = n;
	if (n == 0)
		return MTD_OK;
	const size_t N = (size_t)n;
	CU(ctx->t_tile.reserve(N));
	CU(ctx->t_damage.reserve(N));
	CU(ctx->t_overTime.reserve(N));
	CU(ctx->t_period.reserve(N));
	CU(ctx->t_manaCost.reserve(N));
	CU(ctx->t_delay.reserve(N));
	CU(ctx->t_mana.reserve(N));
	CU(ctx->t_hits.reserve(N));
	CU(ctx->t_range.reserve(N));
	CU(ctx->t_splash.reserve(N));
	CU(ctx->t_invalid.reserve(N));
	CU(ctx->t_element.reserve(N));
	CU(ctx->t_bonus.reserve(N));
	CU(ctx->t_targeting.reserve(N));
	CU(ctx->t_random.reserve(N));
	CU(ctx->t_target.reserve(N));
	const cudaMemcpyKind k = cudaMemcpyHostToDevice;
	CU(cudaMemcpyAsync(ctx->t_tile.p, tile, N * 4, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_damage.p, damage, N * 4, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_overTime.p, overTime, N * 4, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_period.p, firingPeriod, N * 4, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_manaCost.p, manaCost, N * 4, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_delay.p, firingDelay, N * 4, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_mana.p, mana, N * 4, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_range.p, firingRange, N * 4, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_splash.p, splashRadius, N * 4, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_invalid.p, invalidClasses, N, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_element.p, element, N, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_bonus.p, bonus, N, k, ctx->stream));
	CU(cudaMemcpyAsync(ctx->t_targeting.p, targeting, N, k, ctx->stream));
	CU(cudaMemsetAsync(ctx->t_target.p, 0xFF, N * 4, ctx->stream));
	CU(cudaMemsetAsync(ctx->t_hits.p, 0, N * 4, ctx->stream));
	CU(cudaStreamSynchronize(ctx->stream));
	return MTD_OK;
}

static int towers_attack_dev(mtd_ctx *ctx, uint32_t gameTime, const uint64_t *random_words)
{
	if (!ctx->have_grid || ctx->stripe)
		return fail(ctx, MTD_ERR_STATE, "no grid uploaded (or the context holds a row stripe)");
	if (ctx->nmon <= 0 || !ctx->m_has_dots)
		return fail(ctx, MTD_ERR_STATE, "the attack pass needs the monsters and their damage-over-time state (mtd_monsters_upload, mtd_monsters_dots_upload)");
	if (ctx->ntowers == 0)
		return MTD_OK;
	const size_t N = (size_t)ctx->nmon, B = ctx->T + 1;
	CU(ctx->m_pos.reserve(N * 3));
	CU(ctx->sp_count.reserve(B));
	CU(ctx->sp_start.reserve(B));
	CU(ctx->sp_binof.reserve(N));
	CU(ctx->sp_rank.reserve(N));
	CU(ctx->sp_sorted.reserve(N));
	CU(ctx->sp_xz.reserve(N));
	CU(cudaMemsetAsync(ctx->sp_count.p, 0, B * 4, ctx->stream));
	if (random_words)
	{
		CU(cudaMemcpyAsync(ctx->t_random.p, random_words, (size_t)ctx->ntowers * 8, cudaMemcpyHostToDevice, ctx->stream));
	}
	SpatialView S;
	S.W = ctx->W;
	S.H = ctx->H;
	S.ox = ctx->ox;
	S.oy = ctx->oy;
	S.n = ctx->nmon;
	S.pos = ctx->m_pos.p;
	S.bin_count = ctx->sp_count.p;
	S.bin_start = ctx->sp_start.p;
	S.bin_of = ctx->sp_binof.p;
	S.rank = ctx->sp_rank.p;
	S.sorted_idx = ctx->sp_sorted.p;
	S.sorted_xz = ctx->sp_xz.p;
	TowersView Tw;
	Tw.n = ctx->ntowers;
	Tw.tile = ctx->t_tile.p;
	Tw.damage = ctx->t_damage.p;
	Tw.overTime = ctx->t_overTime.p;
	Tw.firingPeriod = ctx->t_period.p;
	Tw.manaCost = ctx->t_manaCost.p;
	Tw.firingRange = ctx->t_range.p;
	Tw.splashRadius = ctx->t_splash.p;
	Tw.invalidClasses = ctx->t_invalid.p;
	Tw.element = ctx->t_element.p;
	Tw.bonus = ctx->t_bonus.p;
	Tw.targeting = ctx->t_targeting.p;
	Tw.firingDelay = ctx->t_delay.p;
	Tw.mana = ctx->t_mana.p;
	Tw.random_words = random_words ? ctx->t_random.p : nullptr;
	Tw.target = ctx->t_target.p;
	Tw.hits = ctx->t_hits.p;
	TargetStats G;
	G.monsterClass = ctx->m_has_class ? ctx->m_class.p : nullptr;
	G.timeToArrive = ctx->m_tta.p;
	G.life = ctx->m_life.p;
	G.dot_duration = ctx->m_dot_duration.p;
	G.dot_damage = ctx->m_dot_damage.p;
	const int mb = (ctx->nmon + 255) / 256;
	k_spatial_bin<<<mb, 256, 0, ctx->stream>>>(layers_view(ctx), monsters_view(ctx), S, gameTime);
	LAUNCHED();
	k_spatial_scan<<<1, 1024, 0, ctx->stream>>>(S);
	LAUNCHED();
	k_spatial_scatter<<<mb, 256, 0, ctx->stream>>>(S);
	LAUNCHED();
	k_towers_attack<<<(ctx->ntowers + 3) / 4, 128, 0, ctx->stream>>>(Tw, S, G, ctx->W, ctx->ox, ctx->oy);
	LAUNCHED();
	// MonsterComponent::affected(Fire / Water) is read live by the next steering tick (monsters.cpp:13-16, 92-95)
	k_dots_flags<<<mb, 256, 0, ctx->stream>>>(dots_view(ctx));
	LAUNCHED();
	return MTD_OK;
}

extern "C" int mtd_towers_attack_dev(mtd_ctx *ctx, uint32_t gameTime, const uint64_t *random_words)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	return towers_attack_dev(ctx, gameTime, random_words);
}

extern "C" int mtd_towers_download(mtd_ctx *ctx, uint32_t *firingDelay, uint32_t *mana, int32_t *target, uint32_t *hits)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	const size_t N = (size_t)ctx->ntowers;
	if (N == 0)
		return MTD_OK;
	if (firingDelay)
		CU(cudaMemcpyAsync(firingDelay, ctx->t_delay.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (mana)
		CU(cudaMemcpyAsync(mana, ctx->t_mana.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (target)
		CU(cudaMemcpyAsync(target, ctx->t_target.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (hits)
		CU(cudaMemcpyAsync(hits, ctx->t_hits.p, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
	return sync_and_status(ctx);
}

// =====================================================================================================
// C3 across GPUs: the layers of one rank are broadcast INTO the resident layer buffers of the others
// =====================================================================================================
extern "C" int mtd_waypoints_layers_refresh(mtd_ctx *ctx)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (ctx->nlayers <= 0)
		return fail(ctx, MTD_ERR_STATE, "no resident waypoint layers (mtd_waypoints_update_dev once, to size them)");
	if (ctx->nlayers <= FIND_MAXN)
	{
		k_build_dtab<<<1, 64, 0, ctx->stream>>>(layers_view(ctx), ctx->dtab.p);
		LAUNCHED();
	}
	// the imported layers belong to the grid as it is now
	ctx->inc_dirty = 0;
	ctx->inc_layers_ok = ctx->inc_n == ctx->nlayers && !ctx->h_blocking.empty();
	return MTD_OK;
}

// =====================================================================================================
// Allocation-free variant (no counterpart in the reference)
// =====================================================================================================
// Full-path lengths / distances / aggregates of a few resident maps, with the kernel's scratch -- the distance and
// next-hop layers the walker warps chase, and the map counter -- in a fixed array that is part of the library image
// instead of cudaMalloc'ed buffers. Every GPU that loads the library holds MTD_NOALLOC_SCRATCH_BYTES for it until exit.
// All calls share the array: calls that overlap on one GPU (two streams, two contexts) race.
static __device__ __align__(256) unsigned char g_noalloc_scratch[MTD_NOALLOC_SCRATCH_BYTES];

namespace
{
	constexpr int NOALLOC_MAX_DEVICES = 64;
	constexpr size_t NOALLOC_HEAD = 256; // control words in front of the layers (word 0: the kernel's map counter)
	std::mutex g_noalloc_mu;
	unsigned char *g_noalloc_ptr[NOALLOC_MAX_DEVICES]; // device address of g_noalloc_scratch, per device (the symbol's name is only a host-side handle)
	bool g_noalloc_cfg[NOALLOC_MAX_DEVICES];           // the kernel was opted in to its shared memory size on that device

	// once per device, with that device current (the caller holds a DeviceGuard)
	int noalloc_setup(mtd_ctx *ctx, unsigned char **scratch)
	{
		if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES)
			return fail(ctx, MTD_ERR_CAPACITY, "device ordinal %d beyond the %d the allocation-free entry point keeps state for", ctx->device, NOALLOC_MAX_DEVICES);
		std::lock_guard<std::mutex> lk(g_noalloc_mu);
		if (!g_noalloc_ptr[ctx->device])
		{
			void *p = nullptr;
			const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
			if (e != cudaSuccess || !p)
			{
				(void)cudaGetLastError();
				return fail(ctx, MTD_ERR_CUDA, "cudaGetSymbolAddress(scratch) failed: %s", cudaGetErrorString(e));
			}
			g_noalloc_ptr[ctx->device] = (unsigned char *)p;
		}
		if (!g_noalloc_cfg[ctx->device])
		{
			const cudaError_t e = cudaFuncSetAttribute((const void *)k_waypoints_small<SMALL_GROUP_WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MAX_SMEM);
			if (e != cudaSuccess)
			{
				(void)cudaGetLastError();
				return fail(ctx, MTD_ERR_CUDA, "kernel configuration failed: %s", cudaGetErrorString(e));
			}
			g_noalloc_cfg[ctx->device] = true;
		}
		*scratch = g_noalloc_ptr[ctx->device];
		return MTD_OK;
	}
}

extern "C" int mtd_waypoints_paths_noalloc(mtd_ctx *ctx, int32_t W, int32_t H, int32_t nmaps, int32_t n, const int16_t *d_elev, const uint16_t *d_flags, const uint32_t *d_wp_tiles, uint32_t *d_fullpath_len, uint32_t *d_full_distance, uint32_t *d_agg)
{
	if (!ctx)
		return MTD_ERR_ARG;
	std::lock_guard<std::mutex> lk(ctx->mu);
	DeviceGuard g(ctx->device);
	if (W < 1 || H < 1 || nmaps < 0 || n < 1 || n > FIND_MAXN)
		return fail(ctx, MTD_ERR_ARG, "bad shape: W, H >= 1, nmaps >= 0, 1 <= n <= %d waypoints", FIND_MAXN);
	if (nmaps == 0)
		return MTD_OK;
	if (!d_elev || !d_flags || !d_wp_tiles || !d_fullpath_len || !d_full_distance || !d_agg)
		return fail(ctx, MTD_ERR_ARG, "every input and output pointer is required");
	const int R = small_resident(W, H, n);
	if (R < 1)
		return fail(ctx, MTD_ERR_CAPACITY, "a %d x %d map does not fit the one-CTA path", W, H);
	const size_t T = (size_t)W * H;
	const size_t layers = (size_t)nmaps * n * T; // words per array (distances, next hops)
	const size_t need = NOALLOC_HEAD + layers * 8;
	if (need > (size_t)MTD_NOALLOC_SCRATCH_BYTES)
		return fail(ctx, MTD_ERR_CAPACITY, "nmaps * n * W * H * 8 + %zu = %zu bytes of layers exceed the fixed scratch of %zu bytes (MTD_NOALLOC_SCRATCH_BYTES): at most %zu tiles x layers per call",
					NOALLOC_HEAD, need, (size_t)MTD_NOALLOC_SCRATCH_BYTES, ((size_t)MTD_NOALLOC_SCRATCH_BYTES - NOALLOC_HEAD) / 8);
	unsigned char *scratch = nullptr;
	const int rc = noalloc_setup(ctx, &scratch);
	if (rc)
		return rc;

	constexpr int GW = SMALL_GROUP_WARPS;
	WaypointsSmallParams P;
	memset(&P, 0, sizeof(P));
	P.W = W;
	P.H = H;
	P.n = n;
	P.R = R;
	P.nmaps = nmaps;
	P.nsrc = 1;
	P.elev = d_elev;
	P.flags = d_flags;
	P.wp = d_wp_tiles;
	P.extra_blocked = MTD_M;
	P.inc_tile = MTD_M;
	P.dist = (uint32_t *)(scratch + NOALLOC_HEAD);
	P.tiles = P.dist + layers;
	P.fullpath_len = d_fullpath_len;
	P.full_distance = d_full_distance;
	P.agg = d_agg;
	P.status = ctx->d_status.p; // allocated by mtd_create; read by mtd_synchronize
	P.map_counter = (uint32_t *)scratch;
	P.smem_budget = (uint32_t)MAX_SMEM;
	P.walker = 1; // the walker warps chase the layers in the scratch
	const SmallLayout L = small_layout(W, H, R, P.smem_budget, 0, GW);
	// the array keeps the previous call's contents: the map counter must start at zero; the layers are written before they are read
	CU(cudaMemsetAsync(scratch, 0, NOALLOC_HEAD, ctx->stream));
	const int per_sm = (size_t)L.total * 2 + 2048 <= MAX_SMEM ? 2 : 1;
	const int blocks = nmaps < ctx->sms * per_sm ? nmaps : ctx->sms * per_sm;
	k_waypoints_small<GW><<<blocks, SMALL_BLOCK, L.total, ctx->stream>>>(P);
	LAUNCHED();
	return MTD_OK;
}
