// waypoints_small.cuh -- fused Waypoints::update (sources/map/paths.cpp:24-139) for maps that fit one CTA's
// shared memory: the game's 101x101 map and batches of them.
//
// One persistent CTA (512 threads) per SM pulls maps from a global counter. Per map:
//   1. TMA bulk copies (cp.async.bulk, mbarrier completion) stage the map's elevations and flags in shared memory
//   2. edge weights = Grid::neighborDistance of every right/down edge (sources/map/grid.cpp:141-145), blocking bitmap
//   3. up to SMALL_MAX_RESIDENT distance layers are relaxed concurrently, FOUR WARPS PER LAYER:
//        bucketed wavefront, bucket width 255 = lower bound of every edge weight, so a settled label is final
//        (Dijkstra order without a heap); each warp appends the cells it reaches to its own queue segment with
//        ballot compaction (no atomics on the queue), the group meets at one named barrier per bucket
//   4. next-hop tiles (closed form of the heap's tie-break, see sssp_warp.cuh) + coalesced streaming write-out,
//      overlapped with the full-path walks (paths.cpp:87-108) that one lane per spawner runs on the resident layers
//   5. aggregates (paths.cpp:122-138)
// Queue overflow (pathological frontiers) switches the layer to a queue-less dense sweep that needs no capacity.
#pragma once

#include "common.cuh"
#include "sssp_warp.cuh"
#include "walk.cuh"

namespace mtd
{
	enum : int
	{
		ST_GUARD = 1,      // a distance reached the uint32 guard band
		ST_NO_VALID = 2,   // generated map has no valid tile (reference would spin forever in findSpawnPoint)
		ST_PATH_TRUNC = 4, // a full path exceeded fullpath_cap (length is still exact)
	};

	constexpr int SMALL_THREADS = 512;
	constexpr int SMALL_GROUP_WARPS = 2;   // default number of warps cooperating on one layer (template parameter GW: 1, 2 or 4)
	constexpr int SMALL_WALKER_WARP = SMALL_THREADS / 32; // the extra warp that walks the previous map's paths
	constexpr int SMALL_MAX_RESIDENT = 4;  // layers concurrently in shared memory
	constexpr int MTD_MAXSRC_SMALL = 32;
	constexpr int SMALL_OUT_CHUNK = 128;   // cells per write-out work item

	struct WaypointsSmallParams
	{
		int W, H;
		int n;       // layers per map
		int R;       // resident layers per pass (<= SMALL_MAX_RESIDENT, chosen by the host to fit shared memory)
		int nmaps;
		int nsrc;    // sources per layer: 1 = reference; > 1 = multi-source extension (then n must be 1)
		const int16_t *elev;    // [nmaps][T]
		const uint16_t *flags;  // [nmaps][T]
		const uint32_t *wp;     // [nmaps][n] (or [nsrc] when nsrc > 1)
		uint32_t extra_blocked; // MTD_M = none
		uint32_t *tiles;        // [nmaps][n][T] or null
		uint32_t *dist;         // [nmaps][n][T]; required when n > R and paths are wanted
		uint32_t *fullpath;     // [nmaps][n][cap] or null
		uint32_t cap;
		uint32_t *fullpath_len;  // [nmaps][n] or null: null skips walks + aggregates
		uint32_t *full_distance; // [nmaps][n]
		uint32_t *agg;           // [nmaps][3]
		int *status;
		uint32_t *map_counter;   // zeroed before the launch
		uint32_t smem_budget;    // dynamic shared memory of the launch
		int qseg_cap;            // > 0 caps the queue segment size (tests force the dense fallback with it)
		int walker;              // != 0: paths are walked by the walker warp on the layers in global memory (needs tiles + dist),
		                         //       one map behind the workers; 0: the workers walk the resident layers themselves
	};

	// ---- shared memory layout (byte offsets), identical on host and device ---------------------------------
	struct SmallLayout
	{
		uint32_t dist, wr, wd, blk, ctl, dtab, q, total;
		int qseg;
	};

	__host__ __device__ inline uint32_t align16u(uint32_t v)
	{
		return (v + 15u) & ~15u;
	}

	__host__ __device__ inline SmallLayout small_layout(int T, int R, uint32_t budget, int qseg_cap, int gw = SMALL_GROUP_WARPS)
	{
		SmallLayout L;
		uint32_t o = 0;
		const uint32_t stage = 2u * align16u((uint32_t)(T + 8) * 2u); // elev + flags staging aliases the distance layers
		uint32_t dist_bytes = (uint32_t)R * (uint32_t)T * 4u;
		if (dist_bytes < stage)
			dist_bytes = stage;
		L.dist = o;
		o += align16u(dist_bytes);
		L.wr = o;
		o += align16u((uint32_t)T * 2u);
		L.wd = o;
		o += align16u((uint32_t)T * 2u);
		L.blk = o;
		o += align16u((uint32_t)((T + 31) >> 5) * 4u);
		L.ctl = o;
		o += (uint32_t)R * 64u; // [parity 2][counts 4 | mins 4] uint32
		L.dtab = o;
		o += FIND_MAXN * FIND_MAXN * 4u;
		L.q = o;
		const uint32_t per_entry = (uint32_t)R * 2u * (uint32_t)gw * 2u; // layers x parities x segments x sizeof(u16)
		int qseg = budget > o ? (int)((budget - o) / per_entry) : 0;
		qseg &= ~7;
		if (qseg > 2048)
			qseg = 2048;
		if (qseg_cap > 0 && qseg > qseg_cap)
			qseg = qseg_cap < MTD_MAXSRC_SMALL ? MTD_MAXSRC_SMALL : qseg_cap; // the seeds must always fit (dense rounds need thr_prev > 0)
		L.qseg = qseg;
		L.total = o + per_entry * (uint32_t)qseg;
		return L;
	}

	// ---- small PTX helpers -----------------------------------------------------------------------------------
	__device__ __forceinline__ uint32_t smem_u32(const void *p)
	{
		return (uint32_t)__cvta_generic_to_shared(p);
	}

	__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
	{
		asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
	}

	__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
	{
		asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
	}

	__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t phase)
	{
		asm volatile(
			"{\n"
			".reg .pred p;\n"
			"WAIT_%=:\n"
			"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
			"@p bra DONE_%=;\n"
			"bra WAIT_%=;\n"
			"DONE_%=:\n"
			"}\n" ::"r"(smem_u32(bar)),
			"r"(phase)
			: "memory");
	}

	// TMA bulk copy global -> shared (1-D), completion counted in bytes on the mbarrier. 16-byte aligned, size % 16 == 0.
	__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
	{
		asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
	}

	__device__ __forceinline__ void fence_proxy_async()
	{
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	}

	__device__ __forceinline__ void group_barrier(int id, int threads)
	{
		asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
	}

	// i / W for i < 65536, W <= 256 with magicW = ceil(2^32 / W); W == 1 has no 32-bit magic and passes 0
	__device__ __forceinline__ uint32_t div_w(uint32_t i, uint32_t magicW)
	{
		return magicW ? __umulhi(i, magicW) : i;
	}

	// stage T int16/uint16 elements of `src` (2-byte aligned) at `dst_base + (src & 15)` so that the 16-byte aligned
	// body goes through the TMA; returns the body size. Called by ONE thread; head/tail elements are copied by
	// stage_edges() on other threads.
	struct StagePlan
	{
		uint32_t head, body, tail; // bytes
	};

	__device__ __forceinline__ StagePlan stage_plan(const void *src, uint32_t bytes)
	{
		StagePlan s;
		const uint32_t mis = (uint32_t)((uintptr_t)src & 15u);
		s.head = (16u - mis) & 15u;
		if (s.head > bytes)
			s.head = bytes;
		s.body = (bytes - s.head) & ~15u;
		s.tail = bytes - s.head - s.body;
		return s;
	}

	// ---- per-cell relaxation shared by the queued and the dense round -----------------------------------------
	// Settles cell c (distance d, final because d < dmin + 255) and relaxes its four neighbours with shared-memory
	// atomicMin. fresh[k] != 0 iff neighbour k was reached for the first time (old label was "unreached"), which
	// is the only way a cell enters the queue: labels of settled cells never improve again. Returns min of the
	// improved labels (MTD_INF if none).
	struct RelaxOut
	{
		uint32_t cell[4]; // packed y<<8|x of the neighbour
		bool fresh[4];
		uint32_t newmin;
		bool guard;
	};

	__device__ __forceinline__ RelaxOut relax_cell(uint32_t *dist, const uint16_t *wr, const uint16_t *wd, int W, int H, bool settle, int x, int y, uint32_t c, uint32_t d)
	{
		RelaxOut o;
		o.newmin = MTD_INF;
		o.guard = false;
		// 0: up, 1: left, 2: right, 3: down
		const bool ok[4] = { settle && y > 0, settle && x > 0, settle && x < W - 1, settle && y < H - 1 };
		const uint32_t nb[4] = { c - (uint32_t)W, c - 1u, c + 1u, c + (uint32_t)W };
		uint32_t wv[4];
		wv[0] = ok[0] ? wd[nb[0]] : 0u;
		wv[1] = ok[1] ? wr[nb[1]] : 0u;
		wv[2] = ok[2] ? wr[c] : 0u;
		wv[3] = ok[3] ? wd[c] : 0u;
		uint32_t old[4], nd[4];
#pragma unroll
		for (int k = 0; k < 4; k++)
		{
			nd[k] = d + MIN_EDGE + wv[k];
			old[k] = 0u;
			if (ok[k])
			{
				if (nd[k] >= MTD_GUARD)
					o.guard = true;
				else
					old[k] = atomicMin(&dist[nb[k]], nd[k]);
			}
		}
		const uint32_t e = ((uint32_t)y << 8) | (uint32_t)x;
		o.cell[0] = e - 256u;
		o.cell[1] = e - 1u;
		o.cell[2] = e + 1u;
		o.cell[3] = e + 256u;
#pragma unroll
		for (int k = 0; k < 4; k++)
		{
			const bool improved = ok[k] && nd[k] < old[k];
			o.fresh[k] = improved && old[k] == MTD_INF;
			if (improved)
				o.newmin = min(o.newmin, nd[k]);
		}
		return o;
	}

	// ---- one layer, GW warps ---------------------------------------------------------------------------------------
	// ctl: [parity][0..3] = entries each warp appended for the round that reads `parity`, [parity][4..7] = min pending
	// label each warp saw. q: [parity][warp][qseg] packed cells. Initial state (parity 0) is written by the caller.
	// Returns true if a label hit the guard band.
	template<int GW>
	__device__ bool group_relax(uint32_t *dist, const uint16_t *wr, const uint16_t *wd, uint16_t *q, uint32_t *ctl, int W, int H, int T, int qseg, uint32_t magicW, int gw, uint32_t lane, int bar_id)
	{
		const uint32_t lt = lanemask_lt();
		uint32_t parity = 0, thr_prev = 0;
		bool guard = false, dense = false;
		for (;;)
		{
			const uint4 cn = *reinterpret_cast<const uint4 *>(ctl + parity * 8);
			const uint4 mn = *reinterpret_cast<const uint4 *>(ctl + parity * 8 + 4);
			const uint32_t dmin = min(min(mn.x, mn.y), min(mn.z, mn.w));
			if (dmin == MTD_INF)
				break; // nothing pending
			const uint32_t thr = dmin >= MTD_GUARD ? MTD_INF : dmin + MIN_EDGE;
			const uint32_t cap = (uint32_t)qseg;
			dense = dense || cn.x > cap || cn.y > cap || cn.z > cap || cn.w > cap; // an append was dropped: the queue is incomplete from now on
			uint32_t newmin = MTD_INF, outn = 0;
			uint16_t *wq = q + ((parity ^ 1u) * GW + (uint32_t)gw) * cap;
			if (!dense)
			{
				const uint32_t p1 = cn.x, p2 = p1 + cn.y, p3 = p2 + cn.z, n = p3 + cn.w;
				const uint16_t *rq = q + parity * GW * cap;
				for (uint32_t base = (uint32_t)gw * 32u; base < n; base += GW * 32u)
				{
					const uint32_t idx = base + lane;
					const bool active = idx < n;
					const uint32_t seg = (idx >= p1) + (idx >= p2) + (idx >= p3);
					const uint32_t off = idx - (seg == 0 ? 0u : seg == 1 ? p1 : seg == 2 ? p2 : p3);
					const uint32_t e = active ? rq[seg * cap + off] : 0u;
					const int y = (int)(e >> 8), x = (int)(e & 255u);
					const uint32_t c = (uint32_t)(y * W + x);
					const uint32_t d = active ? dist[c] : MTD_INF;
					const bool settle = active && d < thr;
					const bool keep = active && !settle;
					const RelaxOut r = relax_cell(dist, wr, wd, W, H, settle, x, y, c, d);
					guard |= r.guard;
					newmin = min(newmin, keep ? d : r.newmin);
					// ---- append survivors and newly reached cells to this warp's segment (ballot compaction) ----
					const uint32_t bk = __ballot_sync(FULL_MASK, keep);
					const uint32_t b0 = __ballot_sync(FULL_MASK, r.fresh[0]);
					const uint32_t b1 = __ballot_sync(FULL_MASK, r.fresh[1]);
					const uint32_t b2 = __ballot_sync(FULL_MASK, r.fresh[2]);
					const uint32_t b3 = __ballot_sync(FULL_MASK, r.fresh[3]);
					uint32_t pos = outn;
					if (keep && pos + __popc(bk & lt) < cap)
						wq[pos + __popc(bk & lt)] = (uint16_t)e;
					pos += __popc(bk);
					if (r.fresh[0] && pos + __popc(b0 & lt) < cap)
						wq[pos + __popc(b0 & lt)] = (uint16_t)r.cell[0];
					pos += __popc(b0);
					if (r.fresh[1] && pos + __popc(b1 & lt) < cap)
						wq[pos + __popc(b1 & lt)] = (uint16_t)r.cell[1];
					pos += __popc(b1);
					if (r.fresh[2] && pos + __popc(b2 & lt) < cap)
						wq[pos + __popc(b2 & lt)] = (uint16_t)r.cell[2];
					pos += __popc(b2);
					if (r.fresh[3] && pos + __popc(b3 & lt) < cap)
						wq[pos + __popc(b3 & lt)] = (uint16_t)r.cell[3];
					pos += __popc(b3);
					outn = pos; // may exceed cap: that is the overflow signal read at the next round
				}
			}
			else
			{
				// dense round: every pending cell has thr_prev <= label < INF (settled ones are below thr_prev, parked
				// blocked cells are 0), so the bucket is found by scanning the labels -- no queue, no capacity
				outn = cap + 1u; // stay dense
				for (int i = gw * 32 + (int)lane; i < T; i += GW * 32)
				{
					const uint32_t d = dist[i];
					const bool settle = d >= thr_prev && d < thr;
					if (d >= thr && d != MTD_INF)
						newmin = min(newmin, d);
					if (__any_sync(__activemask(), settle))
					{
						const int y = (int)div_w((uint32_t)i, magicW), x = i - y * W;
						const RelaxOut r = relax_cell(dist, wr, wd, W, H, settle, x, y, (uint32_t)i, d);
						guard |= r.guard;
						newmin = min(newmin, r.newmin);
					}
				}
			}
			newmin = warp_min_u32(newmin);
			if (lane == 0)
			{
				ctl[(parity ^ 1u) * 8 + (uint32_t)gw] = outn;
				ctl[(parity ^ 1u) * 8 + 4 + (uint32_t)gw] = newmin;
			}
			if (GW == 1)
				__syncwarp();
			else
				group_barrier(bar_id, GW * 32);
			thr_prev = thr;
			parity ^= 1u;
		}
		return __any_sync(FULL_MASK, guard);
	}

	// next-hop with the four edge weights already loaded (wu = w(c-W,c), wl = w(c-1,c), wrr = w(c,c+1), wdn = w(c,c+W), all - 255)
	__device__ __forceinline__ uint32_t next_hop_w(const uint32_t *dl, int W, int H, int x, int y, uint32_t c, uint32_t wu, uint32_t wl, uint32_t wrr, uint32_t wdn)
	{
		const uint32_t dv = dl[c];
		if (dv == MTD_INF || dv == 0u)
			return MTD_INF;
		const uint32_t du0 = y > 0 ? dl[c - W] : MTD_INF;
		const uint32_t du1 = x > 0 ? dl[c - 1] : MTD_INF;
		const uint32_t du2 = x < W - 1 ? dl[c + 1] : MTD_INF;
		const uint32_t du3 = y < H - 1 ? dl[c + W] : MTD_INF;
		uint32_t best = MTD_INF, bestd = MTD_INF;
		if (du0 != MTD_INF && du0 + MIN_EDGE + wu == dv)
		{
			bestd = du0;
			best = c - (uint32_t)W;
		}
		if (du1 != MTD_INF && du1 + MIN_EDGE + wl == dv && du1 < bestd)
		{
			bestd = du1;
			best = c - 1u;
		}
		if (du2 != MTD_INF && du2 + MIN_EDGE + wrr == dv && du2 < bestd)
		{
			bestd = du2;
			best = c + 1u;
		}
		if (du3 != MTD_INF && du3 + MIN_EDGE + wdn == dv && du3 < bestd)
		{
			bestd = du3;
			best = c + (uint32_t)W;
		}
		return best;
	}

	// layers held in shared memory: next-hop tiles are derived from the final distances on the fly
	struct SmemLayersAcc
	{
		const uint32_t *sdist; // [R][T]
		const uint16_t *wr, *wd;
		int W, H, T;
		uint32_t magicW;
		__device__ __forceinline__ uint32_t dist(uint32_t l, uint32_t t) const { return sdist[(size_t)l * T + t]; }
		__device__ __forceinline__ uint32_t hop(uint32_t l, uint32_t t) const
		{
			const int y = (int)div_w(t, magicW), x = (int)t - y * W;
			return next_hop(sdist + (size_t)l * T, wr, wd, W, H, x, y);
		}
		__device__ __forceinline__ uint32_t edge(uint32_t a, uint32_t b) const
		{
			const uint32_t lo = a < b ? a : b, hi = a < b ? b : a;
			return MIN_EDGE + (hi - lo == 1u ? wr[lo] : wd[lo]);
		}
	};

	// barrier of the 512 worker threads (the walker warp never takes part)
	__device__ __forceinline__ void workers_sync()
	{
		group_barrier(1, SMALL_THREADS);
	}

	// Full-path walks + aggregates of one map on layers resident in global memory, by ONE warp: lane s walks spawner s
	// (the dependent loads of the n chases overlap), then all lanes pad the path buffers. wp/D/fd: scratch in shared memory.
	__device__ void walk_map_global(const WaypointsSmallParams &P, int mapi, uint32_t *wp, uint32_t *D, uint32_t *fd, uint32_t lane)
	{
		const int n = P.n;
		const size_t T = (size_t)P.W * P.H;
		GlobalLayersAcc acc;
		acc.gdist = P.dist + (size_t)mapi * n * T;
		acc.gtiles = P.tiles + (size_t)mapi * n * T;
		acc.gelev = P.elev + (size_t)mapi * T;
		acc.T = T;
		if ((int)lane < n)
			wp[lane] = P.wp[(size_t)mapi * n + lane];
		__syncwarp();
		for (int i = (int)lane; i < n * n; i += 32)
		{
			const int a = i / n, b = i % n;
			D[a * FIND_MAXN + b] = acc.dist((uint32_t)b, wp[a]);
		}
		__syncwarp();
		const bool table_fast = walk_table_fast(n, D);
		WalkResult r;
		r.len = 0;
		r.distance = 0;
		if ((int)lane < n)
		{
			uint32_t *path = P.fullpath ? P.fullpath + ((size_t)mapi * n + lane) * P.cap : nullptr;
			r = walk_path(n, wp, D, table_fast, (int)lane, acc, path, P.cap, (uint32_t)T * (uint32_t)n + 8u);
			if (P.fullpath && r.len > P.cap)
				atomicOr(P.status, ST_PATH_TRUNC);
			P.fullpath_len[(size_t)mapi * n + lane] = r.len;
			P.full_distance[(size_t)mapi * n + lane] = r.distance;
			fd[lane] = r.distance;
		}
		__syncwarp();
		if (P.fullpath)
			for (int s = 0; s < n; s++)
			{
				uint32_t *path = P.fullpath + ((size_t)mapi * n + s) * P.cap;
				for (uint32_t i = __shfl_sync(FULL_MASK, r.len, s) + lane; i < P.cap; i += 32u)
					st_stream_u32(path + i, MTD_INF);
			}
		if (lane == 0)
			walk_aggregates(n, fd, P.agg + (size_t)mapi * 3);
		__syncwarp();
	}

	// =============================================================================================================
	// 512 worker threads + 1 walker warp. GW = warps cooperating on one layer during the relaxation.
	template<int GW>
	__global__ void __launch_bounds__(SMALL_THREADS + 32, 1) k_waypoints_small(const WaypointsSmallParams P)
	{
		extern __shared__ __align__(16) unsigned char smem_raw[];
		const int W = P.W, H = P.H, T = W * H, n = P.n, R = P.R;
		const SmallLayout L = small_layout(T, R, P.smem_budget, P.qseg_cap, GW);
		uint32_t *sdist = (uint32_t *)(smem_raw + L.dist);
		uint16_t *wr = (uint16_t *)(smem_raw + L.wr);
		uint16_t *wd = (uint16_t *)(smem_raw + L.wd);
		uint32_t *blk = (uint32_t *)(smem_raw + L.blk);
		uint32_t *ctl = (uint32_t *)(smem_raw + L.ctl);
		uint32_t *Dtab = (uint32_t *)(smem_raw + L.dtab);
		uint16_t *q = (uint16_t *)(smem_raw + L.q);
		const int qseg = L.qseg;
		__shared__ __align__(8) uint64_t s_mbar;
		__shared__ uint32_t s_fd[FIND_MAXN];
		__shared__ uint32_t s_wp[MTD_MAXSRC_SMALL];
		__shared__ int s_guard;
		__shared__ uint32_t s_map, s_chunk;
		// the walker warp's private copies (it works on the PREVIOUS map while the workers overwrite the above)
		__shared__ uint32_t w_wp[FIND_MAXN], w_D[FIND_MAXN * FIND_MAXN], w_fd[FIND_MAXN];

		const int tid = threadIdx.x;
		const int warp = tid >> 5;
		const uint32_t lane = lane_id();
		const bool resident_all = n <= R;
		const int nsrc_layer = P.nsrc > 1 ? P.nsrc : 1;
		const int nwp = P.nsrc > 1 ? P.nsrc : n;
		const uint32_t magicW = W > 1 ? (uint32_t)((0x100000000ull + (uint32_t)W - 1u) / (uint32_t)W) : 0u; // y = umulhi(i, magicW) for i < 65536, W <= 256
		const uint32_t stage_half = align16u((uint32_t)(T + 8) * 2u);
		const bool want_walks = P.fullpath_len && P.nsrc <= 1;
		uint32_t mbar_phase = 0;
		int walker_prev = -1;

		if (tid == 0)
		{
			mbar_init(&s_mbar, 1);
			asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
		}

		for (;;)
		{
			__syncthreads(); // all 544 threads: the previous map is completely written out and the walker finished the one before
			if (tid == 0)
				s_map = atomicAdd(P.map_counter, 1u);
			__syncthreads();
			const int mapi = (int)s_map;
			if (warp == SMALL_WALKER_WARP)
			{
				// ---- walker warp: paths of the map the workers finished last, from its layers in global memory (still in L2) ----
				if (P.walker && want_walks && walker_prev >= 0)
					walk_map_global(P, walker_prev, w_wp, w_D, w_fd, lane);
				if (mapi >= P.nmaps)
					break;
				walker_prev = mapi;
				continue;
			}
			if (mapi >= P.nmaps)
				break;
			const int16_t *ge = P.elev + (size_t)mapi * T;
			const uint16_t *gf = P.flags + (size_t)mapi * T;
			const uint32_t *gwp = P.wp + (size_t)mapi * nwp;
			uint32_t *gtiles = P.tiles ? P.tiles + (size_t)mapi * n * T : nullptr;
			uint32_t *gdist = P.dist ? P.dist + (size_t)mapi * n * T : nullptr;

			// ---- 1. stage elevations + flags (TMA bulk copies for the 16-byte aligned bodies) ----
			unsigned char *st_e = smem_raw + L.dist + ((uintptr_t)ge & 15u);
			unsigned char *st_f = smem_raw + L.dist + stage_half + ((uintptr_t)gf & 15u);
			const int16_t *selev = (const int16_t *)st_e;
			const uint16_t *sflags = (const uint16_t *)st_f;
			{
				const StagePlan pe = stage_plan(ge, (uint32_t)T * 2u), pf = stage_plan(gf, (uint32_t)T * 2u);
				if (tid == 0)
				{
					fence_proxy_async(); // earlier generic-proxy accesses to this shared memory are ordered before the async writes
					mbar_expect_tx(&s_mbar, pe.body + pf.body);
					if (pe.body)
						tma_bulk_g2s(st_e + pe.head, (const unsigned char *)ge + pe.head, pe.body, &s_mbar);
					if (pf.body)
						tma_bulk_g2s(st_f + pf.head, (const unsigned char *)gf + pf.head, pf.body, &s_mbar);
				}
				else if (warp == 1)
				{
					// unaligned head / tail elements (at most 7 each)
					for (uint32_t b = lane * 2u; b < pe.head; b += 64u)
						*(int16_t *)(st_e + b) = *(const int16_t *)((const unsigned char *)ge + b);
					for (uint32_t b = lane * 2u; b < pe.tail; b += 64u)
						*(int16_t *)(st_e + pe.head + pe.body + b) = *(const int16_t *)((const unsigned char *)ge + pe.head + pe.body + b);
					for (uint32_t b = lane * 2u; b < pf.head; b += 64u)
						*(uint16_t *)(st_f + b) = *(const uint16_t *)((const unsigned char *)gf + b);
					for (uint32_t b = lane * 2u; b < pf.tail; b += 64u)
						*(uint16_t *)(st_f + pf.head + pf.body + b) = *(const uint16_t *)((const unsigned char *)gf + pf.head + pf.body + b);
				}
				else if (warp == 2)
				{
					for (int i = (int)lane; i < nwp && i < MTD_MAXSRC_SMALL; i += 32)
						s_wp[i] = gwp[i];
					if (lane == 0)
					{
						s_guard = 0;
						s_chunk = 0;
					}
				}
				mbar_wait(&s_mbar, mbar_phase);
				mbar_phase ^= 1u;
				workers_sync();
			}

			// ---- 2. edge weights (Grid::neighborDistance of the right and down edge, grid.cpp:141-145) + blocking bitmap ----
			for (int base = 0; base < T; base += SMALL_THREADS)
			{
				const int i = base + tid;
				bool blocked = false;
				if (i < T)
				{
					const int y = (int)div_w((uint32_t)i, magicW), x = i - y * W;
					const int16_t e = selev[i];
					wr[i] = x < W - 1 ? (uint16_t)(edge_weight(e, selev[i + 1]) - MIN_EDGE) : (uint16_t)0;
					wd[i] = y < H - 1 ? (uint16_t)(edge_weight(e, selev[i + W]) - MIN_EDGE) : (uint16_t)0;
					blocked = (sflags[i] & F_BLOCKING) != 0 || (uint32_t)i == P.extra_blocked;
				}
				const uint32_t bal = __ballot_sync(FULL_MASK, blocked);
				if (lane == 0 && base + warp * 32 < T)
					blk[(base >> 5) + warp] = bal;
			}
			workers_sync(); // staging is dead from here on: the distance layers take its place

			for (int l0 = 0; l0 < n; l0 += R)
			{
				const int lb = (n - l0) < R ? (n - l0) : R; // layers in this pass
				// ---- 3a. init: paths.cpp:40-43; blocked tiles are parked at 0 during the run (never relaxed: >= 255 cannot lower 0) ----
				for (int i = tid; i < T; i += SMALL_THREADS)
				{
					const uint32_t v = ((blk[i >> 5] >> (i & 31)) & 1u) ? 0u : MTD_INF;
					for (int r = 0; r < lb; r++)
						sdist[(size_t)r * T + i] = v;
				}
				workers_sync();
				// seeds: paths.cpp:48-49 (a blocked source still expands: the popped tile is never tested, paths.cpp:63-67)
				if (warp < lb * GW && (warp % GW) == 0 && lane == 0)
				{
					const int g = warp / GW;
					uint32_t *dl = sdist + (size_t)g * T;
					uint32_t *c = ctl + g * 16;
					uint16_t *q0 = q + (size_t)g * 2 * GW * qseg;
					uint32_t nseed = 0;
					for (int s = 0; s < nsrc_layer; s++)
					{
						const uint32_t src = P.nsrc > 1 ? s_wp[s] : s_wp[l0 + g];
						bool dup = false;
						for (int s2 = 0; s2 < s; s2++)
							dup |= P.nsrc > 1 && s_wp[s2] == src;
						if (dup)
							continue;
						dl[src] = 0u;
						const uint32_t y = div_w(src, magicW);
						if ((int)nseed < qseg)
							q0[nseed] = (uint16_t)((y << 8) | (src - y * (uint32_t)W));
						nseed++;
					}
					c[0] = nseed;
					c[1] = c[2] = c[3] = 0u;
					c[4] = 0u;
					c[5] = c[6] = c[7] = MTD_INF;
					// slots of warps that do not exist (GW < 4) stay empty in both parities
					c[8] = c[9] = c[10] = c[11] = 0u;
					c[12] = c[13] = c[14] = c[15] = MTD_INF;
				}
				workers_sync();

				// ---- 3b. relaxation: GW warps per layer, one named barrier per bucket ----
				if (warp < lb * GW)
				{
					const int g = warp / GW;
					if (group_relax<GW>(sdist + (size_t)g * T, wr, wd, q + (size_t)g * 2 * GW * qseg, ctl + g * 16, W, H, T, qseg, magicW, warp % GW, lane, 2 + g) && lane == 0)
						s_guard = 1;
				}
				workers_sync();

				// ---- 3c. blocked, non-source tiles read "unreachable" (they were never relaxed: paths.cpp:65-67) ----
				for (int i = tid; i < T; i += SMALL_THREADS)
				{
					if (!((blk[i >> 5] >> (i & 31)) & 1u))
						continue;
					for (int r = 0; r < lb; r++)
					{
						bool is_src = false;
						if (P.nsrc > 1)
							for (int s = 0; s < nsrc_layer; s++)
								is_src |= s_wp[s] == (uint32_t)i;
						else
							is_src = s_wp[l0 + r] == (uint32_t)i;
						if (!is_src)
							sdist[(size_t)r * T + i] = MTD_INF;
					}
				}
				const bool walk_now = want_walks && !P.walker && resident_all; // single map: walk the resident layers while they are written out
				if (walk_now && tid < n * n)
				{
					// inter-waypoint table D[a][b] = dist_b[wp_a]; a waypoint tile is either open or the source of its own layer,
					// except blocked waypoints of OTHER layers, which the loop above may be rewriting concurrently: read the
					// final value explicitly
					const int a = tid / n, b = tid % n;
					const uint32_t t = s_wp[a];
					uint32_t v = sdist[(size_t)b * T + t];
					if (((blk[t >> 5] >> (t & 31)) & 1u) && s_wp[b] != t)
						v = MTD_INF;
					Dtab[a * FIND_MAXN + b] = v;
				}
				workers_sync();

				// ---- 4. full-path walks (one lane per spawner) overlapped with next-hop tiles + write-out of this pass ----
				uint32_t walk_len = 0;
				if (walk_now && warp < n && lane == 0)
				{
					SmemLayersAcc acc;
					acc.sdist = sdist;
					acc.wr = wr;
					acc.wd = wd;
					acc.W = W;
					acc.H = H;
					acc.T = T;
					acc.magicW = magicW;
					uint32_t *path = P.fullpath ? P.fullpath + ((size_t)mapi * n + warp) * P.cap : nullptr;
					const WalkResult r = walk_path(n, s_wp, Dtab, walk_table_fast(n, Dtab), warp, acc, path, P.cap, (uint32_t)T * (uint32_t)n + 8u);
					if (P.fullpath && r.len > P.cap)
						atomicOr(P.status, ST_PATH_TRUNC);
					P.fullpath_len[(size_t)mapi * n + warp] = r.len;
					P.full_distance[(size_t)mapi * n + warp] = r.distance;
					s_fd[warp] = r.distance;
					walk_len = r.len;
				}
				__syncwarp();
				if (walk_now && warp < n && P.fullpath)
				{
					// pad the unused tail of the path buffer with m, all lanes
					uint32_t *path = P.fullpath + ((size_t)mapi * n + warp) * P.cap;
					for (uint32_t i = __shfl_sync(FULL_MASK, walk_len, 0) + lane; i < P.cap; i += 32u)
						st_stream_u32(path + i, MTD_INF);
				}
				if (gdist || gtiles)
				{
					const int nchunks = (T + SMALL_OUT_CHUNK - 1) / SMALL_OUT_CHUNK;
					for (;;)
					{
						uint32_t item = 0;
						if (lane == 0)
							item = atomicAdd(&s_chunk, 1u);
						item = __shfl_sync(FULL_MASK, item, 0);
						if (item >= (uint32_t)nchunks)
							break;
#pragma unroll
						for (int k = 0; k < SMALL_OUT_CHUNK / 32; k++)
						{
							const int i = (int)item * SMALL_OUT_CHUNK + k * 32 + (int)lane;
							if (i >= T)
								break;
							const int y = (int)div_w((uint32_t)i, magicW), x = i - y * W;
							uint32_t wu = 0, wl = 0, wrr = 0, wdn = 0;
							if (gtiles)
							{
								wu = y > 0 ? wd[i - W] : 0u;
								wl = x > 0 ? wr[i - 1] : 0u;
								wrr = wr[i];
								wdn = wd[i];
							}
							for (int r = 0; r < lb; r++)
							{
								const uint32_t *dl = sdist + (size_t)r * T;
								if (gdist)
									st_stream_u32(gdist + (size_t)(l0 + r) * T + i, dl[i]);
								if (gtiles)
									st_stream_u32(gtiles + (size_t)(l0 + r) * T + i, next_hop_w(dl, W, H, x, y, (uint32_t)i, wu, wl, wrr, wdn));
							}
						}
					}
				}
				workers_sync();
				if (tid == 0)
					s_chunk = 0;
			}

			if (s_guard && tid == 0)
				atomicOr(P.status, ST_GUARD);
			if (!want_walks || P.walker)
				continue;

			if (!resident_all)
			{
				// single map with more layers than fit: walk the layers just written to global memory (L2), warp 0
				workers_sync(); // every warp's stores of the last pass are done
				if (warp == 0)
					walk_map_global(P, mapi, w_wp, w_D, w_fd, lane);
				continue;
			}

			// ---- 5. aggregates: paths.cpp:122-138 ----
			if (tid == 0)
				walk_aggregates(n, s_fd, P.agg + (size_t)mapi * 3);
		}
	}
}
