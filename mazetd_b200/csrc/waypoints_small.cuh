// waypoints_small.cuh -- fused Waypoints::update (sources/map/paths.cpp:24-139) for maps that fit one CTA's
// shared memory: the game's 101x101 map and batches of them.
//
// One persistent CTA (512 worker threads + 1 walker warp) per SM pulls maps from a global counter. Per map:
//   1. TMA bulk copies (cp.async.bulk, mbarrier completion) stage the map's elevations and flags in shared memory
//   2. edge weights = Grid::neighborDistance of every right/down edge (sources/map/grid.cpp:141-145), blocking bitmap
//   3. up to SMALL_MAX_RESIDENT distance layers are relaxed concurrently, GW warps per layer:
//        bucketed wavefront, bucket width 255 = lower bound of every edge weight, so a settled label is final
//        (Dijkstra order without a heap); each warp appends the cells it reaches to its own queue segment with
//        ballot compaction (no atomics on the queue), the group meets at one named barrier per bucket.
//        The layers live in a PADDED frame (one parked column on the right, one parked row above and below): the four
//        neighbours of a cell are p-1, p+1, p-Wp, p+Wp with no bounds test, and the queue holds linear indices.
//   4. next-hop tiles (closed form of the heap's tie-break, see sssp_warp.cuh) + coalesced streaming write-out
//   5. full paths + aggregates (paths.cpp:87-139): single map -> the workers walk the resident layers while they are
//      written out; batches -> the walker warp chases the PREVIOUS map's layers in global memory (L2) while the
//      workers relax the next map
// Queue overflow (pathological frontiers) switches the layer to a queue-less dense sweep that needs no capacity.
#pragma once

#include "common.cuh"
#include "mapgen_steps.cuh"
#include "sssp_warp.cuh"
#include "walk.cuh"

namespace mtd
{
	enum : int
	{
		ST_GUARD = 1,      // a distance reached the uint32 guard band
		ST_NO_VALID = 2,   // generated map has no valid tile (reference would spin forever in findSpawnPoint)
		ST_PATH_TRUNC = 4, // a full path exceeded fullpath_cap (length is still exact)
		ST_PEER_TIMEOUT = 8, // peer-coupled stripes: the other ranks did not reach a round barrier
		ST_ASYNC_TIMEOUT = 16, // the dataflow relaxation hit its time limit
		ST_MONSTER_STUCK = 32, // a monster stands on a tile outside the grid or cannot reach its next waypoint: its state was left untouched
	};

	constexpr int SMALL_THREADS = 512;
	constexpr int SMALL_GROUP_WARPS = 2;   // default number of warps cooperating on one layer (template parameter GW: 1, 2 or 4)
	constexpr int SMALL_WALKER_WARP = SMALL_THREADS / 32; // first of the extra warps that walk the finished maps' paths
	constexpr int SMALL_WALKERS = 2;                      // walker warps (finished maps alternate between them)
	constexpr int SMALL_BLOCK = SMALL_THREADS + SMALL_WALKERS * 32;
	constexpr uint32_t SMALL_RING = 8;                    // finished maps the workers may be ahead of the walkers
	constexpr int SMALL_MAX_RESIDENT = 4;  // layers concurrently in shared memory
	constexpr int MTD_MAXSRC_SMALL = 32;
	constexpr int SMALL_OUT_CHUNK = 128;   // cells per write-out work item
	// a settled label this close to the guard band is reported instead of relaxed: labels + one edge (< 2^17) stay below MTD_GUARD
	constexpr uint32_t SMALL_SETTLE_LIMIT = MTD_GUARD - (1u << 17);

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
		// incremental re-pathing: the layers of the previous update are still valid except for ONE tile
		// whose blocking state flipped since (sources/player/buildingsPlacement.cpp:120-130, 143-164, trial :18-42)
		const uint32_t *prev_dist; // [n][T] distances before the flip (may alias `dist`), or null
		uint32_t inc_tile;         // the flipped tile, MTD_M = none (full update). Needs nmaps == 1, n <= R, nsrc == 1
		// fused generator (newProcedural + newGrid + spawn picking, mapgen_steps.cuh): the kernel WRITES elev / flags / wp of
		// every map before it relaxes it. Needs nsrc == 1, n <= R, small_gen_fits()
		int gen;
		int gen_warps; // warps (of those without a layer) that compute the next map's elevations during the relaxation
		int gen_ox, gen_oy, gen_radius, gen_first;
		uint64_t gen_s1, gen_s2;
	};

	// ---- geometry of the padded frame -------------------------------------------------------------------------
	// grid cell (x, y) lives at p = (y + 1) * Wp + x with Wp = W + 1; column W and rows 0 and H + 1 are parked at 0
	struct SmallGeom
	{
		int W, H, T, Wp, Tp;
		uint32_t magicW, magicWp; // ceil(2^32 / d): q = umulhi(i, magic) is exact for i < 65536, d <= 257
		__host__ __device__ static uint32_t magic(uint32_t d) { return d > 1 ? (uint32_t)((0x100000000ull + d - 1u) / d) : 0u; }
		__host__ __device__ SmallGeom(int w, int h) : W(w), H(h), T(w * h), Wp(w + 1), Tp((w + 1) * (h + 2)), magicW(magic((uint32_t)w)), magicWp(magic((uint32_t)w + 1u)) {}
#ifdef __CUDACC__
		__device__ __forceinline__ uint32_t row_of(uint32_t i) const { return magicW ? __umulhi(i, magicW) : i; }
		__device__ __forceinline__ uint32_t g2p(uint32_t i) const
		{
			const uint32_t y = row_of(i);
			return i + y + (uint32_t)Wp; // (y + 1) * Wp + (i - y * W)
		}
		__device__ __forceinline__ uint32_t p2g(uint32_t p) const
		{
			const uint32_t yp = __umulhi(p, magicWp); // Wp >= 2
			return p - yp - (uint32_t)W;              // (yp - 1) * W + (p - yp * Wp)
		}
#endif
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

	__host__ __device__ inline SmallLayout small_layout(int W, int H, int R, uint32_t budget, int qseg_cap, int gw = SMALL_GROUP_WARPS)
	{
		SmallLayout L;
		const uint32_t T = (uint32_t)(W * H), Tp = (uint32_t)((W + 1) * (H + 2));
		uint32_t o = 0;
		const uint32_t stage = 2u * align16u((T + 8u) * 2u); // elev + flags staging aliases the distance layers
		uint32_t dist_bytes = (uint32_t)R * Tp * 4u;
		if (dist_bytes < stage)
			dist_bytes = stage;
		L.dist = o;
		o += align16u(dist_bytes);
		L.wr = o;
		o += align16u(Tp * 2u);
		L.wd = o;
		o += align16u(Tp * 2u);
		L.blk = o;
		o += align16u(((T + 31u) >> 5) * 4u);
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
	// fused generator: elevations, flags and the union-find parents of one map live in the distance layers' memory
	// before the layers are initialised
	__host__ __device__ inline uint32_t small_gen_parent_offset(int W, int H)
	{
		return 2u * align16u(((uint32_t)(W * H) + 8u) * 2u) + 32u;
	}
	__host__ __device__ inline bool small_gen_fits(int W, int H, int R)
	{
		return small_gen_parent_offset(W, H) + (uint32_t)(W * H) * 4u <= (uint32_t)R * (uint32_t)((W + 1) * (H + 2)) * 4u;
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

#ifdef MTD_TIMING
	// developer probe (never in the shipped build): cycle split of one warp's relaxation rounds, printed by CTA 0
	__device__ unsigned long long g_timing[8];
#define MTD_TICK(k)                                           \
	do                                                        \
	{                                                         \
		const long long now__ = clock64();                    \
		acc__[k] += (unsigned long long)(now__ - last__);     \
		last__ = now__;                                       \
	} while (0)
	__device__ unsigned long long g_phase[24];
#define MTD_PHASE(k)                                                   \
	do                                                                 \
	{                                                                  \
		if (blockIdx.x == 0 && threadIdx.x == 0)                       \
		{                                                              \
			const long long now__ = clock64();                         \
			g_phase[k] += (unsigned long long)(now__ - phase_last__);  \
			phase_last__ = now__;                                      \
		}                                                              \
	} while (0)
#else
#define MTD_TICK(k)
#define MTD_PHASE(k)
#endif

	// ---- one layer, GW warps ---------------------------------------------------------------------------------------
	// dist/wr/wd are in the padded frame: neighbours of p are p-1, p+1, p-Wp, p+Wp; parked cells (blocked, padding) hold 0
	// and can never be lowered. ctl: [parity][0..3] = entries each warp appended for the round that reads `parity`,
	// [parity][4..7] = min pending label each warp saw. q: [parity][warp][qseg] linear cell indices. Initial state
	// (parity 0) is written by the caller. Returns true if a label hit the guard band.
	template<int GW>
	__device__ bool group_relax(uint32_t *dist, const uint16_t *wr, const uint16_t *wd, uint16_t *q, uint32_t *ctl, int Wp, int Tp, int qseg, int gw, uint32_t lane, int bar_id)
	{
		const uint32_t lt = lanemask_lt();
		uint32_t parity = 0, thr_prev = 0;
		bool guard = false, dense = false;
#ifdef MTD_TIMING
		const bool probe__ = blockIdx.x == 0 && bar_id == 2 && gw == 0 && lane == 0;
		long long last__ = clock64();
		unsigned long long acc__[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
#endif
		for (;;)
		{
			MTD_TICK(4);
			const uint4 cn = *reinterpret_cast<const uint4 *>(ctl + parity * 8);
			const uint4 mn = *reinterpret_cast<const uint4 *>(ctl + parity * 8 + 4);
			const uint32_t dmin = min(min(mn.x, mn.y), min(mn.z, mn.w));
			if (dmin == MTD_INF)
				break; // nothing pending
			const uint32_t thr = dmin + MIN_EDGE; // dmin < SMALL_SETTLE_LIMIT or every pending cell is reported below
			const uint32_t cap = (uint32_t)qseg;
			dense = dense || cn.x > cap || cn.y > cap || cn.z > cap || cn.w > cap; // an append was dropped: the queue is incomplete from now on
			uint32_t newmin = MTD_INF, outn = 0;
			uint16_t *wq = q + ((parity ^ 1u) * GW + (uint32_t)gw) * cap;
			MTD_TICK(0);
#ifdef MTD_TIMING
			acc__[5] += 1;
			acc__[6] += cn.x + cn.y + cn.z + cn.w;
#endif
			if (!dense)
			{
				const uint32_t p1 = cn.x, p2 = p1 + cn.y, p3 = p2 + cn.z, n = p3 + cn.w;
				const uint16_t *rq = q + parity * GW * cap;
				for (uint32_t base = (uint32_t)gw * 32u; base < n; base += GW * 32u)
				{
					const uint32_t idx = base + lane;
					const bool active = idx < n;
					uint32_t slot = idx;
					if (GW > 1)
					{
						const uint32_t seg = (idx >= p1) + (GW > 2 ? (idx >= p2) + (idx >= p3) : 0u);
						slot = seg * cap + idx - (seg == 0 ? 0u : seg == 1 ? p1 : seg == 2 ? p2 : p3);
					}
					// idle lanes read an in-frame cell too (loads without a branch); the label and the four edge weights are
					// independent loads
					const uint32_t c = active ? rq[slot] : (uint32_t)Wp + 1u;
					const uint32_t dc = dist[c];
					// 0: up, 1: left, 2: right, 3: down -- weights are stored on the upper / left cell of an edge
					const uint32_t w0 = wd[c - Wp], w1 = wr[c - 1], w2 = wr[c], w3 = wd[c];
					const uint32_t d = active ? dc : MTD_INF;
					bool settle = d < thr;
					if (settle && d >= SMALL_SETTLE_LIMIT)
					{
						guard = true; // reference arithmetic is about to wrap: report, do not expand
						settle = false;
					}
					const bool keep = active && d >= thr && d < SMALL_SETTLE_LIMIT;
					// no branch around the atomics: a lane with nothing to settle offers MTD_INF (which lowers nothing) to a parked
					// cell of the top padding row, one per lane so that the idle lanes do not serialise on one word
					const uint32_t base_nd = d + MIN_EDGE;
					const uint32_t n0 = settle ? base_nd + w0 : MTD_INF, n1 = settle ? base_nd + w1 : MTD_INF;
					const uint32_t n2 = settle ? base_nd + w2 : MTD_INF, n3 = settle ? base_nd + w3 : MTD_INF;
					const uint32_t spare = lane < (uint32_t)Wp ? lane : 0u;
					const uint32_t o0 = atomicMin(&dist[settle ? c - Wp : spare], n0);
					const uint32_t o1 = atomicMin(&dist[settle ? c - 1u : spare], n1);
					const uint32_t o2 = atomicMin(&dist[settle ? c + 1u : spare], n2);
					const uint32_t o3 = atomicMin(&dist[settle ? c + Wp : spare], n3);
					// a cell enters the queue exactly once: when its label leaves "unreached" (labels of queued cells may
					// still improve before they settle, settled labels are final)
					const bool f0 = settle && o0 == MTD_INF, f1 = settle && o1 == MTD_INF, f2 = settle && o2 == MTD_INF, f3 = settle && o3 == MTD_INF;
					uint32_t lm = keep ? d : MTD_INF;
					lm = n0 < o0 ? min(lm, n0) : lm; // never true for an idle lane: it offered MTD_INF
					lm = n1 < o1 ? min(lm, n1) : lm;
					lm = n2 < o2 ? min(lm, n2) : lm;
					lm = n3 < o3 ? min(lm, n3) : lm;
					newmin = min(newmin, lm);
					// ---- append survivors and newly reached cells to this warp's segment (ballot compaction) ----
					const uint32_t bk = __ballot_sync(FULL_MASK, keep);
					const uint32_t b0 = __ballot_sync(FULL_MASK, f0);
					const uint32_t b1 = __ballot_sync(FULL_MASK, f1);
					const uint32_t b2 = __ballot_sync(FULL_MASK, f2);
					const uint32_t b3 = __ballot_sync(FULL_MASK, f3);
					uint32_t pos = outn + __popc(bk & lt);
					if (keep && pos < cap)
						wq[pos] = (uint16_t)c;
					outn += __popc(bk);
					pos = outn + __popc(b0 & lt);
					if (f0 && pos < cap)
						wq[pos] = (uint16_t)(c - Wp);
					outn += __popc(b0);
					pos = outn + __popc(b1 & lt);
					if (f1 && pos < cap)
						wq[pos] = (uint16_t)(c - 1u);
					outn += __popc(b1);
					pos = outn + __popc(b2 & lt);
					if (f2 && pos < cap)
						wq[pos] = (uint16_t)(c + 1u);
					outn += __popc(b2);
					pos = outn + __popc(b3 & lt);
					if (f3 && pos < cap)
						wq[pos] = (uint16_t)(c + Wp);
					outn += __popc(b3); // may exceed cap: that is the overflow signal read at the next round
				}
			}
			else
			{
				// dense round: every pending cell has thr_prev <= label < INF (settled ones are below thr_prev, parked
				// cells are 0), so the bucket is found by scanning the labels -- no queue, no capacity
				outn = cap + 1u; // stay dense
				for (int c = gw * 32 + (int)lane; c < Tp; c += GW * 32)
				{
					const uint32_t d = dist[c];
					bool settle = d >= thr_prev && d < thr;
					if (settle && d >= SMALL_SETTLE_LIMIT)
					{
						guard = true;
						settle = false;
					}
					if (d >= thr && d < SMALL_SETTLE_LIMIT)
						newmin = min(newmin, d);
					else if (d >= SMALL_SETTLE_LIMIT && d != MTD_INF)
						guard = true;
					if (settle)
					{
						// thr_prev > 0 in every dense round, so parked cells (and the padding they include) never settle:
						// c - Wp ... c + Wp stay inside the frame
						const uint32_t base_nd = d + MIN_EDGE;
						const uint32_t n0 = base_nd + wd[c - Wp], n1 = base_nd + wr[c - 1], n2 = base_nd + wr[c], n3 = base_nd + wd[c];
						if (n0 < atomicMin(&dist[c - Wp], n0))
							newmin = min(newmin, n0);
						if (n1 < atomicMin(&dist[c - 1], n1))
							newmin = min(newmin, n1);
						if (n2 < atomicMin(&dist[c + 1], n2))
							newmin = min(newmin, n2);
						if (n3 < atomicMin(&dist[c + Wp], n3))
							newmin = min(newmin, n3);
					}
				}
			}
			MTD_TICK(1);
			newmin = warp_min_u32(newmin);
			if (lane == 0)
			{
				ctl[(parity ^ 1u) * 8 + (uint32_t)gw] = outn;
				ctl[(parity ^ 1u) * 8 + 4 + (uint32_t)gw] = newmin;
			}
			MTD_TICK(2);
			if (GW == 1)
				__syncwarp();
			else
				group_barrier(bar_id, GW * 32);
			MTD_TICK(3);
			thr_prev = thr;
			parity ^= 1u;
		}
#ifdef MTD_TIMING
		if (probe__)
			for (int k = 0; k < 8; k++)
				g_timing[k] += acc__[k];
#endif
		return __any_sync(FULL_MASK, guard);
	}

	// ---- incremental re-pathing on one layer in the padded frame (one warp) -------------------------------------------
	// Labels of cells that belong to the subtree under repair carry INC_MARK; inc_lab() strips it. All finite labels are
	// below INC_MARK in this mode (checked when the previous layers are loaded).
	constexpr uint32_t INC_MARK = 0x80000000u;
	constexpr uint32_t INC_MAX_LEVELS = 20;

	__device__ __forceinline__ uint32_t inc_lab(uint32_t v)
	{
		return v == MTD_INF ? MTD_INF : (v & ~INC_MARK);
	}

	// is u the next hop of v (label dv)? Closed form of the heap order as in next_hop_p, on labels that may be marked.
	__device__ __forceinline__ bool inc_is_child(const uint32_t *dl, const uint16_t *wr, const uint16_t *wd, int Wp, uint32_t v, uint32_t dv, uint32_t u, uint32_t srcp)
	{
		const uint32_t u0 = v - (uint32_t)Wp, u1 = v - 1u, u2 = v + 1u, u3 = v + (uint32_t)Wp;
		const uint32_t d0 = inc_lab(dl[u0]), d1 = inc_lab(dl[u1]), d2 = inc_lab(dl[u2]), d3 = inc_lab(dl[u3]);
		const uint32_t w0 = wd[u0], w1 = wr[u1], w2 = wr[v], w3 = wd[v];
		uint32_t best = MTD_INF, bestd = MTD_INF;
		if (d0 != MTD_INF && d0 + MIN_EDGE + w0 == dv && (d0 != 0u || u0 == srcp))
		{
			bestd = d0;
			best = u0;
		}
		if (d1 != MTD_INF && d1 + MIN_EDGE + w1 == dv && d1 < bestd && (d1 != 0u || u1 == srcp))
		{
			bestd = d1;
			best = u1;
		}
		if (d2 != MTD_INF && d2 + MIN_EDGE + w2 == dv && d2 < bestd && (d2 != 0u || u2 == srcp))
		{
			bestd = d2;
			best = u2;
		}
		if (d3 != MTD_INF && d3 + MIN_EDGE + w3 == dv && d3 < bestd && (d3 != 0u || u3 == srcp))
		{
			bestd = d3;
			best = u3;
		}
		return best == u;
	}

	// smallest label the unmarked neighbours offer cell v (MTD_INF: none). A parked neighbour holds 0 like the source does.
	__device__ __forceinline__ uint32_t inc_offer(const uint32_t *dl, const uint16_t *wr, const uint16_t *wd, int Wp, uint32_t v, uint32_t srcp)
	{
		const uint32_t u[4] = { v - (uint32_t)Wp, v - 1u, v + 1u, v + (uint32_t)Wp };
		const uint32_t w[4] = { wd[u[0]], wr[u[1]], wr[v], wd[v] };
		uint32_t best = MTD_INF;
#pragma unroll
		for (int k = 0; k < 4; k++)
		{
			const uint32_t d = dl[u[k]];
			if (d == MTD_INF || (d & INC_MARK) || (d == 0u && u[k] != srcp))
				continue;
			best = min(best, d + MIN_EDGE + w[k]);
		}
		return best;
	}

	// Repairs layer `dl` after tile bp flipped. now_blocked: the subtree of bp in the next-hop tree (every tile whose path ran
	// through bp) is cut out and its rim seeds the bucket queue that group_relax then settles; labels elsewhere cannot
	// change. !now_blocked: bp gets its label from its neighbours and the improvement is propagated (label-correcting, one
	// warp) -- labels only drop, and only behind bp. Returns false when the repair does not fit (lists full, labels too large,
	// bp is the source): the caller then recomputes the layer from scratch. list: cap entries of scratch (the layer's queue).
	__device__ bool inc_repair(uint32_t *dl, const uint16_t *wr, const uint16_t *wd, int Wp, uint32_t bp, uint32_t srcp, bool now_blocked, uint32_t d_b_old, uint16_t *list, uint32_t cap, uint32_t qseg, uint32_t *c, uint32_t lane)
	{
		const uint32_t lt = lanemask_lt();
		uint32_t nseed = 0, seedmin = MTD_INF;
		bool ok = true;
		if (bp == srcp)
			return false;
		if (now_blocked)
		{
			if (d_b_old != MTD_INF)
			{
				// ---- subtree of bp, breadth first; list[0 .. n) collects it ----
				if (lane == 0)
				{
					dl[bp] = d_b_old | INC_MARK; // parked by the caller; its old label is what its children were computed from
					list[0] = (uint16_t)bp;
				}
				__syncwarp();
				uint32_t f0 = 0, n = 1, levels = 0;
				while (f0 < n && ok)
				{
					// a level costs ~1 us of dependent shared-memory accesses on one warp, a bucket of the full update ~0.6 us
					// for the whole layer: a subtree deeper than this is cheaper to recompute (r02m: the first path tile)
					if (++levels > INC_MAX_LEVELS)
						return false;
					const uint32_t f1 = n;
					for (uint32_t base = f0; base < f1; base += 32u)
					{
						const uint32_t idx = base + lane;
						const bool act = idx < f1;
						const uint32_t u = act ? list[idx] : bp;
#pragma unroll
						for (int k = 0; k < 4; k++)
						{
							const uint32_t v = k == 0 ? u - (uint32_t)Wp : k == 1 ? u - 1u : k == 2 ? u + 1u : u + (uint32_t)Wp;
							const uint32_t dv = dl[v];
							const bool child = act && dv != 0u && dv != MTD_INF && !(dv & INC_MARK) && inc_is_child(dl, wr, wd, Wp, v, dv, u, srcp);
							const uint32_t bal = __ballot_sync(FULL_MASK, child);
							if (child)
							{
								const uint32_t pos = n + __popc(bal & lt);
								if (pos < cap)
									list[pos] = (uint16_t)v;
								dl[v] = dv | INC_MARK; // a tile has one next hop: nobody else marks v
							}
							n += __popc(bal);
							__syncwarp();
						}
					}
					if (n > cap)
						ok = false;
					f0 = f1;
				}
				if (!ok)
					return false;
				// ---- rim: what the untouched neighbours offer each cut-out tile (marks stay on while everybody reads) ----
				for (uint32_t base = 1; base < n; base += 32u)
				{
					const uint32_t idx = base + lane;
					if (idx < n)
					{
						const uint32_t v = list[idx];
						const uint32_t nl = inc_offer(dl, wr, wd, Wp, v, srcp);
						if (nl != MTD_INF && nl >= SMALL_SETTLE_LIMIT)
							ok = false;
						dl[v] = nl == MTD_INF ? MTD_INF : (nl | INC_MARK);
					}
				}
				__syncwarp();
				if (!__all_sync(FULL_MASK, ok))
					return false;
				// ---- marks off; tiles with a label are the seeds (compacted in place: seeds never outrun the reads) ----
				for (uint32_t base = 1; base < n; base += 32u)
				{
					const uint32_t idx = base + lane;
					uint32_t v = 0, val = MTD_INF;
					if (idx < n)
					{
						v = list[idx];
						val = inc_lab(dl[v]);
						dl[v] = val;
					}
					const bool seed = val != MTD_INF;
					const uint32_t bal = __ballot_sync(FULL_MASK, seed);
					__syncwarp();
					if (seed)
					{
						const uint32_t pos = nseed + __popc(bal & lt);
						if (pos < qseg)
							list[pos] = (uint16_t)v;
						seedmin = min(seedmin, val);
					}
					nseed += __popc(bal);
					__syncwarp();
				}
				if (lane == 0)
					dl[bp] = 0u; // parked again
				if (nseed > qseg)
					return false;
				seedmin = warp_min_u32(seedmin);
			}
		}
		else
		{
			// bp was blocked (label m in the previous layers), now it is a tile like any other
			if (d_b_old != MTD_INF)
				return false; // it already had a label: not the single flip this mode is for
			const uint32_t nl = inc_offer(dl, wr, wd, Wp, bp, srcp);
			if (nl != MTD_INF)
			{
				if (nl >= SMALL_SETTLE_LIMIT)
					return false;
				if (lane == 0)
				{
					dl[bp] = nl;
					list[0] = (uint16_t)bp;
				}
				__syncwarp();
				// label-correcting propagation, two lists in the halves of the scratch
				const uint32_t half = cap / 2u;
				uint32_t cur = 0, ncur = 1;
				while (ncur != 0u)
				{
					const uint16_t *rq = list + cur * half;
					uint16_t *wq = list + (cur ^ 1u) * half;
					uint32_t nn = 0;
					for (uint32_t base = 0; base < ncur; base += 32u)
					{
						const uint32_t idx = base + lane;
						const bool act = idx < ncur;
						const uint32_t cc = act ? rq[idx] : bp;
						const uint32_t d = dl[cc];
						const uint32_t base_nd = d + MIN_EDGE;
						if (act && base_nd >= SMALL_SETTLE_LIMIT)
							ok = false;
#pragma unroll
						for (int k = 0; k < 4; k++)
						{
							const uint32_t v = k == 0 ? cc - (uint32_t)Wp : k == 1 ? cc - 1u : k == 2 ? cc + 1u : cc + (uint32_t)Wp;
							const uint32_t w = k == 0 ? wd[v] : k == 1 ? wr[v] : k == 2 ? wr[cc] : wd[cc];
							const uint32_t nd = base_nd + w;
							bool imp = false;
							if (act && ok)
								imp = nd < atomicMin(&dl[v], nd); // parked cells hold 0: never improved
							const uint32_t bal = __ballot_sync(FULL_MASK, imp);
							if (imp)
							{
								const uint32_t pos = nn + __popc(bal & lt);
								if (pos < half)
									wq[pos] = (uint16_t)v;
							}
							nn += __popc(bal);
						}
					}
					__syncwarp();
					if (nn > half || !__all_sync(FULL_MASK, ok))
						return false;
					ncur = nn;
					cur ^= 1u;
				}
			}
		}
		if (lane == 0)
		{
			c[0] = nseed;
			c[1] = c[2] = c[3] = 0u;
			c[4] = seedmin;
			c[5] = c[6] = c[7] = MTD_INF;
			c[8] = c[9] = c[10] = c[11] = 0u;
			c[12] = c[13] = c[14] = c[15] = MTD_INF;
		}
		return true;
	}

	// tiles[v] from final labels in the padded frame (closed form of the heap's tie-break, sssp_warp.cuh): the neighbour u
	// with d[u] + w(u, v) == d[v] minimising (d[u], u); neighbours in ascending tile index (up, left, right, down), replace
	// only on a strictly smaller d[u]. dv = dl[p] must be a reached, non-source label (not 0, not INF). A parked neighbour
	// holds 0 like a source does: is_src(u) tells them apart. Returns a padded index or MTD_INF.
	template<class IsSrc>
	__device__ __forceinline__ uint32_t next_hop_p(const uint32_t *dl, const uint16_t *wr, const uint16_t *wd, int Wp, uint32_t p, uint32_t dv, IsSrc is_src)
	{
		const uint32_t u0 = p - (uint32_t)Wp, u1 = p - 1u, u2 = p + 1u, u3 = p + (uint32_t)Wp;
		const uint32_t d0 = dl[u0], d1 = dl[u1], d2 = dl[u2], d3 = dl[u3];
		const uint32_t w0 = wd[u0], w1 = wr[u1], w2 = wr[p], w3 = wd[p];
		uint32_t best = MTD_INF, bestd = MTD_INF;
		if (d0 != MTD_INF && d0 + MIN_EDGE + w0 == dv && (d0 != 0u || is_src(u0)))
		{
			bestd = d0;
			best = u0;
		}
		if (d1 != MTD_INF && d1 + MIN_EDGE + w1 == dv && d1 < bestd && (d1 != 0u || is_src(u1)))
		{
			bestd = d1;
			best = u1;
		}
		if (d2 != MTD_INF && d2 + MIN_EDGE + w2 == dv && d2 < bestd && (d2 != 0u || is_src(u2)))
		{
			bestd = d2;
			best = u2;
		}
		if (d3 != MTD_INF && d3 + MIN_EDGE + w3 == dv && d3 < bestd && (d3 != 0u || is_src(u3)))
		{
			bestd = d3;
			best = u3;
		}
		return best;
	}

	// layers held in shared memory (padded frame, blocked tiles still parked at 0), single-source layers:
	// srcp[l] = padded index of layer l's source
	struct SmemLayersAcc
	{
		const uint32_t *sdist; // [R][Tp]
		const uint16_t *wr, *wd;
		const uint32_t *srcp;
		SmallGeom G;
		__device__ __forceinline__ SmemLayersAcc(const SmallGeom &g) : G(g) {}
		__device__ __forceinline__ uint32_t dist(uint32_t l, uint32_t t) const
		{
			const uint32_t p = G.g2p(t);
			const uint32_t v = sdist[(size_t)l * G.Tp + p];
			return v == 0u && p != srcp[l] ? MTD_INF : v; // parked = blocked, never relaxed (paths.cpp:65-67)
		}
		__device__ __forceinline__ uint32_t hop(uint32_t l, uint32_t t) const
		{
			const uint32_t p = G.g2p(t);
			const uint32_t *dl = sdist + (size_t)l * G.Tp;
			const uint32_t dv = dl[p];
			if (dv == MTD_INF || dv == 0u)
				return MTD_INF;
			const uint32_t sp = srcp[l];
			const uint32_t q = next_hop_p(dl, wr, wd, G.Wp, p, dv, [&](uint32_t u) { return u == sp; });
			return q == MTD_INF ? MTD_INF : G.p2g(q);
		}
		__device__ __forceinline__ uint32_t edge(uint32_t a, uint32_t b) const
		{
			const uint32_t pa = G.g2p(a), pb = G.g2p(b);
			const uint32_t lo = pa < pb ? pa : pb, hi = pa < pb ? pb : pa;
			return MIN_EDGE + (hi - lo == 1u ? wr[lo] : wd[lo]);
		}
	};

	__device__ __forceinline__ uint32_t ld_volatile_shared(const uint32_t *p)
	{
		return *(const volatile uint32_t *)p;
	}
	__device__ __forceinline__ void st_volatile_shared(uint32_t *p, uint32_t v)
	{
		*(volatile uint32_t *)p = v;
	}
	// barrier of the 512 worker threads (the walker warps never take part)
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
	// 512 worker threads + SMALL_WALKERS walker warps. GW = warps cooperating on one layer during the relaxation.
	// The walkers are decoupled from the workers: a finished map (layers written to global memory) is published in a small
	// ring in shared memory, entry k is walked by walker k % SMALL_WALKERS while the workers go on with the next maps.
	template<int GW>
	__global__ void __launch_bounds__(SMALL_BLOCK, 1) k_waypoints_small(const WaypointsSmallParams P)
	{
		extern __shared__ __align__(16) unsigned char smem_raw[];
		const SmallGeom G(P.W, P.H);
		const int W = G.W, H = G.H, T = G.T, Wp = G.Wp, Tp = G.Tp, n = P.n, R = P.R;
		const SmallLayout L = small_layout(W, H, R, P.smem_budget, P.qseg_cap, GW);
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
		__shared__ uint32_t s_wp[MTD_MAXSRC_SMALL], s_srcp[MTD_MAXSRC_SMALL]; // waypoint / source tiles: grid index, padded index
		__shared__ int s_guard, s_inc_fail;
		__shared__ uint32_t s_map, s_chunk;
		// the walker warps' private copies (they work on EARLIER maps while the workers overwrite the above)
		__shared__ uint32_t w_wp[SMALL_WALKERS][FIND_MAXN], w_D[SMALL_WALKERS][FIND_MAXN * FIND_MAXN], w_fd[SMALL_WALKERS][FIND_MAXN];
		__shared__ uint32_t s_ring[SMALL_RING];      // finished maps, entry k in slot k % SMALL_RING
		__shared__ uint32_t s_head, s_stop;          // entries published so far / the workers are done
		__shared__ uint32_t s_walked[SMALL_WALKERS]; // entries each walker completed
		// fused generator
		__shared__ float2 s_grad[128];
		__shared__ noise::ElevationParams s_ep; // of the map whose elevations are computed next
		__shared__ Rng s_rng;                   // the current map's stream after the generator's draws
		__shared__ unsigned long long s_best_key;
		__shared__ uint32_t s_next;

		const int tid = threadIdx.x;
		const int warp = tid >> 5;
		const uint32_t lane = lane_id();
		const bool resident_all = n <= R;
		const int nsrc_layer = P.nsrc > 1 ? P.nsrc : 1;
		const int nwp = P.nsrc > 1 ? P.nsrc : n;
		const uint32_t stage_half = align16u((uint32_t)(T + 8) * 2u);
		const bool want_walks = P.fullpath_len && P.nsrc <= 1;
		uint32_t mbar_phase = 0;
		uint32_t published = 0; // thread 0: entries in the ring so far

		if (tid == 0)
		{
			mbar_init(&s_mbar, 1);
			asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
			s_head = 0;
			s_stop = 0;
			for (int w = 0; w < SMALL_WALKERS; w++)
				s_walked[w] = 0;
		}
		__syncthreads();

		if (warp >= SMALL_WALKER_WARP)
		{
			// ---- walker warps: paths of finished maps, from their layers in global memory (still in L2) ----
			const int w = warp - SMALL_WALKER_WARP;
			for (uint32_t k = (uint32_t)w;; k += SMALL_WALKERS)
			{
				uint32_t head;
				for (;;)
				{
					const uint32_t stop = ld_volatile_shared(&s_stop);
					head = ld_volatile_shared(&s_head); // read after the stop flag: a set flag means the last entry is visible
					if (head > k || stop)
						break;
					__nanosleep(200);
				}
				if (head <= k)
					break;
				__threadfence(); // acquire: the workers' stores of this map's layers
				walk_map_global(P, (int)ld_volatile_shared(&s_ring[k % SMALL_RING]), w_wp[w], w_D[w], w_fd[w], lane);
				if (lane == 0)
					st_volatile_shared(&s_walked[w], k / SMALL_WALKERS + 1u);
			}
			return;
		}

		// ---- fused generator: the first map's elevations are computed by all workers, every later map's by the warps that
		//      have no layer to relax while the map before it is relaxed ----
		const bool gen = P.gen != 0;
		MapgenShape S;
		S.W = W;
		S.H = H;
		S.ox = P.gen_ox;
		S.oy = P.gen_oy;
		S.radius = P.gen_radius;
		S.magicW = G.magicW;
		int16_t *const gen_elev = const_cast<int16_t *>(P.elev);
		uint16_t *const gen_flags = const_cast<uint16_t *>(P.flags);
		uint32_t *const gen_wp = const_cast<uint32_t *>(P.wp);
		int gen_cur = 0;
		const int gen_warps = min(max(P.gen_warps, 0), SMALL_THREADS / 32 - min(n, R) * GW); // warps that compute elevations during the relaxation
		if (gen)
		{
			for (int i = tid; i < 128; i += SMALL_THREADS)
				s_grad[i] = noise::gradient(i);
			if (tid == 0)
			{
				const uint32_t m = atomicAdd(P.map_counter, 1u);
				s_next = m;
				if (m < (uint32_t)P.nmaps)
				{
					Rng g;
					noise::ElevationParams e;
					mapgen_draws(P.gen_s1, P.gen_s2, P.gen_first + (int)m, e, g);
					s_ep = e;
				}
			}
			workers_sync();
			gen_cur = (int)s_next;
			if (gen_cur < P.nmaps)
			{
				const noise::ElevationParams lp = s_ep;
				mapgen_elevations(s_grad, lp, S, gen_elev + (size_t)gen_cur * T, tid, SMALL_THREADS);
			}
		}

#ifdef MTD_TIMING
		long long phase_last__ = clock64();
#endif
		uint32_t claim = 0; // thread 0: the next map of this CTA
		if (tid == 0 && !gen)
			claim = atomicAdd(P.map_counter, 1u);
		for (;;)
		{
			workers_sync(); // the previous map is completely written out (and this map's elevations, when they are generated here)
			if (tid == 0)
			{
				if (!gen)
					s_map = claim; // issued during the previous map's write-out: the L2 round trip is off the critical path
				else if (gen_cur < P.nmaps)
				{
					// claim the map after this one now: its elevations are computed while this one is relaxed
					Rng g;
					noise::ElevationParams e;
					mapgen_draws(P.gen_s1, P.gen_s2, P.gen_first + gen_cur, e, g);
					s_rng = g;
					const uint32_t m = atomicAdd(P.map_counter, 1u);
					s_next = m;
					if (m < (uint32_t)P.nmaps)
					{
						mapgen_draws(P.gen_s1, P.gen_s2, P.gen_first + (int)m, e, g);
						s_ep = e;
					}
				}
			}
			workers_sync();
			const int mapi = gen ? gen_cur : (int)s_map;
			MTD_PHASE(0); // claiming a map
			if (mapi >= P.nmaps)
				break;
			const int gen_next = gen ? (int)s_next : P.nmaps;
			gen_cur = gen_next;
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
			if (gen)
			{
				// ---- 1g. the rest of the generator, in the memory the distance layers take afterwards ----
				int16_t *se = (int16_t *)st_e;
				uint16_t *sf = (uint16_t *)st_f;
				uint32_t *parent = (uint32_t *)(smem_raw + L.dist + small_gen_parent_offset(W, H));
				for (int i0 = tid; i0 < T; i0 += 4 * SMALL_THREADS)
				{
					// written by this CTA one map ago; four independent loads in flight per thread
					int16_t v[4];
#pragma unroll
					for (int k = 0; k < 4; k++)
						v[k] = i0 + k * SMALL_THREADS < T ? __ldcg(ge + i0 + k * SMALL_THREADS) : (int16_t)0;
#pragma unroll
					for (int k = 0; k < 4; k++)
						if (i0 + k * SMALL_THREADS < T)
							se[i0 + k * SMALL_THREADS] = v[k];
				}
				if (warp == 2 && lane == 0)
				{
					s_guard = 0;
					s_chunk = 0;
				}
				else if (warp == 3)
				{
					for (int k = (int)lane; k < Wp; k += 32)
					{
						wr[k] = 0;
						wd[k] = 0;
						wr[(H + 1) * Wp + k] = 0;
						wd[(H + 1) * Wp + k] = 0;
					}
					for (int y = 1 + (int)lane; y <= H; y += 32)
					{
						wr[y * Wp + W] = 0;
						wd[y * Wp + W] = 0;
					}
				}
				workers_sync();
				MTD_PHASE(6); // generator: elevations back from L2
				const bool any_valid = mapgen_flags(S, se, sf, parent, &s_best_key, tid, SMALL_THREADS, [] { workers_sync(); });
				MTD_PHASE(7); // generator: flags, components, border
				if (tid == 0)
				{
					mapgen_pick_spawns(s_rng, T, sf, nwp, any_valid, s_wp, P.status);
					for (int k = 0; k < nwp; k++)
					{
						const uint32_t t = s_wp[k];
						gen_wp[(size_t)mapi * nwp + k] = t;
						s_srcp[k] = G.g2p(t < (uint32_t)T ? t : 0u); // a map without valid tiles is reported (ST_NO_VALID), not relaxed into
					}
				}
				workers_sync();
				MTD_PHASE(8); // generator: spawn picking
				uint16_t *gfo = gen_flags + (size_t)mapi * T;
				for (int i = tid; i < T; i += SMALL_THREADS)
					gfo[i] = sf[i];
			}
			else
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
					{
						const uint32_t t = gwp[i];
						s_wp[i] = t;
						s_srcp[i] = G.g2p(t < (uint32_t)T ? t : 0u); // the host validates waypoints; MTD_INF comes from a generated map without valid tiles (reported)
					}
					if (lane == 0)
					{
						s_guard = 0;
						s_chunk = 0;
					}
				}
				else if (warp == 3)
				{
					// weights of the padding (never matter: their far end is parked) are zeroed once per map for determinism
					for (int k = (int)lane; k < Wp; k += 32)
					{
						wr[k] = 0;
						wd[k] = 0;
						wr[(H + 1) * Wp + k] = 0;
						wd[(H + 1) * Wp + k] = 0;
					}
					for (int y = 1 + (int)lane; y <= H; y += 32)
					{
						wr[y * Wp + W] = 0;
						wd[y * Wp + W] = 0;
					}
				}
				mbar_wait(&s_mbar, mbar_phase);
				mbar_phase ^= 1u;
				workers_sync();
			}
			MTD_PHASE(1); // staging

			// ---- 2. edge weights (Grid::neighborDistance of the right and down edge, grid.cpp:141-145) + blocking bitmap ----
			for (int base = 0; base < T; base += SMALL_THREADS)
			{
				const int i = base + tid;
				bool blocked = false;
				if (i < T)
				{
					const int y = (int)G.row_of((uint32_t)i), x = i - y * W;
					const int p = i + y + Wp;
					const int16_t e = selev[i];
					wr[p] = x < W - 1 ? (uint16_t)(edge_weight(e, selev[i + 1]) - MIN_EDGE) : (uint16_t)0;
					wd[p] = y < H - 1 ? (uint16_t)(edge_weight(e, selev[i + W]) - MIN_EDGE) : (uint16_t)0;
					blocked = (sflags[i] & F_BLOCKING) != 0 || (uint32_t)i == P.extra_blocked;
				}
				const uint32_t bal = __ballot_sync(FULL_MASK, blocked);
				if (lane == 0 && base + warp * 32 < T)
					blk[(base >> 5) + warp] = bal;
			}
			workers_sync(); // staging is dead from here on: the distance layers take its place
			MTD_PHASE(2); // weights

			for (int l0 = 0; l0 < n; l0 += R)
			{
				const int lb = (n - l0) < R ? (n - l0) : R; // layers in this pass
				// ---- 3a'. incremental: the previous layers are the starting point, one warp per layer repairs what the flipped
				//      tile changed (inc_repair); anything it cannot handle falls back to the full update below ----
				bool inc = P.inc_tile != MTD_M && P.prev_dist && resident_all && P.nsrc <= 1;
				if (inc)
				{
					if (tid == 0)
						s_inc_fail = 0;
					workers_sync();
					bool big = false;
					for (int p = tid; p < Tp; p += SMALL_THREADS)
					{
						const uint32_t yp = __umulhi((uint32_t)p, G.magicWp), x = (uint32_t)p - yp * (uint32_t)Wp;
						const bool cell = yp >= 1u && yp <= (uint32_t)H && x < (uint32_t)W;
						const uint32_t i = cell ? (uint32_t)p - yp - (uint32_t)W : 0u;
						const bool blocked = cell && ((blk[i >> 5] >> (i & 31u)) & 1u);
						for (int r = 0; r < lb; r++)
						{
							uint32_t v = 0u;
							if (cell && !blocked)
							{
								v = __ldcg(P.prev_dist + (size_t)(l0 + r) * T + i);
								big |= v != MTD_INF && v >= INC_MARK - (1u << 17);
							}
							sdist[(size_t)r * Tp + p] = v;
						}
					}
					if (big)
						s_inc_fail = 1;
					workers_sync();
					if (warp < lb * GW && (warp % GW) == 0 && !s_inc_fail)
					{
						const int g = warp / GW;
						const uint32_t b = P.inc_tile;
						const bool now_blocked = (blk[b >> 5] >> (b & 31u)) & 1u;
						const uint32_t d_b_old = __ldcg(P.prev_dist + (size_t)(l0 + g) * T + b);
						if (!inc_repair(sdist + (size_t)g * Tp, wr, wd, Wp, G.g2p(b), s_srcp[l0 + g], now_blocked, d_b_old, q + (size_t)g * 2 * GW * qseg, (uint32_t)(2 * GW * qseg), (uint32_t)qseg, ctl + g * 16, lane) && lane == 0)
							s_inc_fail = 1;
					}
					workers_sync();
					inc = !s_inc_fail;
				}
				if (!inc)
				{
					// ---- 3a. init: paths.cpp:40-43; blocked tiles and the padding are parked at 0 (never relaxed: >= 255 cannot lower 0) ----
					for (int p = tid; p < Tp; p += SMALL_THREADS)
					{
						const uint32_t yp = __umulhi((uint32_t)p, G.magicWp), x = (uint32_t)p - yp * (uint32_t)Wp;
						uint32_t v = 0u;
						if (yp >= 1u && yp <= (uint32_t)H && x < (uint32_t)W)
						{
							const uint32_t i = (uint32_t)p - yp - (uint32_t)W;
							v = ((blk[i >> 5] >> (i & 31u)) & 1u) ? 0u : MTD_INF;
						}
						for (int r = 0; r < lb; r++)
							sdist[(size_t)r * Tp + p] = v;
					}
					workers_sync();
					// seeds: paths.cpp:48-49 (a blocked source still expands: the popped tile is never tested, paths.cpp:63-67)
					if (warp < lb * GW && (warp % GW) == 0 && lane == 0)
					{
						const int g = warp / GW;
						uint32_t *dl = sdist + (size_t)g * Tp;
						uint32_t *c = ctl + g * 16;
						uint16_t *q0 = q + (size_t)g * 2 * GW * qseg;
						uint32_t nseed = 0;
						for (int s = 0; s < nsrc_layer; s++)
						{
							const uint32_t sp = P.nsrc > 1 ? s_srcp[s] : s_srcp[l0 + g];
							bool dup = false;
							for (int s2 = 0; s2 < s; s2++)
								dup |= P.nsrc > 1 && s_srcp[s2] == sp;
							if (dup)
								continue;
							dl[sp] = 0u;
							if ((int)nseed < qseg)
								q0[nseed] = (uint16_t)sp;
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

				}
				MTD_PHASE(3); // init + seeds (or incremental repair)
				// ---- 3b. relaxation: GW warps per layer, one named barrier per bucket ----
				if (warp < lb * GW)
				{
					const int g = warp / GW;
					if (group_relax<GW>(sdist + (size_t)g * Tp, wr, wd, q + (size_t)g * 2 * GW * qseg, ctl + g * 16, Wp, Tp, qseg, warp % GW, lane, 2 + g) && lane == 0)
						s_guard = 1;
				}
				else if (gen_next < P.nmaps && warp < lb * GW + gen_warps)
				{
					// fused generator: warps without a layer compute the next map's elevations (pure arithmetic, no shared
					// memory beyond the gradient table) into its slot of the output, where step 1g picks them up. Not all of
					// them: the relaxation is a chain of dependent instructions and every ready warp next to it delays its issue
					const noise::ElevationParams lp = s_ep;
					mapgen_elevations(s_grad, lp, S, gen_elev + (size_t)gen_next * T, tid - lb * GW * 32, gen_warps * 32);
				}
				workers_sync();
				if (gen_next < P.nmaps && gen_warps == 0)
				{
					// every warp had a layer: the elevations are computed by all of them now
					const noise::ElevationParams lp = s_ep;
					mapgen_elevations(s_grad, lp, S, gen_elev + (size_t)gen_next * T, tid, SMALL_THREADS);
				}
				MTD_PHASE(4); // relaxation (slowest layer)
				if (tid == 0 && !gen && l0 + R >= n)
					claim = atomicAdd(P.map_counter, 1u); // consumed at the top of the loop

				// ---- 4. full-path walks (one lane per spawner) overlapped with next-hop tiles + write-out of this pass ----
				// Blocked tiles are still parked at 0 in the resident layers: every reader below maps "0 and not this layer's
				// source" to m (they were never relaxed: paths.cpp:65-67).
				const bool walk_now = want_walks && !P.walker && resident_all; // single map: walk the resident layers while they are written out
				SmemLayersAcc acc(G);
				acc.sdist = sdist;
				acc.wr = wr;
				acc.wd = wd;
				acc.srcp = s_srcp;
				if (walk_now)
				{
					if (tid < n * n)
					{
						const int a = tid / n, b = tid % n;
						Dtab[a * FIND_MAXN + b] = acc.dist((uint32_t)b, s_wp[a]); // inter-waypoint table D[a][b] = dist_b[wp_a]
					}
					workers_sync();
				}
				uint32_t walk_len = 0;
				if (walk_now && warp < n && lane == 0)
				{
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
							const uint32_t bw = blk[i >> 5]; // warp-uniform: i - lane is a multiple of 32
							if (bw == 0xFFFFFFFFu && P.nsrc <= 1)
							{
								// 32 blocked tiles: m everywhere except a blocked tile that is the layer's own source
								for (int r = 0; r < lb; r++)
								{
									const bool is_src = s_wp[l0 + r] == (uint32_t)i;
									if (gdist)
										st_stream_u32(gdist + (size_t)(l0 + r) * T + i, is_src ? 0u : MTD_INF);
									if (gtiles)
										st_stream_u32(gtiles + (size_t)(l0 + r) * T + i, MTD_INF);
								}
								continue;
							}
							const bool blocked = (bw >> (i & 31)) & 1u;
							const uint32_t p = G.g2p((uint32_t)i);
							if (lb == 4 && P.nsrc <= 1 && gdist && gtiles)
							{
								// the usual case, four single-source layers: the tile's four edge weights are loaded once and the 20
								// label loads of the four layers are independent of each other (the per-layer loop below serialises them)
								uint32_t dv[4], hop[4];
								const uint32_t u0 = p - (uint32_t)Wp, u1 = p - 1u, u2 = p + 1u, u3 = p + (uint32_t)Wp;
								const uint32_t w0 = MIN_EDGE + wd[u0], w1 = MIN_EDGE + wr[u1], w2 = MIN_EDGE + wr[p], w3 = MIN_EDGE + wd[p];
								uint32_t c[4], d0[4], d1[4], d2[4], d3[4];
#pragma unroll
								for (int r = 0; r < 4; r++)
								{
									const uint32_t *dl = sdist + (size_t)r * Tp;
									c[r] = dl[p];
									d0[r] = dl[u0];
									d1[r] = dl[u1];
									d2[r] = dl[u2];
									d3[r] = dl[u3];
								}
#pragma unroll
								for (int r = 0; r < 4; r++)
								{
									const uint32_t sp = s_srcp[l0 + r];
									uint32_t best = MTD_INF, bestd = MTD_INF;
									// same order and strict '<' as next_hop_p: up, left, right, down
									if (d0[r] != MTD_INF && d0[r] + w0 == c[r] && (d0[r] != 0u || u0 == sp))
									{
										bestd = d0[r];
										best = u0;
									}
									if (d1[r] != MTD_INF && d1[r] + w1 == c[r] && d1[r] < bestd && (d1[r] != 0u || u1 == sp))
									{
										bestd = d1[r];
										best = u1;
									}
									if (d2[r] != MTD_INF && d2[r] + w2 == c[r] && d2[r] < bestd && (d2[r] != 0u || u2 == sp))
									{
										bestd = d2[r];
										best = u2;
									}
									if (d3[r] != MTD_INF && d3[r] + w3 == c[r] && d3[r] < bestd && (d3[r] != 0u || u3 == sp))
									{
										bestd = d3[r];
										best = u3;
									}
									const bool reach = !blocked && c[r] != MTD_INF && c[r] != 0u;
									hop[r] = reach && best != MTD_INF ? G.p2g(best) : MTD_INF;
									dv[r] = blocked ? (sp == p ? 0u : MTD_INF) : c[r];
								}
#pragma unroll
								for (int r = 0; r < 4; r++)
								{
									st_stream_u32(gdist + (size_t)(l0 + r) * T + i, dv[r]);
									st_stream_u32(gtiles + (size_t)(l0 + r) * T + i, hop[r]);
								}
								continue;
							}
							for (int r = 0; r < lb; r++)
							{
								const uint32_t *dl = sdist + (size_t)r * Tp;
								uint32_t dv = dl[p], hop = MTD_INF;
								if (blocked)
								{
									// parked at 0: a blocked source keeps 0, every other blocked tile is unreachable
									bool is_src = false;
									if (P.nsrc > 1)
										for (int s = 0; s < nsrc_layer; s++)
											is_src |= s_srcp[s] == p;
									else
										is_src = s_srcp[l0 + r] == p;
									dv = is_src ? 0u : MTD_INF;
								}
								else if (gtiles && dv != MTD_INF && dv != 0u)
								{
									uint32_t hp;
									if (P.nsrc > 1)
										hp = next_hop_p(dl, wr, wd, Wp, p, dv, [&](uint32_t u)
											{
												bool s_ = false;
												for (int s = 0; s < nsrc_layer; s++)
													s_ |= s_srcp[s] == u;
												return s_;
											});
									else
									{
										const uint32_t sp = s_srcp[l0 + r];
										hp = next_hop_p(dl, wr, wd, Wp, p, dv, [&](uint32_t u) { return u == sp; });
									}
									hop = hp == MTD_INF ? MTD_INF : G.p2g(hp);
								}
								if (gdist)
									st_stream_u32(gdist + (size_t)(l0 + r) * T + i, dv);
								if (gtiles)
									st_stream_u32(gtiles + (size_t)(l0 + r) * T + i, hop);
							}
						}
					}
				}
				workers_sync();
				MTD_PHASE(5); // next hops + write-out
				if (tid == 0)
					s_chunk = 0;
			}

			if (s_guard && tid == 0)
				atomicOr(P.status, ST_GUARD);
			if (want_walks && P.walker)
			{
				// hand the map to a walker. The last workers_sync() ordered every worker's stores before this thread; the fence
				// makes them visible to the walker that reads the new head
				if (tid == 0)
				{
					if (published >= SMALL_RING)
					{
						const uint32_t old = published - SMALL_RING; // its slot is reused: wait until that entry has been walked
						while (ld_volatile_shared(&s_walked[old % SMALL_WALKERS]) <= old / SMALL_WALKERS)
							__nanosleep(200);
					}
					st_volatile_shared(&s_ring[published % SMALL_RING], (uint32_t)mapi);
					published++;
					__threadfence();
					st_volatile_shared(&s_head, published);
				}
				continue;
			}
			if (!want_walks)
				continue;

			if (!resident_all)
			{
				// single map with more layers than fit: walk the layers just written to global memory (L2), warp 0
				if (warp == 0)
					walk_map_global(P, mapi, w_wp[0], w_D[0], w_fd[0], lane);
				continue;
			}

			// ---- 5. aggregates: paths.cpp:122-138 ----
			if (tid == 0)
				walk_aggregates(n, s_fd, P.agg + (size_t)mapi * 3);
		}
		if (tid == 0)
		{
			__threadfence();
			st_volatile_shared(&s_stop, 1u); // after the last head update: the walkers drain the ring and leave
		}
#ifdef MTD_TIMING
		if (blockIdx.x == 0 && tid == 0)
		{
			printf("TIMING rounds=%llu items=%llu | ctl+thr=%llu body=%llu redux+sts=%llu barrier=%llu looptop=%llu (cycles, warp 0 of layer 0, CTA 0)\n", g_timing[5], g_timing[6], g_timing[0], g_timing[1], g_timing[2], g_timing[3], g_timing[4]);
			printf("PHASES wait=%llu stage=%llu weights=%llu init=%llu relax=%llu writeout=%llu | generator: load=%llu flags=%llu spawns=%llu (cycles, CTA 0)\n", g_phase[0], g_phase[1], g_phase[2], g_phase[3], g_phase[4], g_phase[5], g_phase[6], g_phase[7], g_phase[8]);
			for (int k = 0; k < 8; k++)
				g_timing[k] = g_phase[k] = 0;
			g_phase[8] = 0;
		}
#endif
	}
}
