// tile_warp.cuh -- one global round of the large-grid relaxation with ONE WARP per (layer, tile) work item.
//
// Same contract as the block-per-tile round it replaces (Directions::update, sources/map/paths.cpp:51-77, as tile-local
// label correction; see big_grid.cuh for the data structures): a work item is staged with a 1-cell apron, relaxed to
// local convergence, written back, and the neighbour tiles whose adjacent cells it can still improve are queued for the
// next round. What changed is the execution shape:
//   * a tile is 32 x 32 cells and belongs to one warp; the 4 warps of a CTA work on 4 different items and never meet:
//     no __syncthreads, no named barriers, no shared-memory atomics on queue tails -- the bucket loop's control state
//     (queue length, smallest pending label) lives in warp-uniform registers (ballot / popc / redux.sync);
//   * per warp 12.9 KB of shared memory (labels, packed weights, two 512-entry queues), 16 items in flight per SM;
//   * measured on the block-per-tile version: ~1200 cycles per bucket for ~38 pending cells, i.e. the visit was a chain of
//     barrier + shared-atomic latencies, not work. A warp-synchronous bucket is a few dependent shared-memory accesses.
#pragma once

#include <cooperative_groups.h>

#include "common.cuh"
#include "sssp_warp.cuh"
// included at the end of big_grid.cuh: BigParams, activate_item, cp_async_* and the BIG_* constants come from there

namespace mtd
{
	template<int TS>
	struct WarpFrame
	{
		static constexpr int RW = TS + 2;    // frame rows / columns (interior + apron)
		static constexpr int PITCH = TS + 8; // words per frame row; frame column rx lives at word rx + 3 (interior starts 16-byte aligned)
		static constexpr int CELLS = RW * PITCH;
		static constexpr int QCAP = 512;
		uint32_t d[CELLS];     // labels
		uint32_t w[CELLS];     // packed edge weights of the same cells: (w_right - 255) | (w_down - 255) << 16
		uint16_t q[2][QCAP];   // pending cells (frame word index), double buffered
	};

	__device__ __forceinline__ uint32_t warp_redux_min(uint32_t v)
	{
		uint32_t r;
		asm volatile("redux.sync.min.u32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(v));
		return r;
	}

	// ---- peer-coupled stripes: a barrier across the ranks, on a cumulative arrive counter in rank 0's peer block ----
	__device__ __forceinline__ uint32_t ld_sys(const uint32_t *p)
	{
		uint32_t v;
		asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
		return v;
	}

	__device__ __forceinline__ unsigned long long global_timer_ns()
	{
		unsigned long long t;
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
		return t;
	}

	__device__ __forceinline__ void st_sys(uint32_t *p, uint32_t v)
	{
		asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
	}

	__device__ __forceinline__ void st_sys64(unsigned long long *p, unsigned long long v)
	{
		asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
	}

	__device__ __forceinline__ unsigned long long ld_sys64(const unsigned long long *p)
	{
		unsigned long long v;
		asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
		return v;
	}

	// Phase A, one thread per rank: tell every rank that all I wrote into its block has landed, wait until every rank
	// has told me. false: a rank did not get here within `limit_ns` (it died or never launched).
	__device__ inline bool peer_arrive_wait(const BigParams &P, uint32_t &epoch, bool &dead, unsigned long long limit_ns = 4000000000ull)
	{
		if (dead)
			return false; // one timeout is final: do not wait again
		__threadfence_system();
		epoch++;
		for (int j = 0; j < P.world; j++)
			st_sys(P.peers[j] + PEER_SLOT_A + P.rank, epoch);
		const unsigned long long t0 = global_timer_ns();
		for (int j = 0; j < P.world; j++)
			while ((int32_t)(ld_sys(P.ctl + PEER_SLOT_A + j) - epoch) < 0)
				if (global_timer_ns() - t0 > limit_ns)
				{
					dead = true;
					return false;
				}
		__threadfence_system();
		return true;
	}

	// Phase B: every rank learns the total queue length and the smallest key of all ranks
	__device__ inline bool peer_exchange(const BigParams &P, uint32_t &epoch, bool &dead, uint32_t count, uint32_t minkey, uint32_t &total, uint32_t &gmin)
	{
		total = 0u;
		gmin = MTD_INF;
		if (dead)
			return false;
		epoch++;
		const unsigned long long tag = epoch & 0xFFull;
		const unsigned long long v = ((unsigned long long)minkey << 32) | ((unsigned long long)(count < 0xFFFFFFu ? count : 0xFFFFFFu) << 8) | tag;
		for (int j = 0; j < P.world; j++)
			st_sys64(reinterpret_cast<unsigned long long *>(P.peers[j] + PEER_SLOT_B) + P.rank, v);
		const unsigned long long t0 = global_timer_ns();
		for (int j = 0; j < P.world; j++)
		{
			unsigned long long w;
			while (((w = ld_sys64(reinterpret_cast<const unsigned long long *>(P.ctl + PEER_SLOT_B) + j)) & 0xFFull) != tag)
				if (global_timer_ns() - t0 > 4000000000ull)
				{
					dead = true;
					return false;
				}
			total += (uint32_t)((w >> 8) & 0xFFFFFFull);
			gmin = min(gmin, (uint32_t)(w >> 32));
		}
		return true;
	}

	// Runs rounds `round0`, `round0 + 1`, ... until the work-list is empty or `max_rounds` rounds are done. COOPERATIVE launch:
	// every CTA is resident and the rounds are separated by grid.sync(). out[0] = items queued for the next round,
	// out[1] = index of the next round.
	template<int TS, int WARPS>
	__global__ void __launch_bounds__(WARPS * 32) k_tile_rounds(const BigParams P, int round0, int max_rounds, uint32_t *__restrict__ out)
	{
		static_assert(TS == 32, "one lane per border cell: the tile side is the warp width");
		extern __shared__ __align__(16) unsigned char smem_raw[];
		cooperative_groups::grid_group grid = cooperative_groups::this_grid();
		using WT = WarpFrame<TS>;
		constexpr int RW = WT::RW, PITCH = WT::PITCH, QCAP = WT::QCAP;
		const int warp = threadIdx.x >> 5;
		const uint32_t lane = lane_id();
		const uint32_t lt = lanemask_lt();
		WT &S = ((WT *)smem_raw)[warp];
		const int W = P.W;
		const int Hl = P.y1 - P.y0;
		const uint32_t ntiles = (uint32_t)(P.tilesX * P.tilesY);
		bool guard = false;
		int round = round0;
		uint32_t pending = 0;
#ifdef MTD_TIMING
		unsigned long long tacc[16] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 }; // developer probe: cycles per phase, buckets, queue entries, longest visit, visits over 60k / 100k cycles
#endif

		// Peer-coupled stripes: between two rounds every rank (1) tells all ranks that its boundary rows and keys have landed
		// and waits for the same from all of them, (2) queues the border tiles its neighbours offered a label to, (3) stores
		// its queue length and smallest key into every rank's block and reads everybody's: all ranks see the same wavefront
		// position and stop in the same round.
		uint32_t grid_gen = 0;
		auto peer_grid_sync = [&]()
		{
			if (!P.soft_grid)
			{
				grid.sync();
				return;
			}
			__syncthreads();
			if (threadIdx.x == 0)
			{
				uint32_t *bar = P.count + 11;
				__threadfence();
				grid_gen++;
				atomicAdd(bar, 1u);
				const uint32_t target = grid_gen * gridDim.x;
				while ((int32_t)(*(volatile uint32_t *)bar - target) < 0)
					;
				__threadfence();
			}
			__syncthreads();
		};
		uint32_t epoch = 0;
		bool peer_dead = false;
		auto peer_round_sync = [&](int r_next)
		{
			__threadfence_system();
			peer_grid_sync();
			if (blockIdx.x == 0)
			{
				const int slot = r_next % 3;
				bool ok = true;
				if (threadIdx.x == 0)
					ok = peer_arrive_wait(P, epoch, peer_dead);
				__syncthreads();
				for (int tx = (int)threadIdx.x; tx < P.tilesX; tx += (int)blockDim.x)
				{
					if (P.act_up)
					{
						const uint32_t k = atomicExch(P.act_up + tx, MTD_INF);
						if (k != MTD_INF)
							activate_item(P, 0, tx, 0, r_next - 1, k);
					}
					if (P.act_down)
					{
						const uint32_t k = atomicExch(P.act_down + tx, MTD_INF);
						if (k != MTD_INF)
							activate_item(P, 0, tx, P.tilesY - 1, r_next - 1, k);
					}
				}
				__threadfence();
				__syncthreads();
				if (threadIdx.x == 0)
				{
					const uint32_t cnt = __ldcg(P.count + slot), mk = __ldcg(P.minkey + slot);
					uint32_t gp, gm;
					ok = peer_exchange(P, epoch, peer_dead, cnt, mk, gp, gm) && ok;
					P.minkey[slot] = gm;
					P.ctl[PEER_GPENDING] = ok ? gp : 0u;
					P.ctl[PEER_EPOCH] = epoch;
					if (!ok)
						atomicOr(P.status, 8); // ST_PEER_TIMEOUT
				}
			}
			__threadfence();
			peer_grid_sync();
		};
		if (P.ctl0)
		{
			epoch = P.ctl[PEER_EPOCH];
			if (blockIdx.x == 0)
			{
				// every rank has entered this run (so nobody still reads the halo copies of the previous one): hand my
				// boundary rows as initialised (m, parked blocked tiles and sources 0) to the neighbours
				if (threadIdx.x == 0 && !peer_arrive_wait(P, epoch, peer_dead, P.soft_grid ? 4000000000ull : 20000000000ull)) // the ranks' hosts may reach the launch seconds apart
					atomicOr(P.status, 8);
				__syncthreads();
				for (int x = (int)threadIdx.x; x < W; x += (int)blockDim.x)
				{
					if (P.peer_up_halo)
						P.peer_up_halo[x] = __ldcg(P.dist + x);
					if (P.peer_down_halo)
						P.peer_down_halo[x] = __ldcg(P.dist + (size_t)(Hl - 1) * W + x);
				}
				__threadfence_system();
				__syncthreads();
			}
			peer_round_sync(round);
			if (__ldcg(P.ctl + PEER_GPENDING) == 0u)
				max_rounds = 0;
		}

		for (int done = 0; done < max_rounds; done++)
		{
		const int slot = round % 3;
		const uint32_t *list = P.list[slot];
		const uint32_t count = __ldcg(P.count + slot);
		uint32_t *cursor = P.cursor + slot;
		if (blockIdx.x == 0 && threadIdx.x == 0)
		{
			// the list consumed in the previous round is free: nobody reads or appends to it during this round
			P.count[(round + 2) % 3] = 0u;
			P.cursor[(round + 2) % 3] = 0u;
			P.minkey[(round + 2) % 3] = MTD_INF;
		}
		// Tile-level priority: only items whose best offer is within tile_delta of the smallest offer on the list are
		// visited now; the others are re-queued. Visiting far-away tiles early floods them with labels that the wavefront
		// proper undercuts later (measured without it: 6.7 visits per tile on 16384^2 / 16 sources).
		const uint32_t kmin = __ldcg(P.minkey + slot);
		const uint32_t kthr = kmin > MTD_INF - P.tile_delta ? MTD_INF : kmin + P.tile_delta;
		for (;;)
		{
			__syncwarp();
			uint32_t item = 0;
			if (lane == 0)
				item = atomicAdd(cursor, 1u);
			item = __shfl_sync(FULL_MASK, item, 0);
			if (item >= count)
				break;
			const uint32_t it = __ldcg(list + item);
			const int layer = (int)(it / ntiles);
			const uint32_t tile = it - (uint32_t)layer * ntiles;
			const int tx = (int)(tile % (uint32_t)P.tilesX), ty = (int)(tile / (uint32_t)P.tilesX);
			{
				// what is on offer to this item? nothing (a stale entry: the offer was consumed by an earlier visit) -> drop;
				// too far ahead of the wavefront -> wait; else consume the offer (a concurrent activation that is wiped here
				// wrote its labels before it raised the key, so the staging below sees them)
				uint32_t k = 0;
				if (lane == 0)
				{
					k = __ldcg(P.key + it);
					if (k != MTD_INF && k >= kthr)
						activate_item(P, layer, tx, ty, round, k);
					else if (k != MTD_INF)
					{
						atomicExch(P.key + it, MTD_INF); // reads the latest offer: whoever raised it wrote its labels first
						__threadfence();
					}
				}
				k = __shfl_sync(FULL_MASK, k, 0);
#ifdef MTD_TIMING
				tacc[12] += k == MTD_INF ? 1ull : 0ull;                 // stale entries dropped
				tacc[13] += (k != MTD_INF && k >= kthr) ? 1ull : 0ull;  // entries deferred by the priority window
#endif
				if (k == MTD_INF || k >= kthr)
					continue;
			}
			const int gx0 = tx * TS - 1; // global x of frame column 0
			const int ly0 = ty * TS - 1; // local row of frame row 0
			uint32_t *ldist = P.dist + (size_t)layer * P.layer_stride;
			const uint32_t *lsrc = P.sources + (size_t)layer * P.nsrc;
			const int brow = min(TS + 1, Hl - ly0); // frame row just below the tile's last row that exists in this stripe
			const bool interior = P.aligned16 && ly0 >= 0 && ly0 + RW <= Hl && gx0 >= 0 && gx0 + RW <= W;
#ifdef MTD_TIMING
			long long last__ = clock64();
			const long long visit0__ = last__;
			unsigned long long vsteps__ = 0, vstage__ = 0;
#define MTD_WTICK(k)                          \
	do                                        \
	{                                         \
		const long long now__ = clock64();    \
		tacc[k] += (unsigned long long)(now__ - last__); \
		last__ = now__;                       \
	} while (0)
#else
#define MTD_WTICK(k)
#endif

			// ---- stage the frame: labels + packed weights; cells outside the grid are parked at 0 ----
			if (interior)
			{
				// 34 rows x (8 + 8) sixteen-byte copies (labels, weights) + 34 x 2 x 2 apron words, spread over the lanes
				// the two apron columns are 4-byte accesses a row apart: all of a lane's loads are issued before anything waits
				// for one (L2 loads: the kernel lives across rounds, an L1 line would go stale), the row copies fly meanwhile
				constexpr int APN = (RW * 4 + 31) / 32;
				uint32_t apv[APN];
#pragma unroll
				for (int j = 0; j < APN; j++)
				{
					const int k = (int)lane + 32 * j;
					const int ry = k >> 2, sel = k & 3; // sel bit0: right column, bit1: weights
					const int rx = (sel & 1) ? RW - 1 : 0;
					apv[j] = 0u;
					if (k < RW * 4)
						apv[j] = __ldcg(((sel & 2) ? P.wpack : ldist) + (size_t)(ly0 + ry) * W + gx0 + rx);
				}
				for (int k = (int)lane; k < RW * 16; k += 32)
				{
					const int ry = k >> 4, ch = k & 15;
					const uint32_t *src = (ch < 8 ? ldist : P.wpack) + (size_t)(ly0 + ry) * W + gx0 + 1 + (ch & 7) * 4;
					uint32_t *dst = (ch < 8 ? S.d : S.w) + ry * PITCH + 4 + (ch & 7) * 4;
					cp_async_16(dst, src);
				}
#pragma unroll
				for (int j = 0; j < APN; j++)
				{
					const int k = (int)lane + 32 * j;
					const int ry = k >> 2, sel = k & 3;
					const int rx = (sel & 1) ? RW - 1 : 0;
					if (k < RW * 4)
						(((sel & 2) ? S.w : S.d) + ry * PITCH + 3 + rx)[0] = apv[j];
				}
				cp_async_wait_all();
			}
			else
			{
				for (int k = (int)lane; k < RW * RW; k += 32)
				{
					const int ry = k / RW, rx = k - ry * RW;
					const int ly = ly0 + ry, gx = gx0 + rx;
					uint32_t dv = 0u, wv = 0u;
					if (gx >= 0 && gx < W)
					{
						if (ly >= 0 && ly < Hl)
						{
							dv = __ldcg(ldist + (size_t)ly * W + gx);
							wv = P.wpack[(size_t)ly * W + gx];
						}
						else if (ly == -1 && P.up_row)
						{
							dv = __ldcg(P.up_row + gx);
							wv = P.up_wpack[gx];
						}
						else if (ly == Hl && P.down_row)
							dv = __ldcg(P.down_row + gx); // weights of the row below are never needed (edges are stored on the upper / left cell)
					}
					S.d[ry * PITCH + 3 + rx] = dv;
					S.w[ry * PITCH + 3 + rx] = wv;
				}
			}
			__syncwarp();
			MTD_WTICK(0);
#ifdef MTD_TIMING
			vstage__ = (unsigned long long)(last__ - visit0__);
#endif

			// ---- seeds: pending cells enter q[0]; every lowered label is appended (duplicates are harmless) ----
			uint32_t tail = 0;            // entries in the queue being filled (warp-uniform)
			uint32_t my_min = MTD_INF;    // smallest label this lane appended
			uint16_t *wq = S.q[0];
			auto relax_into = [&](bool want, uint32_t v, uint32_t nd)
			{
				bool improved = false;
				if (want && nd >= BIG_SETTLE_LIMIT)
					guard = true; // the reference's uint32 arithmetic is about to wrap: report, do not expand
				else if (want)
					improved = nd < atomicMin(&S.d[v], nd);
				const uint32_t bal = __ballot_sync(FULL_MASK, improved);
				if (improved)
				{
					const uint32_t slot = tail + __popc(bal & lt);
					if (slot < (uint32_t)QCAP)
						wq[slot] = (uint16_t)v;
					my_min = min(my_min, nd);
				}
				tail += __popc(bal);
			};
			// (1) apron: every finite, non-parked apron label relaxes its single interior neighbour; a label of 0 is a parked
			//     cell unless the source list says otherwise. Lane i owns cell i of each side; what the neighbour tile held is
			//     kept in registers to decide at the end whether it needs a visit.
			uint32_t ap_old[4];
#pragma unroll
			for (int side = 0; side < 4; side++)
			{
				const int i = (int)lane;
				uint32_t a, v, wsel;
				if (side == 0)
				{
					a = (uint32_t)(0 * PITCH + 4 + i); // top
					v = a + PITCH;
					wsel = S.w[a] >> 16;
				}
				else if (side == 1)
				{
					a = (uint32_t)(brow * PITCH + 4 + i); // bottom (the stripe's halo row when the tile is cut short)
					v = a - PITCH;
					wsel = S.w[v] >> 16;
				}
				else if (side == 2)
				{
					a = (uint32_t)((1 + i) * PITCH + 3); // left
					v = a + 1u;
					wsel = S.w[a] & 0xFFFFu;
				}
				else
				{
					a = (uint32_t)((1 + i) * PITCH + 4 + TS); // right
					v = a - 1u;
					wsel = S.w[v] & 0xFFFFu;
				}
				const uint32_t da = S.d[a];
				ap_old[side] = da;
				const bool row_exists = side < 2 || 1 + i < brow; // rows at and below brow of a cut-short tile hold the halo / nothing
				relax_into(row_exists && da != 0u && da != MTD_INF, v, da + MIN_EDGE + wsel);
			}
			uint32_t src_sides = 0; // sides on which a source of this tile touches the border
			for (int s0 = 0; s0 < P.nsrc; s0 += 32)
			{
				const int s = s0 + (int)lane;
				bool want = false;
				uint32_t v = 0, nd = 0;
				if (s < P.nsrc)
				{
					const uint32_t src = lsrc[s];
					const int rx = (int)(src % (uint32_t)W) - gx0, ry = (int)(src / (uint32_t)W) - P.y0 - ly0;
					const bool inx = rx >= 1 && rx <= TS, iny = ry >= 1 && ry < brow;
					const uint32_t a = (uint32_t)(ry * PITCH + rx + 3);
					if (ry == 0 && inx)
					{
						want = true;
						v = a + PITCH;
						nd = MIN_EDGE + (S.w[a] >> 16);
					}
					else if (ry == brow && inx)
					{
						want = true;
						v = a - PITCH;
						nd = MIN_EDGE + (S.w[v] >> 16);
					}
					else if (rx == 0 && iny)
					{
						want = true;
						v = a + 1u;
						nd = MIN_EDGE + (S.w[a] & 0xFFFFu);
					}
					else if (rx == TS + 1 && iny)
					{
						want = true;
						v = a - 1u;
						nd = MIN_EDGE + (S.w[v] & 0xFFFFu);
					}
				}
				relax_into(want, v, nd);
			}
			__syncwarp();
			// (2) park the apron: nothing is relaxed outward, and interior cells need no bounds test
			{
				const int i = (int)lane;
				S.d[0 * PITCH + 4 + i] = 0u;
				S.d[brow * PITCH + 4 + i] = 0u;
				S.d[(1 + i) * PITCH + 3] = 0u;
				S.d[(1 + i) * PITCH + 4 + TS] = 0u;
			}
			__syncwarp();
			// (3) sources inside the tile expand from label 0 (also when they are blocked: paths.cpp:63-67 test only the neighbour)
			for (int s0 = 0; s0 < P.nsrc; s0 += 32)
			{
				const int s = s0 + (int)lane;
				bool want = false;
				uint32_t c = (uint32_t)(PITCH + 4);
				if (s < P.nsrc)
				{
					const uint32_t src = lsrc[s];
					const int rx = (int)(src % (uint32_t)W) - gx0, ry = (int)(src / (uint32_t)W) - P.y0 - ly0;
					want = rx >= 1 && rx <= TS && ry >= 1 && ry < brow;
					if (want)
					{
						c = (uint32_t)(ry * PITCH + rx + 3);
						// a source on a border line reaches into the neighbour tile: that tile finds it through the source list
						src_sides |= (ry == 1 ? 4u : 0u) | (ry == TS ? 8u : 0u) | (rx == 1 ? 16u : 0u) | (rx == TS ? 32u : 0u);
					}
				}
				relax_into(want, c - PITCH, MIN_EDGE + (S.w[c - PITCH] >> 16));
				relax_into(want, c - 1u, MIN_EDGE + (S.w[c - 1u] & 0xFFFFu));
				relax_into(want, c + 1u, MIN_EDGE + (S.w[c] & 0xFFFFu));
				relax_into(want, c + PITCH, MIN_EDGE + (S.w[c] >> 16));
			}
			uint32_t dmin = warp_redux_min(my_min);
			__syncwarp();
			MTD_WTICK(1);

			// ---- bucketed wavefront on the frame, warp-synchronous (label-correcting across visits and inside wide buckets) ----
			uint32_t nq = tail, par = 0;
			bool overflow = false;
			while (nq != 0u && dmin != MTD_INF)
			{
				if (nq > (uint32_t)QCAP)
				{
					overflow = true;
					break;
				}
				const uint32_t thr = dmin + BIG_DELTA;
				const uint16_t *rq = S.q[par];
				wq = S.q[par ^ 1u];
				tail = 0;
#ifdef MTD_TIMING
				tacc[5] += 1ull;
				tacc[6] += nq;
				vsteps__ += 1ull;
#endif
				uint32_t newmin = MTD_INF;
				for (uint32_t base = 0; base < nq; base += 32u)
				{
					const uint32_t idx = base + lane;
					const bool active = idx < nq;
					const uint32_t c = active ? rq[idx] : (uint32_t)(PITCH + 4);
					const uint32_t d = active ? S.d[c] : MTD_INF;
					const bool settle = d < thr && d != 0u; // 0 = parked: never expands (sources expanded in (3))
					const bool keep = active && d >= thr;
					uint32_t n0 = 0, n1 = 0, n2 = 0, n3 = 0;
					bool i0 = false, i1 = false, i2 = false, i3 = false;
					if (settle)
					{
						// 0: up, 1: left, 2: right, 3: down -- weights are stored on the upper / left cell of an edge
						const uint32_t wc = S.w[c], base_nd = d + MIN_EDGE;
						n0 = base_nd + (S.w[c - PITCH] >> 16);
						n1 = base_nd + (S.w[c - 1u] & 0xFFFFu);
						n2 = base_nd + (wc & 0xFFFFu);
						n3 = base_nd + (wc >> 16);
						if (base_nd >= BIG_SETTLE_LIMIT)
							guard = true;
						else
						{
							i0 = n0 < atomicMin(&S.d[c - PITCH], n0);
							i1 = n1 < atomicMin(&S.d[c - 1u], n1);
							i2 = n2 < atomicMin(&S.d[c + 1u], n2);
							i3 = n3 < atomicMin(&S.d[c + PITCH], n3);
						}
					}
					uint32_t lm = keep ? d : MTD_INF;
					lm = i0 ? min(lm, n0) : lm;
					lm = i1 ? min(lm, n1) : lm;
					lm = i2 ? min(lm, n2) : lm;
					lm = i3 ? min(lm, n3) : lm;
					newmin = min(newmin, lm);
					const uint32_t bk = __ballot_sync(FULL_MASK, keep);
					const uint32_t b0 = __ballot_sync(FULL_MASK, i0);
					const uint32_t b1 = __ballot_sync(FULL_MASK, i1);
					const uint32_t b2 = __ballot_sync(FULL_MASK, i2);
					const uint32_t b3 = __ballot_sync(FULL_MASK, i3);
					uint32_t pos = tail;
					uint32_t p = pos + __popc(bk & lt);
					if (keep && p < (uint32_t)QCAP)
						wq[p] = (uint16_t)c;
					pos += __popc(bk);
					p = pos + __popc(b0 & lt);
					if (i0 && p < (uint32_t)QCAP)
						wq[p] = (uint16_t)(c - PITCH);
					pos += __popc(b0);
					p = pos + __popc(b1 & lt);
					if (i1 && p < (uint32_t)QCAP)
						wq[p] = (uint16_t)(c - 1u);
					pos += __popc(b1);
					p = pos + __popc(b2 & lt);
					if (i2 && p < (uint32_t)QCAP)
						wq[p] = (uint16_t)(c + 1u);
					pos += __popc(b2);
					p = pos + __popc(b3 & lt);
					if (i3 && p < (uint32_t)QCAP)
						wq[p] = (uint16_t)(c + PITCH);
					tail = pos + __popc(b3);
				}
				dmin = warp_redux_min(newmin);
				nq = tail;
				par ^= 1u;
				__syncwarp(); // queue entries and labels written by other lanes are read in the next bucket
			}
			if (overflow)
			{
				// the pending set outgrew the queue (pathological frontiers): queue-less sweeps to the local fixpoint.
				// Parked cells hold 0 and never expand; sources expanded in (3) before anything could overflow.
				for (;;)
				{
					bool changed = false;
					for (int i = (int)lane; i < TS * TS; i += 32)
					{
						const uint32_t c = (uint32_t)((1 + i / TS) * PITCH + 4 + i % TS);
						const uint32_t d = S.d[c];
						if (d == 0u || d == MTD_INF)
							continue;
						const uint32_t base_nd = d + MIN_EDGE;
						if (base_nd >= BIG_SETTLE_LIMIT)
						{
							guard = true;
							continue;
						}
						const uint32_t wc = S.w[c];
						const uint32_t n0 = base_nd + (S.w[c - PITCH] >> 16), n1 = base_nd + (S.w[c - 1u] & 0xFFFFu), n2 = base_nd + (wc & 0xFFFFu), n3 = base_nd + (wc >> 16);
						changed |= n0 < atomicMin(&S.d[c - PITCH], n0);
						changed |= n1 < atomicMin(&S.d[c - 1u], n1);
						changed |= n2 < atomicMin(&S.d[c + 1u], n2);
						changed |= n3 < atomicMin(&S.d[c + PITCH], n3);
					}
					__syncwarp();
					if (!__any_sync(FULL_MASK, changed))
						break;
				}
			}
			__syncwarp();
			MTD_WTICK(2);

			// ---- write back the interior (labels only ever decrease; parked and unreached cells rewrite their own value) ----
			if (interior)
			{
				for (int k = (int)lane; k < TS * 8; k += 32)
				{
					const int iy = k >> 3, ch = k & 7;
					const uint4 v = *reinterpret_cast<const uint4 *>(S.d + (1 + iy) * PITCH + 4 + ch * 4);
					__stcg(reinterpret_cast<uint4 *>(ldist + (size_t)(ly0 + 1 + iy) * W + gx0 + 1 + ch * 4), v);
				}
			}
			else
			{
				for (int k = (int)lane; k < TS * TS; k += 32)
				{
					const int iy = k / TS, ix = k - iy * TS;
					const int ly = ly0 + 1 + iy, gx = gx0 + 1 + ix;
					if (ly < Hl && gx < W)
						__stcg(ldist + (size_t)ly * W + gx, S.d[(1 + iy) * PITCH + 4 + ix]);
				}
			}

			if (P.ctl0)
			{
				// my boundary rows as they stand now, straight into the neighbour ranks' halo copies
				const int gx = gx0 + 1 + (int)lane;
				if (gx < W)
				{
					if (ty == 0 && P.peer_up_halo)
						P.peer_up_halo[gx] = S.d[1 * PITCH + 4 + (int)lane];
					if (ty == P.tilesY - 1 && P.peer_down_halo)
						P.peer_down_halo[gx] = S.d[(brow - 1) * PITCH + 4 + (int)lane];
				}
			}

			MTD_WTICK(3);
			// ---- activate neighbours ----
			// A neighbour tile needs a visit only if one of our border cells now offers its adjacent cell a smaller label than
			// that cell held when we staged it (labels only decrease, so a stale apron only over-activates; a parked or source
			// apron cell holds 0 and can never be improved). Our own parked cells (0) offer nothing.
			{
				auto offers = [&](uint32_t db, uint32_t w, uint32_t ap) -> bool { return db != 0u && db < BIG_SETTLE_LIMIT && db + MIN_EDGE + w < ap; };
				const int i = (int)lane;
				const uint32_t ct = (uint32_t)(1 * PITCH + 4 + i), cb = (uint32_t)((brow - 1) * PITCH + 4 + i); // first / last row of the tile that exists
				const uint32_t cl = (uint32_t)((1 + i) * PITCH + 4), cr = (uint32_t)((1 + i) * PITCH + 3 + TS);
				const bool row_exists = 1 + i < brow;
				auto offer = [&](bool exists, uint32_t db, uint32_t w, uint32_t ap) -> uint32_t { return exists && offers(db, w, ap) ? db + MIN_EDGE + w : MTD_INF; };
				uint32_t k_top = warp_redux_min(offer(true, S.d[ct], S.w[ct - PITCH] >> 16, ap_old[0]));
				uint32_t k_bot = warp_redux_min(offer(true, S.d[cb], S.w[cb] >> 16, ap_old[1])); // below a cut-short tile: the stripe's halo row or nothing (0)
				uint32_t k_left = warp_redux_min(offer(row_exists, S.d[cl], S.w[cl - 1u] & 0xFFFFu, ap_old[2]));
				uint32_t k_right = warp_redux_min(offer(row_exists, S.d[cr], S.w[cr] & 0xFFFFu, ap_old[3]));
				const uint32_t f = __reduce_or_sync(FULL_MASK, src_sides); // sources on a border line: the neighbour finds them in the list
				if (f & 4u)
					k_top = 0u;
				if (f & 8u)
					k_bot = 0u;
				if (f & 16u)
					k_left = 0u;
				if (f & 32u)
					k_right = 0u;
				if (P.ctl0 && (ty == 0 || ty == P.tilesY - 1))
					__threadfence_system(); // a border tile of the stripe: ... on the neighbour rank too
				else
					__threadfence(); // distances are visible before the neighbour can be scheduled
				__syncwarp();
				if (lane == 0)
				{
					if (P.ctl0)
					{
						if (k_top != MTD_INF && ty == 0 && P.peer_up_act)
							atomicMin_system(P.peer_up_act + tx, k_top);
						if (k_bot != MTD_INF && ty == P.tilesY - 1 && P.peer_down_act)
							atomicMin_system(P.peer_down_act + tx, k_bot);
					}
					atomicAdd(P.total_active, 1u);
				}
				// one lane per neighbour: the four queue insertions (each a chain of dependent atomics) run side by side
				if (lane == 0 && k_top != MTD_INF && ty > 0)
					activate_item(P, layer, tx, ty - 1, round, k_top);
				if (lane == 1 && k_bot != MTD_INF && ty < P.tilesY - 1)
					activate_item(P, layer, tx, ty + 1, round, k_bot);
				if (lane == 2 && k_left != MTD_INF)
					activate_item(P, layer, tx - 1, ty, round, k_left);
				if (lane == 3 && k_right != MTD_INF)
					activate_item(P, layer, tx + 1, ty, round, k_right);
			}
			MTD_WTICK(4);
#ifdef MTD_TIMING
			{
				const unsigned long long dv = (unsigned long long)(last__ - visit0__);
				tacc[7] = dv > tacc[7] ? dv : tacc[7];
				tacc[8] += dv > 60000ull ? 1ull : 0ull;
				tacc[9] += dv > 100000ull ? 1ull : 0ull;
				if (dv > 60000ull)
				{
					tacc[10] += vsteps__; // what the long visits are made of
					tacc[11] += vstage__;
				}
			}
#endif
		}
		// ---- end of the round: everything appended to the next list is visible after the barrier ----
		if (P.ctl0)
		{
			peer_round_sync(round + 1);
			round++;
			pending = __ldcg(P.ctl + PEER_GPENDING); // all ranks
			if (pending == 0u)
				break;
			continue;
		}
		__threadfence();
		grid.sync();
		round++;
		pending = __ldcg(P.count + round % 3);
		if (pending == 0u)
			break;
		}
		if (guard)
			atomicOr(P.status, 1);
#ifdef MTD_TIMING
		if (lane == 0)
			for (int k = 0; k < 16; k++)
				if (tacc[k])
				{
					if (k == 7)
						atomicMax(&g_big_timing[k], tacc[k]);
					else
						atomicAdd(&g_big_timing[k], tacc[k]);
				}
#endif
		if (blockIdx.x == 0 && threadIdx.x == 0)
		{
			out[0] = pending;
			out[1] = (uint32_t)round;
		}
	}
}
