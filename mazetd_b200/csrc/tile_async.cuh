// tile_async.cuh -- the large-grid relaxation (Directions::update, sources/map/paths.cpp:51-77) WITHOUT global rounds.
//
// k_tile_rounds (tile_warp.cuh) stops every warp of the GPU at a grid-wide barrier ~290 times per update of a 16384^2 grid,
// and a round lasts as long as its slowest warp: three quarters of a round were spent waiting (profiles/r01h_*). Here the
// (layer, 32x32 tile) work items flow through the machine as a dataflow:
//   * every item has ONE owner CTA (a fixed function of layer and tile position; tiles that are neighbours in the grid
//     have different owners, so a wavefront -- a curve of tiles -- spreads over all SMs);
//   * an offer to an item is `atomicMin(key[item], label)`; the offer that turns the key from "nothing pending" into a
//     label also appends the item to the owner's inbox ring in global memory (so an item is on at most one list at a
//     time and the ring, sized to the owner's items, cannot overflow);
//   * each CTA = 16 worker warps + 1 scheduler warp. The scheduler drains the inbox into a pending list in shared memory,
//     re-reads the pending keys (offers only ever lower them), publishes the CTA's smallest key, reads every CTA's
//     smallest key (148 words) to get the global wavefront position, and hands the pending items with the smallest keys
//     inside the priority window to idle workers through per-worker mailboxes in shared memory. Workers spin on shared
//     memory only. A tile is never given to two workers at once (write-backs of one tile must not interleave);
//   * termination: a global count of items that are pending or being visited; every increment (an offer that made an item
//     pending) happens before the decrement of the visit that issued it, so the count reaches zero exactly once.
// The labels converge to the unique fixpoint whatever the order of visits, so the result is bit-identical to the
// round-based kernel and to the reference. The priority window (visit only items whose key is within `tile_delta` of the
// smallest key anywhere) is what keeps warps from running ahead of the wavefront and multiplying revisits.
#pragma once

#include <cuda.h> // CUtensorMap (the driver entry point that encodes it is looked up at run time, nothing links libcuda)

#include "common.cuh"
#include "sssp_warp.cuh"
// included after big_grid.cuh and tile_warp.cuh: BigParams, WarpFrame, cp_async_*, warp_redux_min, global_timer_ns

namespace mtd
{
	constexpr int ASYNC_WORKERS = 16;     // worker warps per CTA (one CTA per SM)
#ifndef MTD_ASYNC_BUCKET
#define MTD_ASYNC_BUCKET 384u
#endif
	// bucket width of the engine inside a visit, in label units (edges weigh 255 .. ~600). While throughput bounded the kernel
	// 512 .. 4096 were within 2 % of each other (r02a); now that the chain of dependent visits does, narrower buckets -- fewer
	// cells settled too early and relaxed again -- make a visit shorter: 255 / 384 / 512 / 768 / 1024 / 2048 give 9.08 / 8.20 /
	// 8.28 / 8.52 / 8.63 / 8.78 ms on C5 and 12.62 / 11.71 / 11.79 / 12.01 / 12.11 / 12.27 ms on C4 (profiles/r04f_*)
	constexpr uint32_t ASYNC_BUCKET = MTD_ASYNC_BUCKET;
	constexpr int ASYNC_PCAP = 2048;      // pending items a scheduler can hold (>= items owned by a CTA, checked on the host)
	constexpr int ASYNC_KX = 12;          // owner = (tx + KX * ty + KL * layer) mod CTAs: a 12 x 12 block of tiles has 144 different owners
	constexpr int ASYNC_KL = 53;
	constexpr uint32_t SLOT_IDLE = 0xFFFFFFFFu, SLOT_EXIT = 0xFFFFFFFEu;
	constexpr uint32_t SLOT_CONT = 0x80000000u; // mailbox flag: the worker's frame still holds this item from its last visit (items are < 2^31)
	// control words (AsyncCtl::ctl)
	constexpr int ACTL_OUTSTANDING = 0; // items pending or in flight
	constexpr int ACTL_ABORT = 1;       // a scheduler hit the time limit
	constexpr int ACTL_VISITS = 2;      // statistics
	constexpr int ACTL_IDLE_POLLS = 3;
	// peer-coupled stripes
	constexpr int ACTL_STOP = 4; // set by CTA 0's scheduler when rank 0 announced the end of the run
	constexpr int ACTL_SENT = 5; // offers to a neighbour rank that made its border tile pending
	constexpr int ACTL_RECV = 6; // such offers from the neighbour ranks taken over by this rank's schedulers
	constexpr int ACTL_GO = 7;   // set by CTA 0's scheduler once every rank has entered the run and the initial halo rows have landed
	constexpr int ACTL_WORDS = 32;

	// A worker's frame: the WarpFrame layout (34 rows of 40 words, grid column gx0 + rx at word rx + 3) with both arrays on
	// 128-byte boundaries, so that ONE 2-D TMA box of 40 x 34 words per array (origin tile_x0 - 4, tile_y0 - 1) lands exactly
	// on it; cells outside the grid are zero-filled by the TMA unit, which is the "parked" label the engine expects.
	template<int TS>
	struct alignas(128) AsyncFrame
	{
		static constexpr int RW = TS + 2;
		static constexpr int PITCH = TS + 8;
		static constexpr int CELLS = RW * PITCH;
		static constexpr int QCAP = 512;
		uint32_t d[CELLS];
		uint64_t bar; // mbarrier the two bulk tensor loads of a visit complete on
		alignas(128) uint32_t w[CELLS];
		uint16_t q[2][QCAP];
	};

	// generic-proxy accesses (the atomics on the tile keys, stores of other SMs made visible by them, the previous visit's use
	// of the frame) are ordered before the bulk tensor copies that follow (async proxy), in global and in shared memory
	__device__ __forceinline__ void fence_proxy_async_all()
	{
		asm volatile("fence.proxy.async;" ::: "memory");
	}

	// release / acquire at GPU scope for the message passing between workers, schedulers and offers (__threadfence() is a
	// sequentially consistent fence plus an L1 invalidation: neither is needed, every shared word is read at L2)
	__device__ __forceinline__ void fence_gpu()
	{
		asm volatile("fence.acq_rel.gpu;" ::: "memory");
	}

	__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *tm, int x, int y, int z, uint64_t *bar)
	{
		asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst)), "l"(tm), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
					 : "memory");
	}

	__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *tm, int x, int y, uint64_t *bar)
	{
		asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst)), "l"(tm), "r"(x), "r"(y), "r"(smem_u32(bar))
					 : "memory");
	}

#ifdef MTD_TIMING
	// developer probe (never in the shipped build): cycles per phase summed over all workers
	// 0 wait for work, 1 consume key, 2 stage, 3 seed, 4 engine, 5 write back + offers computed, 6 offers sent + finish, 7 engine steps, 8 cells popped
	__device__ unsigned long long g_async_timing[16];
	__device__ unsigned long long g_async_sched_extra[4];
	__device__ unsigned long long g_async_sched_phase[8]; // cycles of a scheduler iteration: loads | refresh + drain | bookkeeping | hand-out | consume + ring | tail
#define MTD_STICK(k) do { if (lane == 0) { const long long n__ = clock64(); sacc__[k] += (unsigned long long)(n__ - slast__); slast__ = n__; } } while (0)
	// developer probe: length of the longest chain of dependent visits (a visit's depth = 1 + the largest depth among the
	// visits whose offers it was made pending by), and when the deepest visit ended
	__device__ uint32_t g_async_depth[1u << 20];
	__device__ uint32_t g_async_depth_max;
	__device__ uint32_t g_async_sdepth[1u << 20]; // ... counting the previous visit of the same tile as a predecessor too
	__device__ uint32_t g_async_sdepth_off[1u << 20];
	__device__ uint32_t g_async_sdepth_max;
#define MTD_ATICK(k)                                     \
	do                                                   \
	{                                                    \
		const long long now__ = clock64();               \
		tacc[k] += (unsigned long long)(now__ - last__); \
		last__ = now__;                                  \
	} while (0)
#define MTD_ATIMING_ARGS , unsigned long long *tacc, long long &last__
#define MTD_ATIMING_PASS , tacc, last__
#else
#define MTD_ATICK(k)
#define MTD_STICK(k)
#define MTD_ATIMING_ARGS
#define MTD_ATIMING_PASS
#endif

	// hint: bring the frame of the item this worker visits next into L2 while it is busy with the current one
	__device__ __forceinline__ void tma_prefetch_3d(const CUtensorMap *tm, int x, int y, int z)
	{
		asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(tm), "r"(x), "r"(y), "r"(z) : "memory");
	}

	__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap *tm, int x, int y)
	{
		asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(tm), "r"(x), "r"(y) : "memory");
	}

	struct AsyncParams
	{
		unsigned long long *inbox; // [ctas][cap] key << 32 | (item + 1), 0 = empty
		uint32_t *tail;      // [ctas * 32] (one 128-byte line per CTA) entries ever appended to the CTA's inbox
		uint32_t *ctamin;    // [ctas] smallest key among the CTA's pending and in-flight items (MTD_INF: none)
		uint32_t *ctl;       // ACTL_* words
		uint32_t cap_mask;   // cap - 1 (cap is a power of two >= items per owner)
		int ctas;
		int use_tma;         // frames are staged by two bulk tensor copies (needs W % 4 == 0); else by cp.async rows / scalar loads
		uint32_t *fly;       // [items] key an item is being visited at (MTD_INF: not in flight); read by the neighbours' schedulers
		int nb_rule;         // 1: an item does not start while a neighbour tile is in flight or pending with a smaller key
		uint32_t nb_margin;  // ... smaller by more than this
		int depth;           // mailboxes per worker in use: 1 or 2
		int workers;         // worker warps the scheduler hands items to (<= ASYNC_WORKERS; the rest sleep)
		int late_looks;      // how often a visit re-reads its apron from global memory after converging (0 = never)
		int cont;            // 1: a revisit goes to the worker that visited the item last and continues on the frame it left (no staging)
		unsigned long long time_limit_ns;
	};

	__device__ __forceinline__ int async_owner(const AsyncParams &A, int layer, int tx, int ty)
	{
		return (int)(((uint32_t)tx + (uint32_t)ASYNC_KX * (uint32_t)ty + (uint32_t)ASYNC_KL * (uint32_t)layer) % (uint32_t)A.ctas);
	}

	// Offer label `key` to item (layer, tx, ty); the caller has made its labels visible (__threadfence) before AND has
	// already counted the item in ctl[OUTSTANDING] (the count must never be behind). Returns true if the offer made the
	// item pending (it is now on its owner's inbox), false if it already was (the caller takes its count back).
	__device__ __forceinline__ bool async_offer(const BigParams &P, const AsyncParams &A, int layer, int tx, int ty, uint32_t key)
	{
		const uint32_t item = (uint32_t)layer * (uint32_t)(P.tilesX * P.tilesY) + (uint32_t)(ty * P.tilesX + tx);
		const uint32_t old = atomicMin(P.key + item, key);
		if (old != MTD_INF)
			return false;
		const int o = async_owner(A, layer, tx, ty);
		const uint32_t pos = atomicAdd(A.tail + o * 32, 1u);
		__stcg(A.inbox + (size_t)o * (A.cap_mask + 1u) + (pos & A.cap_mask), ((unsigned long long)key << 32) | (unsigned long long)(item + 1u));
		atomicMin(A.ctamin + o, key); // the owner's published wavefront position covers the offer before its scheduler has seen it
		return true;
	}

	// sources := 0, their items pending (one thread per (layer, source)); ctl[OUTSTANDING] counts them
	template<int TS>
	__global__ void k_async_seed(const BigParams P, const AsyncParams A)
	{
		const int s = blockIdx.x * blockDim.x + threadIdx.x;
		if (s >= P.nlayers * P.nsrc)
			return;
		const int layer = s / P.nsrc;
		const uint32_t src = P.sources[s];
		const int x = (int)(src % (uint32_t)P.W), y = (int)(src / (uint32_t)P.W);
		int ty;
		if (y >= P.y0 && y < P.y1)
		{
			P.dist[(size_t)layer * P.layer_stride + (size_t)(y - P.y0) * P.W + x] = 0u;
			ty = (y - P.y0) / TS;
		}
		else if (y == P.y0 - 1)
			ty = 0; // row stripes: a source on the neighbour's boundary row takes its first hop into my border tile (the visit finds it in the source list)
		else if (y == P.y1)
			ty = P.tilesY - 1;
		else
			return;
		atomicAdd(A.ctl + ACTL_OUTSTANDING, 1u);
		__threadfence();
		if (!async_offer(P, A, layer, x / TS, ty, 0u))
			atomicSub(A.ctl + ACTL_OUTSTANDING, 1u); // several sources in one tile: one item
	}

	// One visit of item (layer, tx, ty) by one warp: stage the 34 x 34 frame (labels + packed weights), seed from the apron
	// and the sources, run the warp-synchronous bucketed wavefront to local convergence, write the interior back.
	// Returns in k[4] (warp-uniform) the smallest label on offer to the tile above / below / left / right (MTD_INF: none).
	// Same arithmetic as the visit inside k_tile_rounds; `guard` is raised when a label gets close to the uint32 wrap.
	template<int TS, class WT>
	__device__ __forceinline__ void warp_visit(const BigParams &P, WT &S, int layer, int tx, int ty, uint32_t lane, uint32_t lt, uint32_t (&k)[4], bool &guard, const CUtensorMap *tm_dist, const CUtensorMap *tm_w, uint32_t &tma_phase, uint32_t next_item, int late_looks, bool cont MTD_ATIMING_ARGS)
	{
		constexpr int RW = WT::RW, PITCH = WT::PITCH, QCAP = WT::QCAP;
		const int W = P.W;
		const int Hl = P.y1 - P.y0;
		const int gx0 = tx * TS - 1; // global x of frame column 0
		const int ly0 = ty * TS - 1; // local row of frame row 0
		uint32_t *ldist = P.dist + (size_t)layer * P.layer_stride;
		const uint32_t *lsrc = P.sources + (size_t)layer * P.nsrc;
		const int brow = min(TS + 1, Hl - ly0); // frame row just below the tile's last existing row
		const bool use_tma = tm_dist != nullptr;
		// the whole frame lies inside the grid and rows move as 16-byte vectors (staging without TMA); the write-back needs
		// only the tile itself inside
		const bool interior = P.aligned16 && ly0 >= 0 && ly0 + RW <= Hl && gx0 >= 0 && gx0 + RW <= W;
		const bool interior_wb = P.aligned16 && ly0 + 1 + TS <= Hl && gx0 + 1 + TS <= W;

		// the layer's first 32 sources, requested before the frame so that the list's L2 round trip hides behind the staging
		// (read twice below: r03w showed 7k cycles in the seed phase, two dependent global loads of it)
		const uint32_t src0 = (int)lane < P.nsrc ? __ldg(lsrc + lane) : 0u;
		// ---- stage ---- (not when the frame is still the one this worker left after its last visit of this very item)
		if (!cont)
		{
		if (use_tma)
		{
			// two boxes of 40 x 34 words (labels of this layer, packed weights), origin (tile_x0 - 4, tile_y0 - 1); what lies
			// outside the grid arrives as 0 = parked
			if (lane == 0)
			{
				fence_proxy_async_all();
				mbar_expect_tx(&S.bar, 2u * (uint32_t)(RW * PITCH * 4));
				tma_load_3d(S.d, tm_dist, tx * TS - 4, ly0, layer, &S.bar);
				tma_load_2d(S.w, tm_w, tx * TS - 4, ly0, &S.bar);
				if (next_item < SLOT_EXIT)
				{
					// the item waiting in my second mailbox: its frames are on their way to L2 while this one is relaxed
					const uint32_t nt = (uint32_t)(P.tilesX * P.tilesY);
					const int nl = (int)(next_item / nt);
					const uint32_t ntile = next_item - (uint32_t)nl * nt;
					const int nx = (int)(ntile % (uint32_t)P.tilesX) * TS - 4, ny = (int)(ntile / (uint32_t)P.tilesX) * TS - 1;
					tma_prefetch_3d(tm_dist, nx, ny, nl);
					tma_prefetch_2d(tm_w, nx, ny);
				}
			}
			mbar_wait(&S.bar, tma_phase);
			tma_phase ^= 1u;
		}
		else if (interior)
		{
			constexpr int APN = (RW * 4 + 31) / 32;
			uint32_t apv[APN];
#pragma unroll
			for (int j = 0; j < APN; j++)
			{
				const int q = (int)lane + 32 * j;
				const int ry = q >> 2, sel = q & 3; // sel bit0: right column, bit1: weights
				const int rx = (sel & 1) ? RW - 1 : 0;
				apv[j] = 0u;
				if (q < RW * 4)
					apv[j] = __ldcg(((sel & 2) ? P.wpack : ldist) + (size_t)(ly0 + ry) * W + gx0 + rx);
			}
			for (int q = (int)lane; q < RW * 16; q += 32)
			{
				const int ry = q >> 4, ch = q & 15;
				const uint32_t *src = (ch < 8 ? ldist : P.wpack) + (size_t)(ly0 + ry) * W + gx0 + 1 + (ch & 7) * 4;
				uint32_t *dst = (ch < 8 ? S.d : S.w) + ry * PITCH + 4 + (ch & 7) * 4;
				cp_async_16(dst, src);
			}
#pragma unroll
			for (int j = 0; j < APN; j++)
			{
				const int q = (int)lane + 32 * j;
				const int ry = q >> 2, sel = q & 3;
				const int rx = (sel & 1) ? RW - 1 : 0;
				if (q < RW * 4)
					(((sel & 2) ? S.w : S.d) + ry * PITCH + 3 + rx)[0] = apv[j];
			}
			cp_async_wait_all();
		}
		else
		{
			for (int q = (int)lane; q < RW * RW; q += 32)
			{
				const int ry = q / RW, rx = q - ry * RW;
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
						dv = __ldcg(P.down_row + gx);
				}
				S.d[ry * PITCH + 3 + rx] = dv;
				S.w[ry * PITCH + 3 + rx] = wv;
			}
		}
		}
		__syncwarp();

		MTD_ATICK(2);
		// ---- seeds ----
		uint32_t tail = 0;
		uint32_t my_min = MTD_INF;
		uint16_t *wq = S.q[0];
		auto relax_into = [&](bool want, uint32_t v, uint32_t nd)
		{
			bool improved = false;
			if (want && nd >= BIG_SETTLE_LIMIT)
				guard = true;
			else if (want)
				improved = nd < atomicMin(&S.d[v], nd);
			const uint32_t bal = __ballot_sync(0xFFFFFFFFu, improved);
			if (improved)
			{
				const uint32_t slot = tail + __popc(bal & lt);
				if (slot < (uint32_t)QCAP)
					wq[slot] = (uint16_t)v;
				my_min = min(my_min, nd);
			}
			tail += __popc(bal);
		};
		uint32_t ap_old[4] = { 0u, 0u, 0u, 0u };
		uint32_t src_sides = 0;
		if (!cont)
		{
#pragma unroll
		for (int side = 0; side < 4; side++)
		{
			const int i = (int)lane;
			uint32_t a, v, wsel;
			if (side == 0)
			{
				a = (uint32_t)(0 * PITCH + 4 + i);
				v = a + PITCH;
				wsel = S.w[a] >> 16;
			}
			else if (side == 1)
			{
				a = (uint32_t)(brow * PITCH + 4 + i);
				v = a - PITCH;
				wsel = S.w[v] >> 16;
			}
			else if (side == 2)
			{
				a = (uint32_t)((1 + i) * PITCH + 3);
				v = a + 1u;
				wsel = S.w[a] & 0xFFFFu;
			}
			else
			{
				a = (uint32_t)((1 + i) * PITCH + 4 + TS);
				v = a - 1u;
				wsel = S.w[v] & 0xFFFFu;
			}
			const uint32_t da = S.d[a];
			ap_old[side] = da;
			const bool row_exists = side < 2 || 1 + i < brow;
			relax_into(row_exists && da != 0u && da != MTD_INF, v, da + MIN_EDGE + wsel);
		}
		// is any source inside this frame (tile or apron)? Most tiles have none: the two passes over the list below, with their
		// five ballot-and-store rounds per 32 sources, are skipped then
		bool src_near = false;
		for (int s0 = 0; s0 < P.nsrc; s0 += 32)
		{
			const int s = s0 + (int)lane;
			bool near = false;
			if (s < P.nsrc)
			{
				const uint32_t src = s0 == 0 ? src0 : lsrc[s];
				const int rx = (int)(src % (uint32_t)W) - gx0, ry = (int)(src / (uint32_t)W) - P.y0 - ly0;
				near = rx >= 0 && rx <= TS + 1 && ry >= 0 && ry <= brow;
			}
			src_near |= __any_sync(0xFFFFFFFFu, near);
		}
		for (int s0 = 0; src_near && s0 < P.nsrc; s0 += 32)
		{
			const int s = s0 + (int)lane;
			bool want = false;
			uint32_t v = 0, nd = 0;
			if (s < P.nsrc)
			{
				const uint32_t src = s0 == 0 ? src0 : lsrc[s];
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
		{
			const int i = (int)lane;
			S.d[0 * PITCH + 4 + i] = 0u;
			S.d[brow * PITCH + 4 + i] = 0u;
			S.d[(1 + i) * PITCH + 3] = 0u;
			S.d[(1 + i) * PITCH + 4 + TS] = 0u;
		}
		__syncwarp();
		for (int s0 = 0; src_near && s0 < P.nsrc; s0 += 32)
		{
			const int s = s0 + (int)lane;
			bool want = false;
			uint32_t c = (uint32_t)(PITCH + 4);
			if (s < P.nsrc)
			{
				const uint32_t src = s0 == 0 ? src0 : lsrc[s];
				const int rx = (int)(src % (uint32_t)W) - gx0, ry = (int)(src / (uint32_t)W) - P.y0 - ly0;
				want = rx >= 1 && rx <= TS && ry >= 1 && ry < brow;
				if (want)
				{
					c = (uint32_t)(ry * PITCH + rx + 3);
					src_sides |= (ry == 1 ? 4u : 0u) | (ry == TS ? 8u : 0u) | (rx == 1 ? 16u : 0u) | (rx == TS ? 32u : 0u);
				}
			}
			relax_into(want, c - PITCH, MIN_EDGE + (S.w[c - PITCH] >> 16));
			relax_into(want, c - 1u, MIN_EDGE + (S.w[c - 1u] & 0xFFFFu));
			relax_into(want, c + 1u, MIN_EDGE + (S.w[c] & 0xFFFFu));
			relax_into(want, c + PITCH, MIN_EDGE + (S.w[c] >> 16));
		}
		}
		uint32_t dmin = warp_redux_min(my_min);
		__syncwarp();

		MTD_ATICK(3);
		// ---- bucketed wavefront on the frame, warp-synchronous ----
		// A step settles the queued cells whose label is below `thr` and relaxes their four neighbours; improved cells and
		// cells still above the threshold go to the other queue. The body is branch-free (the ncu capture of the first
		// version, profiles/r02f_*, showed 2.1 cycles of branch-resolving stall per issued instruction): lanes with nothing
		// to settle run the same code with candidates of MTD_INF, positions come from three bit-plane ballots of the
		// per-lane output count instead of five ballots, and the threshold advances by one bucket per step (jumping ahead
		// only when a step settled nothing) instead of tracking the exact minimum.
		const uint32_t spare = (lane + 1u) * (uint32_t)PITCH + ((lane >> 2) % 6u < 3u ? (lane >> 2) % 6u : 34u + (lane >> 2) % 6u); // words 0..2 / 37..39 of frame row lane + 1
		uint32_t par = 0;
		auto engine = [&]()
		{
		uint32_t nq = tail;
		bool overflow = false;
		uint32_t thr = dmin == MTD_INF ? MTD_INF : dmin + ASYNC_BUCKET;
		while (nq != 0u)
		{
			if (nq > (uint32_t)QCAP)
			{
				overflow = true;
				break;
			}
			const uint16_t *rq = S.q[par];
			wq = S.q[par ^ 1u];
			tail = 0;
#ifdef MTD_TIMING
			tacc[7] += 1ull;
			tacc[8] += nq;
#endif
			uint32_t kept_min = MTD_INF;
			bool settled_any = false;
			for (uint32_t base = 0; base < nq; base += 32u)
			{
				const uint32_t idx = base + lane;
				const bool active = idx < nq;
				const uint32_t c = rq[active ? idx : 0u];
				const uint32_t d = S.d[c];
				const bool settle = active && d < thr && d != 0u;
				const bool keep = active && d >= thr;
				const uint32_t wc = S.w[c], wu = S.w[c - PITCH] >> 16, wl = S.w[c - 1u] & 0xFFFFu;
				const uint32_t base_nd = d + MIN_EDGE;
				const bool ok = settle && base_nd < BIG_SETTLE_LIMIT;
				guard |= settle && !ok;
				settled_any |= settle;
				const uint32_t n0 = base_nd + wu, n1 = base_nd + wl, n2 = base_nd + (wc & 0xFFFFu), n3 = base_nd + (wc >> 16);
				// Four UNCONDITIONAL shared-memory atomics: a lane with nothing to settle aims them at its own spare word of the
				// frame (the three words left and right of the apron in every row are never read) with MTD_INF, which changes
				// nothing. `if (ok) atomicMin(...)` -- also as predicated inline PTX -- compiles to a divergent branch per
				// atomic, and branch resolving was the largest single stall reason in profiles/r02f_*. (Reading the four labels
				// first and skipping atomics that cannot improve was measured again in r02i: 812 -> 998 cycles per step.)
				const uint32_t sel = ok ? 0xFFFFFFFFu : 0u;
				const uint32_t a0 = ((c - PITCH) & sel) | (spare & ~sel), a1 = ((c - 1u) & sel) | (spare & ~sel), a2 = ((c + 1u) & sel) | (spare & ~sel), a3 = ((c + PITCH) & sel) | (spare & ~sel);
				const uint32_t o0 = atomicMin(&S.d[a0], n0 | ~sel), o1 = atomicMin(&S.d[a1], n1 | ~sel);
				const uint32_t o2 = atomicMin(&S.d[a2], n2 | ~sel), o3 = atomicMin(&S.d[a3], n3 | ~sel);
				const bool i0 = ok && n0 < o0, i1 = ok && n1 < o1, i2 = ok && n2 < o2, i3 = ok && n3 < o3;
				kept_min = min(kept_min, keep ? d : MTD_INF);
				// every lane has up to five entries for the other queue; exclusive prefix of the counts from three ballots
				const uint32_t cnt = (keep ? 1u : 0u) + (i0 ? 1u : 0u) + (i1 ? 1u : 0u) + (i2 ? 1u : 0u) + (i3 ? 1u : 0u);
				const uint32_t p0 = __ballot_sync(0xFFFFFFFFu, cnt & 1u), p1 = __ballot_sync(0xFFFFFFFFu, cnt & 2u), p2 = __ballot_sync(0xFFFFFFFFu, cnt & 4u);
				uint32_t p = tail + (uint32_t)__popc(p0 & lt) + 2u * (uint32_t)__popc(p1 & lt) + 4u * (uint32_t)__popc(p2 & lt);
				tail += (uint32_t)__popc(p0) + 2u * (uint32_t)__popc(p1) + 4u * (uint32_t)__popc(p2);
				if (keep && p < (uint32_t)QCAP)
					wq[p] = (uint16_t)c;
				p += keep ? 1u : 0u;
				if (i0 && p < (uint32_t)QCAP)
					wq[p] = (uint16_t)(c - PITCH);
				p += i0 ? 1u : 0u;
				if (i1 && p < (uint32_t)QCAP)
					wq[p] = (uint16_t)(c - 1u);
				p += i1 ? 1u : 0u;
				if (i2 && p < (uint32_t)QCAP)
					wq[p] = (uint16_t)(c + 1u);
				p += i2 ? 1u : 0u;
				if (i3 && p < (uint32_t)QCAP)
					wq[p] = (uint16_t)(c + PITCH);
			}
			if (__any_sync(0xFFFFFFFFu, settled_any))
				thr = thr > MTD_INF - ASYNC_BUCKET ? MTD_INF : thr + ASYNC_BUCKET;
			else
			{
				// nothing was below the threshold: jump to the bucket of the smallest pending label
				const uint32_t km = warp_redux_min(kept_min);
				thr = km > MTD_INF - ASYNC_BUCKET ? MTD_INF : km + ASYNC_BUCKET;
			}
			nq = tail;
			par ^= 1u;
			__syncwarp();
		}
		if (overflow)
		{
			// the pending set outgrew the queue: queue-less sweeps to the local fixpoint (parked cells hold 0 and never expand)
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
				if (!__any_sync(0xFFFFFFFFu, changed))
					break;
			}
		}
		__syncwarp();
		};
		// ---- a late look at the apron ----
		// Offers from several neighbours reach a tile within one visit time (the wavefront passes); whatever arrived while
		// this visit was running would otherwise wait for it to end and cost a second full visit -- on the chain of dependent
		// visits that bounds the update that is a wait plus a visit per tile hop. The neighbours write
		// their interiors back before they offer, so their border cells are re-read from global memory here (one L2 round
		// trip), and cells that dropped below what this visit was seeded with seed the engine once more. The pending item
		// such an offer created is still visited later (the accounting stays as it is) and finds nothing to do.
		auto apron_look = [&](bool all)
		{
			wq = S.q[par];
			tail = 0;
			my_min = MTD_INF;
			const int i = (int)lane;
			const int gxi = gx0 + 1 + i, lyi = ly0 + 1 + i;
			uint32_t an[4] = { 0u, 0u, 0u, 0u };
			if (gxi < W)
			{
				if (ly0 >= 0)
					an[0] = __ldcg(ldist + (size_t)ly0 * W + gxi);
				else if (ly0 == -1 && P.up_row)
					an[0] = __ldcg(P.up_row + gxi);
				const int lyb = ly0 + brow;
				if (lyb < Hl)
					an[1] = __ldcg(ldist + (size_t)lyb * W + gxi);
				else if (lyb == Hl && P.down_row)
					an[1] = __ldcg(P.down_row + gxi);
			}
			if (1 + i < brow)
			{
				if (gx0 >= 0)
					an[2] = __ldcg(ldist + (size_t)lyi * W + gx0);
				if (gx0 + TS + 1 < W)
					an[3] = __ldcg(ldist + (size_t)lyi * W + gx0 + TS + 1);
			}
#pragma unroll
			for (int side = 0; side < 4; side++)
			{
				uint32_t a, v, wsel;
				if (side == 0)
				{
					a = (uint32_t)(0 * PITCH + 4 + i);
					v = a + PITCH;
					wsel = S.w[a] >> 16;
				}
				else if (side == 1)
				{
					a = (uint32_t)(brow * PITCH + 4 + i);
					v = a - PITCH;
					wsel = S.w[v] >> 16;
				}
				else if (side == 2)
				{
					a = (uint32_t)((1 + i) * PITCH + 3);
					v = a + 1u;
					wsel = S.w[a] & 0xFFFFu;
				}
				else
				{
					a = (uint32_t)((1 + i) * PITCH + 4 + TS);
					v = a - 1u;
					wsel = S.w[v] & 0xFFFFu;
				}
				const uint32_t da = an[side];
				const bool row_exists = side < 2 || 1 + i < brow;
				const bool lower = row_exists && da != 0u && da != MTD_INF && (all || da < ap_old[side]); // a parked cell (0) stays parked: blocked tiles never change
				relax_into(lower, v, da + MIN_EDGE + wsel);
				if (lower || all)
					ap_old[side] = da;
			}
			dmin = warp_redux_min(my_min);
			__syncwarp();
		};
		// continuing on the frame of my last visit: everything the neighbours hold now is a candidate (their offers are why I am here)
		if (cont)
			apron_look(true);
		engine();
		for (int look = 0; look < late_looks; look++)
		{
			apron_look(false);
			if (tail == 0u)
				break;
			engine();
		}

		MTD_ATICK(4);
		// ---- write back the interior ----
		if (interior_wb)
		{
			for (int q = (int)lane; q < TS * 8; q += 32)
			{
				const int iy = q >> 3, ch = q & 7;
				const uint4 v = *reinterpret_cast<const uint4 *>(S.d + (1 + iy) * PITCH + 4 + ch * 4);
				__stcg(reinterpret_cast<uint4 *>(ldist + (size_t)(ly0 + 1 + iy) * W + gx0 + 1 + ch * 4), v);
			}
		}
		else
		{
			for (int q = (int)lane; q < TS * TS; q += 32)
			{
				const int iy = q / TS, ix = q - iy * TS;
				const int ly = ly0 + 1 + iy, gx = gx0 + 1 + ix;
				if (ly < Hl && gx < W)
					__stcg(ldist + (size_t)ly * W + gx, S.d[(1 + iy) * PITCH + 4 + ix]);
			}
		}

		// ---- what do the border cells offer the neighbour tiles? ----
		{
			auto offer = [&](bool exists, uint32_t db, uint32_t w, uint32_t ap) -> uint32_t
			{ return exists && db != 0u && db < BIG_SETTLE_LIMIT && db + MIN_EDGE + w < ap ? db + MIN_EDGE + w : MTD_INF; };
			const int i = (int)lane;
			const uint32_t ct = (uint32_t)(1 * PITCH + 4 + i), cb = (uint32_t)((brow - 1) * PITCH + 4 + i);
			const uint32_t cl = (uint32_t)((1 + i) * PITCH + 4), cr = (uint32_t)((1 + i) * PITCH + 3 + TS);
			const bool row_exists = 1 + i < brow;
			k[0] = warp_redux_min(offer(true, S.d[ct], S.w[ct - PITCH] >> 16, ap_old[0]));
			k[1] = warp_redux_min(offer(true, S.d[cb], S.w[cb] >> 16, ap_old[1]));
			k[2] = warp_redux_min(offer(row_exists, S.d[cl], S.w[cl - 1u] & 0xFFFFu, ap_old[2]));
			k[3] = warp_redux_min(offer(row_exists, S.d[cr], S.w[cr] & 0xFFFFu, ap_old[3]));
			const uint32_t f = __reduce_or_sync(0xFFFFFFFFu, src_sides); // sources on a border line: the neighbour finds them in the list
			if (f & 4u)
				k[0] = 0u;
			if (f & 8u)
				k[1] = 0u;
			if (f & 16u)
				k[2] = 0u;
			if (f & 32u)
				k[3] = 0u;
		}
	}

	constexpr int ASYNC_SLOTS = 2 * ASYNC_WORKERS; // every worker has a mailbox of two: the item it visits and the one it visits next

	struct AsyncSched
	{
		uint32_t pend_item[ASYNC_PCAP];
		uint32_t pend_key[ASYNC_PCAP];
		uint32_t slot[ASYNC_SLOTS];  // mailbox [worker][2]: SLOT_IDLE, SLOT_EXIT or the item to visit (written by the scheduler, reset by the worker)
		uint32_t ifkey[ASYNC_SLOTS]; // the key each handed-out item was given at
		uint32_t asg_slot[ASYNC_SLOTS], asg_item[ASYNC_SLOTS]; // hand-outs of the current scheduler iteration
	};

	template<int TS>
	constexpr size_t async_smem_bytes()
	{
		return sizeof(AsyncFrame<TS>) * ASYNC_WORKERS + sizeof(AsyncSched);
	}

	// ONE launch per update; cooperative only so that all CTAs are resident together (there is no grid barrier inside).
	template<int TS>
	__global__ void __launch_bounds__((ASYNC_WORKERS + 1) * 32, 1) k_tile_async(const BigParams P, const AsyncParams A, const __grid_constant__ CUtensorMap tm_dist, const __grid_constant__ CUtensorMap tm_w)
	{
		extern __shared__ __align__(128) unsigned char smem_async[];
		using WT = AsyncFrame<TS>;
		const int warp = threadIdx.x >> 5;
		const uint32_t lane = lane_id();
		const uint32_t lt = lanemask_lt();
		AsyncSched &Q = *reinterpret_cast<AsyncSched *>(smem_async + sizeof(WT) * ASYNC_WORKERS);
		volatile uint32_t *slot = Q.slot;
		const uint32_t ntiles = (uint32_t)(P.tilesX * P.tilesY);
		if (threadIdx.x < ASYNC_SLOTS)
		{
			Q.slot[threadIdx.x] = SLOT_IDLE;
			Q.ifkey[threadIdx.x] = MTD_INF;
		}
		__syncthreads();

		if (warp < ASYNC_WORKERS)
		{
			// ---------------- worker ----------------
			WT &S = reinterpret_cast<WT *>(smem_async)[warp];
			bool guard = false;
			uint32_t visits = 0, tma_phase = 0, take = 0, last_item = SLOT_IDLE;
#ifdef MTD_TIMING
			unsigned long long tacc[16] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 };
			long long last__ = clock64();
#endif
			if (lane == 0)
			{
				mbar_init(&S.bar, 1u);
				asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
			}
			__syncwarp();
			for (;;)
			{
				const int mine = warp * 2 + (int)take;
				uint32_t it = 0;
				if (lane == 0)
				{
					while ((it = slot[mine]) == SLOT_IDLE)
						__nanosleep(100);
					fence_gpu(); // acquire: the scheduler consumed the item's key (and fenced) before it wrote the mailbox
				}
				it = __shfl_sync(0xFFFFFFFFu, it, 0);
				if (it == SLOT_EXIT)
					break;
				const bool cont = (it & SLOT_CONT) != 0u && (it & ~SLOT_CONT) == last_item; // the scheduler only flags what it gave me last
				it &= ~SLOT_CONT;
				MTD_ATICK(0);
				const int layer = (int)(it / ntiles);
				const uint32_t tile = it - (uint32_t)layer * ntiles;
				const int tx = (int)(tile % (uint32_t)P.tilesX), ty = (int)(tile / (uint32_t)P.tilesX);
				MTD_ATICK(1);
				uint32_t k[4];
#ifdef MTD_TIMING
				const uint32_t depth__ = it < (1u << 20) ? atomicExch(&g_async_depth[it], 0u) : 0u;
				const uint32_t sdepth__ = it < (1u << 20) ? max(atomicExch(&g_async_sdepth_off[it], 0u), g_async_sdepth[it]) : 0u;
#endif
				uint32_t nxt = slot[mine ^ 1];
				if (nxt & SLOT_CONT)
					nxt = 0xFFFFFFFFu; // nothing to prefetch: a continued visit (or an empty mailbox)
				warp_visit<TS>(P, S, layer, tx, ty, lane, lt, k, guard, A.use_tma ? &tm_dist : nullptr, A.use_tma ? &tm_w : nullptr, tma_phase, nxt, A.late_looks, cont MTD_ATIMING_PASS);
				last_item = it;
				if (P.ctl0)
				{
					// row stripes: the tile above / below a border tile belongs to a neighbour rank. Its copy of my boundary row is
					// in ITS memory: store the row there (NVLink), then post the key in its array of border offers; one of its
					// schedulers takes it from there. Done BEFORE my own item is given up below: an offer that made the
					// neighbour's tile pending is counted in ctl[SENT] while this rank still counts as busy (termination, see the
					// scheduler of CTA 0).
					const int Hl = P.y1 - P.y0;
					const int brow = min(TS + 1, Hl - (ty * TS - 1));
					const int gx = tx * TS + (int)lane;
					// The row goes out after EVERY visit of a border tile, offer or not: the neighbour's next-hop pass breaks ties
					// between optimal predecessors on both sides of the cut by (label, tile), so its copy must end up exact even
					// where my label only equals what it already has (found at 8 ranks: profiles/r03i_*).
					const bool up_row = ty == 0 && P.peer_up_act, down_row = ty == P.tilesY - 1 && P.peer_down_act;
					const bool up = up_row && k[0] != MTD_INF, down = down_row && k[1] != MTD_INF;
					if (up_row && gx < P.W)
						P.peer_up_halo[gx] = S.d[1 * WT::PITCH + 4 + (int)lane];
					if (down_row && gx < P.W)
						P.peer_down_halo[gx] = S.d[(brow - 1) * WT::PITCH + 4 + (int)lane];
					if (up_row || down_row)
					{
						__threadfence_system(); // every lane: its part of the row has landed before the offer, and before this rank can count as idle
						__syncwarp();
						if (lane == 0 && (up || down))
						{
							uint32_t made = 0;
							if (up && atomicMin_system(P.peer_up_act + tx, k[0]) == MTD_INF)
								made++;
							if (down && atomicMin_system(P.peer_down_act + tx, k[1]) == MTD_INF)
								made++;
							if (made)
							{
								atomicAdd(A.ctl + ACTL_SENT, made);
								__threadfence(); // counted before this visit's own count is given up
							}
						}
						__syncwarp();
					}
				}
				// offers to the four neighbours. The items they may create are counted BEFORE the offers are made (the
				// count of pending + in-flight items is never behind, so it cannot read zero early); mine is given up with them.
				const uint32_t kk = lane == 0 ? k[0] : lane == 1 ? k[1] : lane == 2 ? k[2] : k[3];
				const bool have = lane < 4u && kk != MTD_INF && (lane == 0 ? ty > 0 : lane == 1 ? ty < P.tilesY - 1 : lane == 2 ? tx > 0 : tx < P.tilesX - 1);
				const uint32_t noff = (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, have));
				if (lane == 0 && noff != 1u)
					atomicAdd(A.ctl + ACTL_OUTSTANDING, noff - 1u); // noff == 0: this visit's own count goes
				fence_gpu(); // my labels (and the count) are visible before a neighbour can be scheduled
				__syncwarp();
				MTD_ATICK(5);
#ifdef MTD_TIMING
				if (have)
				{
					const int nx__ = lane == 2 ? tx - 1 : lane == 3 ? tx + 1 : tx, ny__ = lane == 0 ? ty - 1 : lane == 1 ? ty + 1 : ty;
					const uint32_t ni__ = (uint32_t)layer * ntiles + (uint32_t)(ny__ * P.tilesX + nx__);
					if (ni__ < (1u << 20))
					{
						atomicMax(&g_async_depth[ni__], depth__ + 1u);
						atomicMax(&g_async_sdepth_off[ni__], sdepth__ + 1u);
					}
				}
				if (lane == 0)
				{
					atomicMax(&g_async_depth_max, depth__ + 1u);
					atomicMax(&g_async_sdepth_max, sdepth__ + 1u);
					if (it < (1u << 20))
						g_async_sdepth[it] = sdepth__ + 1u;
				}
#endif
				bool made = false;
				if (have)
					made = async_offer(P, A, layer, lane == 2 ? tx - 1 : lane == 3 ? tx + 1 : tx, lane == 0 ? ty - 1 : lane == 1 ? ty + 1 : ty, kk);
				const uint32_t stale = (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, have && !made));
				visits++;
				if (lane == 0)
				{
					if (stale)
						atomicSub(A.ctl + ACTL_OUTSTANDING, stale); // those items were pending already
					__stcg(A.fly + it, MTD_INF);
					slot[mine] = SLOT_IDLE;
				}
				take ^= 1u;
				__syncwarp();
				MTD_ATICK(6);
			}
#ifdef MTD_TIMING
			if (lane == 0)
				for (int q = 0; q < 16; q++)
					if (tacc[q])
						atomicAdd(&g_async_timing[q], tacc[q]);
#endif
			if (lane == 0)
			{
				atomicAdd(A.ctl + ACTL_VISITS, visits);
				atomicAdd(P.total_active, visits);
				if (guard)
					atomicOr(P.status, 1);
			}
			return;
		}

		// ---------------- scheduler warp ----------------
		const int me = (int)blockIdx.x;
		const uint32_t cap = A.cap_mask + 1u;
		unsigned long long *ring = A.inbox + (size_t)me * cap;
		uint32_t head = 0, pn = 0, polls = 0;
		uint32_t fill = 0; // lane w < WORKERS: which of worker w's two mailboxes is filled next (the worker empties them in the same order)
		uint32_t lastw = SLOT_IDLE; // lane w < WORKERS: the item worker w was given last (its frame holds it once it is through its mailboxes)
		// ---- peer-coupled stripes: entering the run ----
		const bool peer = P.ctl0 != nullptr;
		bool peer_dead = false;
		uint32_t wave = 1, replied = 0, wm_sent = MTD_INF; // CTA 0, lane 0: termination wave being collected (rank 0) / answered, watermark last published
		unsigned long long prev_reply[PEER_MAX_WORLD];
		bool have_prev = false;
		if (peer)
		{
			if (me == 0)
			{
				if (lane == 0)
				{
					// my words of the protocol start clean (rank 0 writes AREQ / ASTOP, the ranks write AWM and -- into rank 0's block --
					// AREPLY only after the barriers below)
					P.ctl[PEER_AREQ] = 0u;
					P.ctl[PEER_ASTOP] = 0u;
					for (int j = 0; j < PEER_MAX_WORLD; j++)
					{
						P.ctl[PEER_AWM + j] = MTD_INF;
						reinterpret_cast<unsigned long long *>(P.ctl + PEER_AREPLY)[j] = 0ull;
					}
					uint32_t epoch = P.ctl[PEER_EPOCH];
					// every rank has entered this run (nobody still reads the halo copies of the previous one) ...
					bool ok = peer_arrive_wait(P, epoch, peer_dead, P.soft_grid ? 4000000000ull : 20000000000ull);
					P.ctl[PEER_EPOCH] = epoch;
					if (!ok)
						atomicOr(P.status, 8); // ST_PEER_TIMEOUT
				}
				__syncwarp();
				// ... my boundary rows as initialised (m, parked blocked tiles and sources 0) go to the neighbours ...
				const int Hl = P.y1 - P.y0;
				for (int x = (int)lane; x < P.W; x += 32)
				{
					if (P.peer_up_halo)
						P.peer_up_halo[x] = __ldcg(P.dist + x);
					if (P.peer_down_halo)
						P.peer_down_halo[x] = __ldcg(P.dist + (size_t)(Hl - 1) * P.W + x);
				}
				__threadfence_system();
				__syncwarp();
				if (lane == 0)
				{
					// ... and have landed everywhere before any rank reads its copies
					uint32_t epoch = P.ctl[PEER_EPOCH];
					if (!peer_arrive_wait(P, epoch, peer_dead))
						atomicOr(P.status, 8);
					P.ctl[PEER_EPOCH] = epoch;
					if (peer_dead)
						atomicExch(A.ctl + ACTL_ABORT, 1u);
					__threadfence();
					atomicExch(A.ctl + ACTL_GO, 1u);
				}
				__syncwarp();
			}
			else
			{
				while (__ldcg(A.ctl + ACTL_GO) == 0u)
					__nanosleep(500);
				fence_gpu();
			}
		}
		const unsigned long long t0 = global_timer_ns();
		bool stop = false;
#ifdef MTD_TIMING
		unsigned long long sacc__[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
		long long slast__ = clock64();
#endif
		while (!stop)
		{
			// everything the decision needs from global memory is requested at once: ONE L2 round trip per iteration
			//   * the next 32 inbox entries past `head` (an entry is 0 until its store has landed: no tail needed),
			//   * the keys of the pending entries (offers only ever lower them),
			//   * every CTA's smallest key, the count of items alive, the abort flag
			MTD_STICK(5);
			const unsigned long long e = __ldcg(ring + ((head + lane) & A.cap_mask));
			const uint32_t outstanding = __ldcg(A.ctl + ACTL_OUTSTANDING);
			const uint32_t aborted = __ldcg(A.ctl + ACTL_ABORT);
			const uint32_t stopped = peer ? __ldcg(A.ctl + ACTL_STOP) : 0u;
			uint32_t others = MTD_INF;
			for (int c = (int)lane; c < A.ctas; c += 32)
				if (c != me)
					others = min(others, __ldcg(A.ctamin + c));
			uint32_t others_local = others; // CTA 0 publishes this rank's own position
			if (peer && (int)lane < P.world && (int)lane != P.rank)
				others = min(others, ld_sys(P.ctl + PEER_AWM + lane)); // the other ranks' wavefront positions
			if (peer)
			{
				// border offers of the neighbour ranks to the border tiles this CTA owns: taken over and made as local offers
				// (counted first, like every offer; RECV after the count: a rank that reports "nothing outstanding" together with
				// a RECV that includes an offer has really finished with it)
				for (int side = 0; side < 2; side++)
				{
					uint32_t *act = side == 0 ? P.act_up : P.act_down;
					if (!act)
						continue;
					const int ty = side == 0 ? 0 : P.tilesY - 1;
					const int first = (int)(((uint32_t)me + (uint32_t)A.ctas * (uint32_t)(ASYNC_KX * ty) - (uint32_t)(ASYNC_KX * ty)) % (uint32_t)A.ctas); // tx with (tx + KX * ty) % ctas == me
					for (int tx = first + (int)lane * A.ctas; tx < P.tilesX; tx += 32 * A.ctas)
					{
						if (ld_sys(act + tx) == MTD_INF)
							continue;
						const uint32_t bk = atomicExch(act + tx, MTD_INF);
						if (bk == MTD_INF)
							continue;
						atomicAdd(A.ctl + ACTL_OUTSTANDING, 1u);
						__threadfence();
						atomicAdd(A.ctl + ACTL_RECV, 1u);
						__threadfence_system(); // the neighbour stored its row before the key: what a worker stages from here on includes it
						if (!async_offer(P, A, 0, tx, ty, bk))
							atomicSub(A.ctl + ACTL_OUTSTANDING, 1u);
					}
				}
				__syncwarp();
			}
			uint32_t mymin = MTD_INF;
			for (uint32_t i = lane; i < pn; i += 32)
			{
				const uint32_t item = Q.pend_item[i];
				uint32_t kk = __ldcg(P.key + item);
				mymin = min(mymin, kk);
				if (A.nb_rule)
				{
					// Tile-level Dijkstra order, locally: a visit that starts while a neighbour tile is still being relaxed, or
					// waits with a smaller key, is repeated when that neighbour's offer arrives (tools/sim: 4.0 -> 2.7 visits
					// per tile). Such an entry stays pending and is looked at again next iteration.
					const uint32_t layer0 = (item / ntiles) * ntiles, tile = item - layer0;
					const int tx = (int)(tile % (uint32_t)P.tilesX), ty = (int)(tile / (uint32_t)P.tilesX);
					uint32_t nk = MTD_INF, nf = MTD_INF;
					if (ty > 0)
					{
						nk = min(nk, __ldcg(P.key + item - P.tilesX));
						nf = min(nf, __ldcg(A.fly + item - P.tilesX));
					}
					if (ty < P.tilesY - 1)
					{
						nk = min(nk, __ldcg(P.key + item + P.tilesX));
						nf = min(nf, __ldcg(A.fly + item + P.tilesX));
					}
					if (tx > 0)
					{
						nk = min(nk, __ldcg(P.key + item - 1));
						nf = min(nf, __ldcg(A.fly + item - 1));
					}
					if (tx < P.tilesX - 1)
					{
						nk = min(nk, __ldcg(P.key + item + 1));
						nf = min(nf, __ldcg(A.fly + item + 1));
					}
					if (nf != MTD_INF || (nk != MTD_INF && nk + A.nb_margin < kk))
						kk = MTD_INF; // hidden for this iteration (its real key is in mymin)
				}
				Q.pend_key[i] = kk;
			}
			MTD_STICK(0);
			// drain: entries in order up to the first empty one; each comes with the key it was offered at
			{
				const uint32_t room = (uint32_t)ASYNC_PCAP - pn;
				const uint32_t missing = __ballot_sync(0xFFFFFFFFu, e == 0ull);
				uint32_t got = missing ? (uint32_t)(__ffs(missing) - 1) : 32u;
				got = min(got, room);
				if (lane < got)
				{
					__stcg(ring + ((head + lane) & A.cap_mask), 0ull);
					const uint32_t kk = (uint32_t)(e >> 32);
					Q.pend_item[pn + lane] = (uint32_t)(e & 0xFFFFFFFFull) - 1u;
					Q.pend_key[pn + lane] = A.nb_rule ? MTD_INF : kk; // with the neighbour rule: judged next iteration, with its neighbours' state
					mymin = min(mymin, kk);
				}
				head += got;
				pn += got;
			}
			__syncwarp();
			// mailboxes: lane l looks at slot l (worker l / 2)
			uint32_t cur = slot[lane];
			if (cur < SLOT_EXIT)
				mymin = min(mymin, Q.ifkey[lane]);
			const uint32_t busy_slots = __ballot_sync(0xFFFFFFFFu, cur < SLOT_EXIT);
			const uint32_t cmin = warp_redux_min(mymin);
			if (lane == 0)
				__stcg(A.ctamin + me, cmin);
			const uint32_t wm = min(cmin, warp_redux_min(others));
			if (peer && me == 0)
			{
				const uint32_t rank_min = min(cmin, warp_redux_min(others_local));
				if (lane == 0)
				{
					// this rank's wavefront position, for the priority windows of the others
					if (rank_min != wm_sent)
					{
						for (int j = 0; j < P.world; j++)
							if (j != P.rank)
								st_sys(P.peers[j] + PEER_AWM + P.rank, rank_min);
						wm_sent = rank_min;
					}
					// ---- termination (no rounds, no global count): rank 0 collects waves of snapshots. A snapshot of a rank =
					// (nothing outstanding?, border offers made, border offers consumed), read in the order consumed ->
					// outstanding -> made with fences in between, which errs towards "not finished" (see the worker and the
					// take-over above). Two CONSECUTIVE waves -- the second requested after the first was complete -- with
					// identical snapshots, every rank idle and made == consumed in total, describe one instant at which nothing
					// was pending, in flight or on its way between ranks: the counters only grow, so a rank whose snapshot did
					// not change was idle the whole time between its two snapshots, and all these intervals contain the moment
					// the second wave was requested.
					const uint32_t req = ld_sys(P.ctl + PEER_AREQ);
					if (req != replied)
					{
						const uint32_t rcv = __ldcg(A.ctl + ACTL_RECV);
						__threadfence();
						const uint32_t out = __ldcg(A.ctl + ACTL_OUTSTANDING);
						__threadfence();
						const uint32_t snt = __ldcg(A.ctl + ACTL_SENT);
						const unsigned long long word = ((unsigned long long)(req & 0xFFu) << 56) | ((unsigned long long)(out == 0u ? 1u : 0u) << 55) | ((unsigned long long)(snt & 0x7FFFFFFu) << 28) | (unsigned long long)(rcv & 0xFFFFFFFu);
						st_sys64(reinterpret_cast<unsigned long long *>(P.peers[0] + PEER_AREPLY) + P.rank, word);
						replied = req;
					}
					if (P.rank == 0 && !stopped)
					{
						if (ld_sys(P.ctl + PEER_AREQ) != wave)
						{
							for (int j = 0; j < P.world; j++)
								st_sys(P.peers[j] + PEER_AREQ, wave);
						}
						else
						{
							bool complete = true, quiet = true, same = have_prev;
							uint32_t ssum = 0, rsum = 0;
							unsigned long long cur[PEER_MAX_WORLD];
							for (int j = 0; j < P.world; j++)
							{
								const unsigned long long w = ld_sys64(reinterpret_cast<const unsigned long long *>(P.ctl + PEER_AREPLY) + j);
								cur[j] = w & 0x00FFFFFFFFFFFFFFull;
								complete = complete && (uint32_t)(w >> 56) == (wave & 0xFFu);
								quiet = quiet && ((w >> 55) & 1ull);
								ssum += (uint32_t)(w >> 28) & 0x7FFFFFFu;
								rsum += (uint32_t)w & 0xFFFFFFFu;
								same = same && cur[j] == prev_reply[j];
							}
							if (complete)
							{
								quiet = quiet && ((ssum - rsum) & 0x7FFFFFFu) == 0u;
								if (quiet && same)
								{
									for (int j = 0; j < P.world; j++)
										st_sys(P.peers[j] + PEER_ASTOP, 1u);
								}
								for (int j = 0; j < P.world; j++)
									prev_reply[j] = cur[j];
								have_prev = quiet;
								wave++;
								if ((wave & 0xFFu) == 0u)
									wave++; // tag 0 is "no reply yet"
							}
						}
					}
					if (ld_sys(P.ctl + PEER_ASTOP) != 0u)
						atomicExch(A.ctl + ACTL_STOP, 1u);
				}
				__syncwarp();
			}
			MTD_STICK(2);
			const uint32_t kthr = wm > MTD_INF - P.tile_delta ? MTD_INF : wm + P.tile_delta;
			// workers that can take an item: first the ones with nothing to do, then the ones that have no NEXT item yet
			uint32_t idle_w = 0, half_w = 0;
			{
				const uint32_t b0 = (busy_slots >> (2u * (lane & 15u))) & 3u;
				idle_w = __ballot_sync(0xFFFFFFFFu, lane < (uint32_t)A.workers && b0 == 0u);
				half_w = __ballot_sync(0xFFFFFFFFu, lane < (uint32_t)A.workers && (b0 == 1u || b0 == 2u) && A.depth > 1);
			}
			__syncwarp();
			uint32_t nasg = 0, contmask = 0;
#ifdef MTD_TIMING
			const uint32_t idle_before__ = (uint32_t)__popc(idle_w);
			uint32_t blocked__ = 0;
#endif
			// All hand-outs of an iteration in one step, lane i deciding for pending entry i (r03z: the one-at-a-time loop below
			// took 60 % of a scheduler iteration, ~2k cycles per item). Runnable = inside the window and its tile in no
			// mailbox. In the latency-bound regime the workers are mostly idle and every runnable entry finds a mailbox;
			// otherwise the entries with the smallest keys take the free ones. The loop below remains for lists longer than
			// a warp and for the optional modes.
			if (!A.cont && pn != 0u && pn <= 32u && (idle_w | half_w) != 0u)
			{
				const uint32_t my_key = lane < pn ? Q.pend_key[lane] : MTD_INF;
				const uint32_t my_item = lane < pn ? Q.pend_item[lane] : 0xFFFFFFFFu;
				bool inflight = false;
#pragma unroll 8
				for (int sl = 0; sl < ASYNC_SLOTS; sl++)
				{
					const uint32_t c = __shfl_sync(0xFFFFFFFFu, cur, sl);
					inflight |= c == my_item; // an idle or exiting mailbox holds no item number
				}
				const uint32_t runnable = __ballot_sync(0xFFFFFFFFu, my_key < kthr && !inflight);
				const uint32_t nrun = (uint32_t)__popc(runnable), nidle = (uint32_t)__popc(idle_w), nfree = nidle + (uint32_t)__popc(half_w);
				if (nrun != 0u)
				{
					// more runnable entries than free mailboxes: the ones with the smallest keys go (rank by (key, lane))
					uint32_t r = (uint32_t)__popc(runnable & lt);
					if (nrun > nfree)
					{
						r = 0u;
						for (int o = 0; o < 32; o++)
						{
							const uint32_t ok = __shfl_sync(0xFFFFFFFFu, my_key, o);
							r += ((runnable >> o) & 1u) && (ok < my_key || (ok == my_key && (uint32_t)o < lane)) ? 1u : 0u;
						}
					}
					const bool sel = ((runnable >> lane) & 1u) && r < nfree;
					const uint32_t nsel = min(nrun, nfree);
					// the r-th free worker: the idle ones first, then the ones without a next item
					const uint32_t w = sel ? (r < nidle ? __fns(idle_w, 0u, (int)r + 1) : __fns(half_w, 0u, (int)(r - nidle) + 1)) : 0u;
					const uint32_t f = __shfl_sync(0xFFFFFFFFu, fill, (int)w);
					const uint32_t keep = __ballot_sync(0xFFFFFFFFu, lane < pn && !sel);
					__syncwarp();
					if (sel)
					{
						const uint32_t sl = w * 2u + f;
						Q.ifkey[sl] = my_key;
						Q.asg_slot[r] = sl;
						Q.asg_item[r] = my_item;
					}
					else if (lane < pn)
					{
						const uint32_t pos = (uint32_t)__popc(keep & lt);
						Q.pend_item[pos] = my_item;
						Q.pend_key[pos] = my_key;
					}
					// every worker that got an item (one each) moves on to its other mailbox
					uint32_t got = 0;
					if (sel)
						got = 1u << w;
					got = __reduce_or_sync(0xFFFFFFFFu, got);
					if ((got >> lane) & 1u)
						fill ^= 1u;
					nasg = nsel;
					pn = (uint32_t)__popc(keep);
					idle_w = half_w = 0u; // nothing left for the loop below: every runnable entry or every free mailbox is used
					__syncwarp();
				}
			}
			while ((idle_w | half_w) != 0u && pn != 0u)
			{
				uint32_t bk = MTD_INF;
				for (uint32_t i = lane; i < pn; i += 32)
					bk = min(bk, Q.pend_key[i]);
				const uint32_t kmin = warp_redux_min(bk);
				if (kmin >= kthr) // also MTD_INF: nothing selectable
				{
#ifdef MTD_TIMING
					blocked__ = kmin != MTD_INF ? 1u : 0u;
#endif
					break;
				}
				uint32_t bi = MTD_INF;
				for (uint32_t i = lane; i < pn; i += 32)
					if (Q.pend_key[i] == kmin)
						bi = min(bi, i);
				bi = warp_redux_min(bi);
				const uint32_t item = Q.pend_item[bi];
				// A revisit costs least on the worker whose frame still holds the tile: it skips the staging and only looks at
				// the apron again. Most revisits are back to back -- offers from several neighbours reach a wavefront tile within
				// one visit time -- so the item goes into the OTHER mailbox of the worker that is visiting it right now, or, if
				// that visit is over, to that worker if it is idle.
				const uint32_t hit = __ballot_sync(0xFFFFFFFFu, cur < SLOT_EXIT && (cur & ~SLOT_CONT) == item);
				int w = -1;
				bool as_cont = false;
				if (hit)
				{
					const int hw = (__ffs(hit) - 1) >> 1;
					if (A.cont && __popc(hit) == 1 && ((half_w >> hw) & 1u) && __shfl_sync(0xFFFFFFFFu, lastw, hw) == item)
					{
						w = hw;
						as_cont = true;
						half_w &= ~(1u << hw);
					}
					else
					{
						// its tile is being visited and cannot be queued behind that visit: it stays pending; hidden for this iteration
						if (lane == 0)
							Q.pend_key[bi] = MTD_INF;
						__syncwarp();
						continue;
					}
				}
				else
				{
					const uint32_t warm = A.cont ? __ballot_sync(0xFFFFFFFFu, lane < (uint32_t)A.workers && lastw == item) & idle_w : 0u;
					if (warm)
					{
						w = __ffs(warm) - 1;
						as_cont = true;
						idle_w &= ~(1u << w);
					}
					else if (idle_w)
					{
						w = __ffs(idle_w) - 1;
						idle_w &= idle_w - 1u;
					}
					else
					{
						w = __ffs(half_w) - 1;
						half_w &= half_w - 1u;
					}
				}
				const uint32_t f = __shfl_sync(0xFFFFFFFFu, fill, w);
				const uint32_t sl = (uint32_t)w * 2u + f;
				if (lane == (uint32_t)w)
					lastw = item;
				if (as_cont)
					contmask |= 1u << nasg;
				if (lane == (uint32_t)w)
					fill ^= 1u;
				if (lane == sl)
					cur = item;
				if (lane == 0)
				{
					Q.ifkey[sl] = kmin;
					Q.asg_slot[nasg] = sl;
					Q.asg_item[nasg] = item;
					Q.pend_item[bi] = Q.pend_item[pn - 1];
					Q.pend_key[bi] = Q.pend_key[pn - 1];
				}
				nasg++;
				pn--;
				__syncwarp();
			}
			MTD_STICK(3);
			if (nasg)
			{
				// consume the offers of everything handed out (whoever made one wrote its labels first), THEN ring the mailboxes
				if (lane < nasg)
				{
					__stcg(A.fly + Q.asg_item[lane], Q.ifkey[Q.asg_slot[lane]]);
					atomicExch(P.key + Q.asg_item[lane], MTD_INF);
				}
				fence_gpu();
				__syncwarp();
				if (lane < nasg)
					slot[Q.asg_slot[lane]] = Q.asg_item[lane] | (((contmask >> lane) & 1u) ? SLOT_CONT : 0u);
				__syncwarp();
			}
			MTD_STICK(4);
			polls++;
#ifdef MTD_TIMING
			{
				// pending entries that could run right now (inside the window, tile not in flight) but found no free mailbox here
				uint32_t sur = 0, hid = 0, far = 0;
				for (uint32_t i = lane; i < pn; i += 32)
				{
					const uint32_t kk = Q.pend_key[i];
					sur += kk < kthr ? 1u : 0u;
					hid += kk == MTD_INF ? 1u : 0u;
					far += (kk != MTD_INF && kk >= kthr) ? 1u : 0u;
				}
				sur = __reduce_add_sync(0xFFFFFFFFu, sur);
				hid = __reduce_add_sync(0xFFFFFFFFu, hid);
				far = __reduce_add_sync(0xFFFFFFFFu, far);
				if (lane == 0)
				{
					atomicAdd(&g_async_timing[14], (unsigned long long)sur);
					atomicAdd(&g_async_timing[15], (unsigned long long)far);
					atomicAdd(&g_async_timing[8 + 0], 0ull);
					atomicAdd(&g_async_sched_extra[0], (unsigned long long)hid);
					atomicAdd(&g_async_sched_extra[1], (unsigned long long)nasg);
				}
			}
			if (lane == 0)
			{
				// 9 iterations, 10 idle workers seen, 11 idle workers left without an item, 12 ... because the window held the rest back, 13 pending entries seen
				const uint32_t idle_after__ = (uint32_t)__popc(idle_w);
				atomicAdd(&g_async_timing[9], 1ull);
				atomicAdd(&g_async_timing[10], (unsigned long long)idle_before__);
				atomicAdd(&g_async_timing[11], (unsigned long long)idle_after__);
				atomicAdd(&g_async_timing[12], (unsigned long long)(blocked__ ? idle_after__ : 0u));
				atomicAdd(&g_async_timing[13], (unsigned long long)(pn + nasg));
			}
#endif
			if (aborted || (peer ? stopped != 0u : outstanding == 0u))
				stop = true;
			else if (global_timer_ns() - t0 > A.time_limit_ns)
			{
				if (lane == 0)
				{
					atomicExch(A.ctl + ACTL_ABORT, 1u);
					atomicOr(P.status, 16); // ST_ASYNC_TIMEOUT
				}
				stop = true;
			}
			else if (nasg == 0)
				__nanosleep(100);
		}
		// release the workers: a busy one finishes what is in its mailboxes first
		for (;;)
		{
			uint32_t st = slot[lane];
			if (st == SLOT_IDLE)
			{
				slot[lane] = SLOT_EXIT;
				st = SLOT_EXIT;
			}
			if (!__any_sync(0xFFFFFFFFu, st != SLOT_EXIT))
				break;
			__nanosleep(200);
		}
#ifdef MTD_TIMING
		if (lane == 0)
			for (int q = 0; q < 8; q++)
				atomicAdd(&g_async_sched_phase[q], sacc__[q]);
#endif
		if (lane == 0)
			atomicAdd(A.ctl + ACTL_IDLE_POLLS, polls);
	}
}
