// big_grid.cuh -- distance / direction maps for grids that do not fit shared memory (4096^2 ... 16384^2).
//
// Directions::update (sources/map/paths.cpp:24-78) as tile-local wavefront relaxation, all layers of one
// Waypoints::update in flight together (a work item is one (layer, tile) pair):
//   k_edge_weights   Grid::neighborDistance of every right/down edge, once per grid (sources/map/grid.cpp:141-145)
//   k_big_init       distances := m, blocked tiles PARKED at 0 (paths.cpp:40-43, 65-67), work-list cleared
//   k_big_seed       sources := 0 (paths.cpp:48-49), their tiles queued for round 0
//   k_tile_round     one global round: every active 64x64 tile is staged with a 1-cell apron in shared memory by one
//                    4-warp CTA (16-byte cp.async.cg rows, L2-coherent), relaxed to local convergence by a bucketed
//                    wavefront (measured: ~64 buckets per visit = the hop depth of a tile, whatever the bucket width),
//                    written back with 16-byte stores, and the tiles across any improved border line are queued for
//                    the next round
//   k_round_advance  flips the double-buffered work-list
//   k_big_finalize   tiles[v] from final distances (closed form of the heap order, see sssp_warp.cuh); parked blocked
//                    tiles read m again
// The fixpoint is unique, so the result is bit-identical to the reference whatever the schedule.
//
// Frame layout in shared memory: (TS + 2) rows of PITCH = TS + 8 words; grid column gx0 + rx sits at word rx + 3, so the
// 64 interior words of a row start 16-byte aligned. The apron is read once (each finite apron label relaxes its one
// interior neighbour) and then parked at 0 like blocked cells: the four neighbours of an interior cell are p-1, p+1,
// p-PITCH, p+PITCH with no bounds or flag test, and nothing is ever relaxed outward.
//
// Row-striped multi-GPU: each rank runs the same kernels on rows [y0, y1) of the grid; the apron rows just
// outside a stripe are read from the neighbour rank's distance array through a peer pointer (NVLink load) and
// improvements of a stripe's first / last row raise a flag the neighbour rank turns into tile activations.
#pragma once

#include "common.cuh"
#include "sssp_warp.cuh"

namespace mtd
{
	constexpr uint32_t BIG_SETTLE_LIMIT = MTD_GUARD - (1u << 17); // labels this close to the guard band are reported, not expanded
#ifndef MTD_BIG_DELTA
#define MTD_BIG_DELTA 1024u
#endif
	// bucket width of the in-tile wavefront. 255 (= the minimum edge weight) settles only final labels; wider buckets take
	// fewer rounds and re-relax a few labels (the engine re-queues every improvement, so any width is exact).
	// Measured on 4096^2 / 30 % walls / weights 255..570: 255 -> 72 buckets per visit, 1024 -> 64 (the hop depth).
	constexpr uint32_t BIG_DELTA = MTD_BIG_DELTA;

	struct BigParams
	{
		int W, H;               // full grid
		int y0, y1;             // rows owned by this rank (single GPU: 0, H)
		int tilesX, tilesY;     // tiles of the owned stripe
		int nlayers, nsrc;      // independent layers relaxed together; sources per layer
		uint32_t *dist;         // [nlayers][H_local][W] rows y0..y1 (local row r = y - y0)
		size_t layer_stride;    // elements between layers of dist
		const uint32_t *wpack;  // [H_local][W]  (wr - 255) | (wd - 255) << 16 ; wd of the last local row refers to row y1 (valid if y1 < H)
		const uint8_t *blocked; // [H_local][W]
		const uint32_t *sources; // device array [nlayers][nsrc] of global tile indices
		uint32_t *list[2];      // work-lists of items = layer * ntiles + tile
		uint32_t *count;        // [2]
		uint32_t *cursor;       // [1]
		uint32_t *stamp;        // [nlayers * ntiles] last round this item was queued for
		uint32_t *total_active; // [1] running sum of processed items (statistics)
		int *status;
		int aligned16;          // W % 4 == 0 and every base pointer 16-byte aligned: rows move as 16-byte vectors
		// stripe neighbours (null on a single GPU or at the grid edge); nlayers == 1 when striped
		const uint32_t *up_row;    // distance row y0-1 on the rank above (peer memory), [W]
		const uint32_t *down_row;  // distance row y1 on the rank below, [W]
		const uint32_t *up_wpack;  // wpack row y0-1 (for wd of the edge (y0-1) -> y0), local copy
		uint32_t *peer_flag_up;    // [tilesX] set to 1 when this rank improved its first row over that tile column
		uint32_t *peer_flag_down;  // [tilesX] same for the last row
	};

	// ---- edge weights -------------------------------------------------------------------------------
	__global__ void k_edge_weights(int W, int Hl, const int16_t *__restrict__ elev, const int16_t *__restrict__ elev_below /* row after the stripe or null */, const uint16_t *__restrict__ flags, uint32_t *__restrict__ wpack, uint8_t *__restrict__ blocked)
	{
		const size_t total = (size_t)W * Hl;
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
		{
			const int x = (int)(i % (size_t)W);
			const int y = (int)(i / (size_t)W);
			const int16_t e = elev[i];
			uint32_t r = 0, d = 0;
			if (x < W - 1)
				r = edge_weight(e, elev[i + 1]) - MIN_EDGE;
			if (y < Hl - 1)
				d = edge_weight(e, elev[i + W]) - MIN_EDGE;
			else if (elev_below)
				d = edge_weight(e, elev_below[x]) - MIN_EDGE;
			wpack[i] = r | (d << 16);
			blocked[i] = (flags[i] & F_BLOCKING) ? 1 : 0;
		}
	}

	// ---- init ----------------------------------------------------------------------------------------
	// every layer: m everywhere, blocked tiles parked at 0 (never relaxed: >= 255 cannot lower 0)
	__global__ void k_big_init(uint32_t *__restrict__ dist, size_t layer_stride, int nlayers, size_t n, const uint8_t *__restrict__ blocked)
	{
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
		{
			const uint32_t v = blocked[i] ? 0u : MTD_INF;
			for (int l = 0; l < nlayers; l++)
				__stcg(dist + (size_t)l * layer_stride + i, v);
		}
	}

	template<int TS>
	__global__ void k_big_seed(const BigParams P)
	{
		// one thread per (layer, source): distance 0 and queue its tile for round 0
		const int s = blockIdx.x * blockDim.x + threadIdx.x;
		if (s >= P.nlayers * P.nsrc)
			return;
		const int layer = s / P.nsrc;
		const uint32_t src = P.sources[s];
		const int x = (int)(src % (uint32_t)P.W), y = (int)(src / (uint32_t)P.W);
		if (y < P.y0 || y >= P.y1)
			return;
		P.dist[(size_t)layer * P.layer_stride + (size_t)(y - P.y0) * P.W + x] = 0u;
		const uint32_t ntiles = (uint32_t)(P.tilesX * P.tilesY);
		const uint32_t item = (uint32_t)layer * ntiles + (uint32_t)(((y - P.y0) / TS) * P.tilesX + x / TS);
		if (atomicMax(&P.stamp[item], 1u) < 1u) // round r uses stamp r + 1
			P.list[0][atomicAdd(&P.count[0], 1u)] = item;
	}

	__global__ void k_round_advance(const BigParams P, int round, uint32_t *host_flag)
	{
		// after round `round`: reset the consumed list and the cursor; publish the size of the next list
		P.count[round & 1] = 0u;
		*P.cursor = 0u;
		if (host_flag)
			*host_flag = P.count[(round + 1) & 1];
	}

#ifdef MTD_TIMING
	// developer probe (never in the shipped build): cycles per phase of a tile visit, summed over all visits
	__device__ unsigned long long g_big_timing[8];
#define MTD_BTICK(k)                                                      \
	do                                                                    \
	{                                                                     \
		const long long now__ = clock64();                                \
		if (threadIdx.x == 0)                                             \
			atomicAdd(&g_big_timing[k], (unsigned long long)(now__ - last__)); \
		last__ = now__;                                                   \
	} while (0)
#else
#define MTD_BTICK(k)
#endif

	// ---- one round ------------------------------------------------------------------------------------
	template<int TS>
	struct TileSmem
	{
		static constexpr int RW = TS + 2;      // frame rows / columns (interior + apron)
		static constexpr int PITCH = TS + 8;   // words per frame row; column rx lives at word rx + 3
		static constexpr int CELLS = RW * PITCH;
		static constexpr int QCAP = 2048;
		uint32_t d[CELLS];        // labels
		uint32_t w[CELLS];        // packed edge weights of the same cells
		uint32_t apron_old[4][TS]; // top, bottom, left, right apron labels as loaded (the neighbour tiles' border cells)
		uint16_t q[2][QCAP];      // pending cells (frame word index), double buffered
		uint32_t tail[3], newmin[3]; // queue length / smallest pending label, rotating slots (see the engine)
		uint32_t flags;           // bit 1: sweep changed something; bits 2..5: border line improved (top, bottom, left, right)
		uint32_t item;
	};

	__device__ __forceinline__ void cp_async_16(void *smem_dst, const void *gsrc)
	{
		asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
	}

	// 4-byte variant (.ca is the only cache operator it has): used for the two apron columns. A line cached in L1 can
	// be stale only against stores of the SAME round, and every such store re-activates this tile for the next round
	// (a new kernel, L1 invalidated), so a stale read costs nothing but is never final.
	__device__ __forceinline__ void cp_async_4(void *smem_dst, const void *gsrc)
	{
		asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
	}

	__device__ __forceinline__ void cp_async_wait_all()
	{
		asm volatile("cp.async.wait_all;" ::: "memory");
	}

	__device__ __forceinline__ void activate_item(const BigParams &P, int layer, int tx, int ty, int round)
	{
		if (tx < 0 || tx >= P.tilesX || ty < 0 || ty >= P.tilesY)
			return;
		const uint32_t item = (uint32_t)layer * (uint32_t)(P.tilesX * P.tilesY) + (uint32_t)(ty * P.tilesX + tx);
		const uint32_t want = (uint32_t)round + 2u; // stamp for round + 1
		if (atomicMax(&P.stamp[item], want) < want)
			P.list[(round + 1) & 1][atomicAdd(&P.count[(round + 1) & 1], 1u)] = item;
	}

	// One CTA of WARPS warps per (layer, tile) work item.
	template<int TS, int WARPS>
	__global__ void __launch_bounds__(WARPS * 32) k_tile_round(const BigParams P, int round)
	{
		extern __shared__ __align__(16) unsigned char smem_raw[];
		using TSm = TileSmem<TS>;
		TSm &S = *(TSm *)smem_raw;
		constexpr int RW = TSm::RW, PITCH = TSm::PITCH, QCAP = TSm::QCAP, NT = WARPS * 32;
		const int tid = threadIdx.x;
		const int warp = tid >> 5;
		const uint32_t lane = lane_id();
		const uint32_t lt = lanemask_lt();
		const int W = P.W;
		const int Hl = P.y1 - P.y0;
		const uint32_t ntiles = (uint32_t)(P.tilesX * P.tilesY);
		const uint32_t *list = P.list[round & 1];
		const uint32_t count = P.count[round & 1];

		for (;;)
		{
			__syncthreads(); // the previous item's shared memory is free
			if (tid == 0)
				S.item = atomicAdd(P.cursor, 1u);
			__syncthreads();
			const uint32_t item = S.item;
			if (item >= count)
				break;
			const uint32_t it = list[item];
			const int layer = (int)(it / ntiles);
			const uint32_t tile = it - (uint32_t)layer * ntiles;
			const int tx = (int)(tile % (uint32_t)P.tilesX), ty = (int)(tile / (uint32_t)P.tilesX);
			const int gx0 = tx * TS - 1; // global x of frame column 0
			const int ly0 = ty * TS - 1; // local row of frame row 0
			uint32_t *ldist = P.dist + (size_t)layer * P.layer_stride;
			const uint32_t *lsrc = P.sources + (size_t)layer * P.nsrc;
			const bool full_w = P.aligned16 && gx0 + 1 + TS <= W; // the 64 interior columns are inside the grid and 16-byte aligned
			const int brow = min(TS + 1, Hl - ly0); // frame row just below the tile's last row that exists in this stripe
#ifdef MTD_TIMING
			long long last__ = clock64();
#endif
			if (tid < 3)
			{
				S.tail[tid] = 0u;
				S.newmin[tid] = MTD_INF;
			}
			if (tid == 0)
				S.flags = 0u;

			// ---- stage the frame: labels + packed weights. Cells outside the grid are parked at 0. ----
			if (full_w && ly0 >= 0 && ly0 + RW <= Hl && gx0 >= 0 && gx0 + RW <= W)
			{
				// interior tile (the common case): no row or column tests. Per frame row 16 + 16 sixteen-byte copies
				// (labels, weights) and 2 + 2 apron words; one warp instruction moves one row.
				const int k = (int)(lane & 15u) * 4;
				const uint32_t *src = (lane < 16u ? ldist : P.wpack) + (size_t)ly0 * W + gx0 + 1 + k;
				uint32_t *dst = (lane < 16u ? S.d : S.w) + 4 + k;
				const int rxa = (lane & 1u) ? RW - 1 : 0;
				const uint32_t *asrc = ((lane & 2u) ? P.wpack : ldist) + (size_t)ly0 * W + gx0 + rxa;
				uint32_t *adst = ((lane & 2u) ? S.w : S.d) + 3 + rxa;
				for (int ry = warp; ry < RW; ry += WARPS)
				{
					cp_async_16(dst + ry * PITCH, src + (size_t)ry * W);
					if (lane < 4u)
						cp_async_4(adst + ry * PITCH, asrc + (size_t)ry * W);
				}
			}
			else
			{
				for (int ry = warp; ry < RW; ry += WARPS)
				{
					const int ly = ly0 + ry; // local row; may be -1 or Hl (neighbour stripe or outside)
					const uint32_t *drow = nullptr;
					const uint32_t *wrow = nullptr;
					if (ly >= 0 && ly < Hl)
					{
						drow = ldist + (size_t)ly * W;
						wrow = P.wpack + (size_t)ly * W;
					}
					else if (ly < 0 && P.up_row)
					{
						drow = P.up_row;
						wrow = P.up_wpack;
					}
					else if (ly == Hl && P.down_row)
						drow = P.down_row; // weights of the row below are never needed (edges are stored on the upper / left cell)
					uint32_t *sd = S.d + ry * PITCH + 3;
					uint32_t *sw = S.w + ry * PITCH + 3;
					if (drow && full_w)
					{
						const int k = (int)(lane & 15u) * 4;
						if (lane < 16u)
							cp_async_16(sd + 1 + k, drow + gx0 + 1 + k);
						else if (wrow)
							cp_async_16(sw + 1 + k, wrow + gx0 + 1 + k);
						else
							*reinterpret_cast<uint4 *>(sw + 1 + k) = make_uint4(0u, 0u, 0u, 0u);
						if (lane < 4u)
						{
							const int rx = (lane & 1u) ? RW - 1 : 0;
							const int gx = gx0 + rx;
							const bool in = gx >= 0 && gx < W;
							uint32_t *dst = lane < 2u ? sd + rx : sw + rx;
							const uint32_t *src = lane < 2u ? drow : wrow;
							if (in && src)
								cp_async_4(dst, src + gx);
							else
								*dst = 0u;
						}
					}
					else
					{
						for (int rx = (int)lane; rx < RW; rx += 32)
						{
							const int gx = gx0 + rx;
							const bool in = drow && gx >= 0 && gx < W;
							sd[rx] = in ? __ldcg(drow + gx) : 0u;
							sw[rx] = in && wrow ? wrow[gx] : 0u;
						}
					}
				}
			}
			cp_async_wait_all();
			__syncthreads();
			MTD_BTICK(0);

			// ---- seeds ----
			// Pending cells enter q[0]. A label lowered here or later is appended every time it improves (duplicates are
			// harmless: the second visit relaxes nothing). Appends: ballot compaction per warp, one shared-memory atomicAdd
			// on the tail per warp instruction.
			uint32_t my_min = MTD_INF;
			bool guard = false;
			auto relax_into = [&](bool want, uint32_t v, uint32_t nd)
			{
				// want: this lane has a candidate label nd for interior cell v
				bool improved = false;
				if (want && nd >= BIG_SETTLE_LIMIT)
					guard = true; // the reference's uint32 arithmetic is about to wrap: report, do not expand
				else if (want)
					improved = nd < atomicMin(&S.d[v], nd);
				const uint32_t bal = __ballot_sync(FULL_MASK, improved);
				if (bal)
				{
					uint32_t base = 0;
					if (lane == 0)
						base = atomicAdd(&S.tail[0], (uint32_t)__popc(bal));
					base = __shfl_sync(FULL_MASK, base, 0);
					if (improved)
					{
						const uint32_t slot = base + __popc(bal & lt);
						if (slot < (uint32_t)QCAP)
							S.q[0][slot] = (uint16_t)v;
						my_min = min(my_min, nd);
					}
				}
			};
			// (1) apron: every finite, non-parked apron label relaxes its single interior neighbour; a label of 0 is a
			//     parked cell unless the source list says otherwise, so sources on the apron are handled from the list.
			//     4 sides x TS cells, one candidate per thread and step.
			for (int j0 = warp * 32; j0 < 4 * TS; j0 += NT)
			{
				const int j = j0 + (int)lane;
				const int side = j / TS, i = j - side * TS; // warp-uniform side (TS is a multiple of 32)
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
				S.apron_old[side][i] = da; // what the neighbour tile holds there now: decides at the end whether it needs a visit
				relax_into(da != 0u && da != MTD_INF, v, da + MIN_EDGE + wsel);
			}
			for (int s0 = warp * 32; s0 < P.nsrc; s0 += NT)
			{
				const int s = s0 + (int)lane;
				bool want = false;
				uint32_t v = 0, nd = 0;
				if (s < P.nsrc)
				{
					const uint32_t src = lsrc[s];
					const int rx = (int)(src % (uint32_t)W) - gx0, ry = (int)(src / (uint32_t)W) - P.y0 - ly0;
					const bool inx = rx >= 1 && rx <= TS, iny = ry >= 1 && ry <= TS;
					const uint32_t a = (uint32_t)(ry * PITCH + rx + 3);
					if (ry == 0 && inx)
					{
						want = true;
						v = a + PITCH;
						nd = MIN_EDGE + (S.w[a] >> 16);
					}
					else if (ry == TS + 1 && inx)
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
			__syncthreads();
			// (2) park the apron: nothing is relaxed outward, and interior cells need no bounds test
			for (int i = tid; i < TS; i += NT)
			{
				S.d[0 * PITCH + 4 + i] = 0u;
				S.d[brow * PITCH + 4 + i] = 0u;
				S.d[(1 + i) * PITCH + 3] = 0u;
				S.d[(1 + i) * PITCH + 4 + TS] = 0u;
			}
			__syncthreads();
			// (3) sources inside the tile expand from label 0 (also when they are blocked: paths.cpp:63-67 test only the neighbour)
			for (int s0 = warp * 32; s0 < P.nsrc; s0 += NT)
			{
				const int s = s0 + (int)lane;
				bool want = false;
				uint32_t c = (uint32_t)(PITCH + 4);
				if (s < P.nsrc)
				{
					const uint32_t src = lsrc[s];
					const int rx = (int)(src % (uint32_t)W) - gx0, ry = (int)(src / (uint32_t)W) - P.y0 - ly0;
					want = rx >= 1 && rx <= TS && ry >= 1 && ry <= TS;
					if (want)
					{
						c = (uint32_t)(ry * PITCH + rx + 3);
						// a source on a border line reaches into the neighbour tile: that tile finds it through the source list
						const uint32_t sides = (ry == 1 ? 4u : 0u) | (ry == TS ? 8u : 0u) | (rx == 1 ? 16u : 0u) | (rx == TS ? 32u : 0u);
						if (sides)
							atomicOr(&S.flags, sides);
					}
				}
				relax_into(want, c - PITCH, MIN_EDGE + (S.w[c - PITCH] >> 16));
				relax_into(want, c - 1u, MIN_EDGE + (S.w[c - 1u] & 0xFFFFu));
				relax_into(want, c + 1u, MIN_EDGE + (S.w[c] & 0xFFFFu));
				relax_into(want, c + PITCH, MIN_EDGE + (S.w[c] >> 16));
			}
			my_min = warp_min_u32(my_min);
			if (lane == 0 && my_min != MTD_INF)
				atomicMin(&S.newmin[0], my_min);
			__syncthreads();
			MTD_BTICK(1);
			// ---- bucketed wavefront on the frame (label-correcting across visits and inside wide buckets) ----
			// Control words rotate through three slots: round r reads slot r % 3, appends to (r + 1) % 3 and clears
			// (r + 2) % 3, so one barrier per round is enough.
			uint32_t si = 0, par = 0;
			bool overflow = false;
			for (;;)
			{
				const uint32_t nq = S.tail[si], dmin = S.newmin[si];
				if (nq > (uint32_t)QCAP)
				{
					overflow = true;
					break;
				}
				if (nq == 0u || dmin == MTD_INF)
					break;
				const uint32_t so = si == 2u ? 0u : si + 1u, sr = so == 2u ? 0u : so + 1u;
				if (tid == 0)
				{
					S.tail[sr] = 0u;
					S.newmin[sr] = MTD_INF;
				}
				const uint32_t thr = dmin + BIG_DELTA;
				const uint16_t *rq = S.q[par];
				uint16_t *wq = S.q[par ^ 1u];
				uint32_t newmin = MTD_INF;
#ifdef MTD_TIMING
				if (tid == 0)
				{
					atomicAdd(&g_big_timing[5], 1ull);
					atomicAdd(&g_big_timing[6], (unsigned long long)nq);
				}
#endif
				for (uint32_t base = (uint32_t)warp * 32u; base < nq; base += (uint32_t)NT)
				{
					const uint32_t idx = base + lane;
					const bool active = idx < nq;
					const uint32_t c = active ? rq[idx] : (uint32_t)(PITCH + 4);
					const uint32_t d = active ? S.d[c] : MTD_INF;
					const bool settle = d < thr;
					const bool keep = active && !settle;
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
					const uint32_t ck = __popc(bk), c0n = __popc(b0), c1n = __popc(b1), c2n = __popc(b2), c3n = __popc(b3);
					const uint32_t total = ck + c0n + c1n + c2n + c3n;
					if (total)
					{
						uint32_t pos = 0;
						if (lane == 0)
							pos = atomicAdd(&S.tail[so], total);
						pos = __shfl_sync(FULL_MASK, pos, 0);
						uint32_t p = pos + __popc(bk & lt);
						if (keep && p < (uint32_t)QCAP)
							wq[p] = (uint16_t)c;
						pos += ck;
						p = pos + __popc(b0 & lt);
						if (i0 && p < (uint32_t)QCAP)
							wq[p] = (uint16_t)(c - PITCH);
						pos += c0n;
						p = pos + __popc(b1 & lt);
						if (i1 && p < (uint32_t)QCAP)
							wq[p] = (uint16_t)(c - 1u);
						pos += c1n;
						p = pos + __popc(b2 & lt);
						if (i2 && p < (uint32_t)QCAP)
							wq[p] = (uint16_t)(c + 1u);
						pos += c2n;
						p = pos + __popc(b3 & lt);
						if (i3 && p < (uint32_t)QCAP)
							wq[p] = (uint16_t)(c + PITCH);
					}
				}
				newmin = warp_min_u32(newmin);
				if (lane == 0 && newmin != MTD_INF)
					atomicMin(&S.newmin[so], newmin);
				__syncthreads();
				si = so;
				par ^= 1u;
			}
			MTD_BTICK(2);
			if (overflow)
			{
				// the pending set outgrew the queue (pathological frontiers): queue-less sweeps to the local fixpoint.
				// Parked cells hold 0 and never expand; sources expanded in (3) before anything could overflow.
				for (;;)
				{
					__syncthreads();
					if (tid == 0)
						S.flags &= ~2u;
					__syncthreads();
					bool changed = false;
					for (int i = tid; i < TS * TS; i += NT)
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
					if (changed)
						atomicOr(&S.flags, 2u);
					__syncthreads();
					if (!(S.flags & 2u))
						break;
				}
			}
			if (guard)
				atomicOr(P.status, 1);
			__syncthreads();

			// ---- write back the interior (labels only ever decrease; parked and unreached cells rewrite their own value) ----
			if (full_w)
			{
				// NT / 16 rows per instruction, 16 bytes per lane
				for (int r0 = 0; r0 < TS; r0 += NT / 16)
				{
					const int iy = r0 + (tid >> 4);
					const int ly = ly0 + 1 + iy;
					if (iy < TS && ly < Hl)
					{
						const int k = (tid & 15) * 4;
						const uint4 v = *reinterpret_cast<const uint4 *>(S.d + (1 + iy) * PITCH + 4 + k);
						__stcg(reinterpret_cast<uint4 *>(ldist + (size_t)ly * W + gx0 + 1 + k), v);
					}
				}
			}
			else
			{
				for (int iy = warp; iy < TS; iy += WARPS)
				{
					const int ly = ly0 + 1 + iy;
					if (ly >= Hl)
						break;
					for (int ix = (int)lane; ix < TS; ix += 32)
						if (gx0 + 1 + ix < W)
							__stcg(ldist + (size_t)ly * W + gx0 + 1 + ix, S.d[(1 + iy) * PITCH + 4 + ix]);
				}
			}
			MTD_BTICK(3);
			// ---- activate neighbours across improved border lines ----
			{
				// A neighbour tile needs a visit only if one of our border cells now offers its adjacent cell a smaller label
				// than that cell held when we staged it (labels only decrease, so a stale apron only over-activates; a parked
				// or source apron cell holds 0 and can never be improved). Our own parked cells (0) offer nothing.
				uint32_t ch = 0;
				auto offers = [&](uint32_t db, uint32_t w, uint32_t ap) -> bool { return db != 0u && db < BIG_SETTLE_LIMIT && db + MIN_EDGE + w < ap; };
				for (int i = tid; i < TS; i += NT)
				{
					const uint32_t ct = (uint32_t)(1 * PITCH + 4 + i), cb = (uint32_t)(TS * PITCH + 4 + i);
					ch |= offers(S.d[ct], S.w[ct - PITCH] >> 16, S.apron_old[0][i]) ? 4u : 0u;
					ch |= (TS < brow && offers(S.d[cb], S.w[cb] >> 16, S.apron_old[1][i])) ? 8u : 0u;
					if (1 + i < brow) // rows below a cut-short tile hold the parked halo / nothing
					{
						const uint32_t cl = (uint32_t)((1 + i) * PITCH + 4), cr = (uint32_t)((1 + i) * PITCH + 3 + TS);
						ch |= offers(S.d[cl], S.w[cl - 1u] & 0xFFFFu, S.apron_old[2][i]) ? 16u : 0u;
						ch |= offers(S.d[cr], S.w[cr] & 0xFFFFu, S.apron_old[3][i]) ? 32u : 0u;
					}
				}
				if (ch)
					atomicOr(&S.flags, ch);
			}
			__syncthreads();
			if (tid == 0)
			{
				const uint32_t f = S.flags;
				__threadfence(); // distances are visible before the neighbour can be scheduled
				if (f & 4u)
				{
					if (ty > 0)
						activate_item(P, layer, tx, ty - 1, round);
					else if (P.peer_flag_up)
						P.peer_flag_up[tx] = 1u;
				}
				if (f & 8u)
				{
					if (ty < P.tilesY - 1)
						activate_item(P, layer, tx, ty + 1, round);
					else if (P.peer_flag_down)
						P.peer_flag_down[tx] = 1u;
				}
				if (f & 16u)
					activate_item(P, layer, tx - 1, ty, round);
				if (f & 32u)
					activate_item(P, layer, tx + 1, ty, round);
				atomicAdd(P.total_active, 1u);
			}
			MTD_BTICK(4);
		}
	}

	// ---- row stripes (one rank per stripe) ------------------------------------------------------------------
	// Sources on the halo rows (y0-1 and y1) belong to the neighbour stripes, but their first hop lands here: queue the
	// border tile under / over them for round 0 (the visit finds them through the source list).
	template<int TS>
	__global__ void k_stripe_seed_halo(const BigParams P)
	{
		const int s = blockIdx.x * blockDim.x + threadIdx.x;
		if (s >= P.nsrc)
			return;
		const uint32_t src = P.sources[s];
		const int x = (int)(src % (uint32_t)P.W), y = (int)(src / (uint32_t)P.W);
		if (y == P.y0 - 1)
			activate_item(P, 0, x / TS, 0, -1);
		else if (y == P.y1)
			activate_item(P, 0, x / TS, P.tilesY - 1, -1);
	}

	// A boundary row arrived from the neighbour stripe: keep the smaller label per column in the local halo copy and queue
	// the border tiles of the improved columns for round `round`. A label of 0 is a parked blocked tile or a source of the
	// neighbour (sources are handled from the list), never an improvement.
	template<int TS>
	__global__ void k_stripe_merge(const BigParams P, uint32_t *__restrict__ halo, const uint32_t *__restrict__ incoming, int ty, int round, uint32_t *__restrict__ improved)
	{
		for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < P.W; x += gridDim.x * blockDim.x)
		{
			const uint32_t n = incoming[x], o = halo[x];
			if (n < o)
			{
				halo[x] = n;
				if (n != 0u)
				{
					atomicAdd(improved, 1u);
					activate_item(P, 0, x / TS, ty, round - 1);
				}
			}
		}
	}

	// ---- next-hop tiles from final distances; parked blocked tiles read m again ---------------------------
	// One thread per tile of one layer (blockIdx.y = layer); neighbours come through L1/L2. A label of 0 is a source or a
	// parked blocked tile: the layer's source list tells them apart (a 2048-bit filter in shared memory answers almost
	// every query without touching the list).
	__global__ void k_big_finalize(int W, int Hl, int y0, int H, uint32_t *__restrict__ dist_base, size_t layer_stride, const uint32_t *__restrict__ wpack, const uint8_t *__restrict__ blocked, const uint32_t *__restrict__ sources_base, int nsrc, const uint32_t *__restrict__ up_row, const uint32_t *__restrict__ up_wpack, const uint32_t *__restrict__ down_row, uint32_t *__restrict__ tiles_base, size_t tiles_stride)
	{
		__shared__ uint32_t filter[64];
		const int layer = blockIdx.y;
		uint32_t *dist = dist_base + (size_t)layer * layer_stride;
		uint32_t *tiles = tiles_base ? tiles_base + (size_t)layer * tiles_stride : nullptr;
		const uint32_t *sources = sources_base + (size_t)layer * nsrc;
		for (int i = threadIdx.x; i < 64; i += blockDim.x)
			filter[i] = 0u;
		__syncthreads();
		for (int s = threadIdx.x; s < nsrc; s += blockDim.x)
		{
			const uint32_t h = (sources[s] * 2654435761u) >> 21;
			atomicOr(&filter[h >> 5], 1u << (h & 31u));
		}
		__syncthreads();
		auto is_source = [&](uint32_t g) -> bool
		{
			const uint32_t h = (g * 2654435761u) >> 21;
			if (!((filter[h >> 5] >> (h & 31u)) & 1u))
				return false;
			for (int s = 0; s < nsrc; s++)
				if (sources[s] == g)
					return true;
			return false;
		};
		const size_t total = (size_t)W * Hl;
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
		{
			const int x = (int)(i % (size_t)W);
			const int ly = (int)(i / (size_t)W);
			const int gy = ly + y0;
			const uint32_t g = (uint32_t)((size_t)gy * W + x);
			const uint32_t dv = dist[i];
			uint32_t best = MTD_INF, bestd = MTD_INF;
			if (dv == 0u)
			{
				// source (keeps 0, no next hop: paths.cpp:49) or parked blocked tile (never relaxed: paths.cpp:65-67)
				if (blocked[i] && !is_source(g))
					dist[i] = MTD_INF;
			}
			else if (dv != MTD_INF && tiles)
			{
				const uint32_t wp = wpack[i];
				auto cand = [&](uint32_t du, uint32_t w, uint32_t ug)
				{
					if (du != MTD_INF && du + MIN_EDGE + w == dv && du < bestd && (du != 0u || is_source(ug)))
					{
						bestd = du;
						best = ug;
					}
				};
				if (gy > 0)
				{
					const bool local = ly > 0;
					const uint32_t du = local ? dist[i - W] : (up_row ? up_row[x] : MTD_INF);
					const uint32_t wu = local ? wpack[i - W] : (up_wpack ? up_wpack[x] : 0u);
					cand(du, wu >> 16, g - (uint32_t)W);
				}
				if (x > 0)
					cand(dist[i - 1], wpack[i - 1] & 0xFFFFu, g - 1u);
				if (x < W - 1)
					cand(dist[i + 1], wp & 0xFFFFu, g + 1u);
				if (gy < H - 1)
				{
					const bool local = ly < Hl - 1;
					const uint32_t du = local ? dist[i + W] : (down_row ? down_row[x] : MTD_INF);
					cand(du, wp >> 16, g + (uint32_t)W);
				}
			}
			if (tiles)
				__stcs(tiles + i, best);
		}
	}
}
