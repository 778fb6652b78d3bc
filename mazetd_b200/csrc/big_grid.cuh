// big_grid.cuh -- distance / direction maps for grids that do not fit shared memory (4096^2 ... 16384^2).
//
// Directions::update (sources/map/paths.cpp:24-78) as tile-local wavefront relaxation:
//   k_edge_weights   Grid::neighborDistance of every right/down edge, once per grid (sources/map/grid.cpp:141-145)
//   k_big_init       distances := m, sources := 0 (paths.cpp:40-49), seed the tile work-list
//   k_tile_round     one global round: every active TSxTS tile is loaded with a 1-cell apron into shared memory
//                    by ONE WARP, relaxed to local convergence by the warp wavefront engine (sssp_warp.cuh),
//                    written back, and the tiles across any improved edge are queued for the next round
//   k_round_advance  flips the double-buffered work-list
//   k_directions     tiles[v] from final distances (closed form of the heap order, see sssp_warp.cuh)
// The fixpoint is unique, so the result is bit-identical to the reference whatever the schedule.
//
// Row-striped multi-GPU: each rank runs the same kernels on rows [y0, y1) of the grid; the apron rows just
// outside a stripe are read from the neighbour rank's distance array through a peer pointer (NVLink load) and
// tiles of the neighbour stripe are activated by peer stores into its stamp/list -- see BigParams::peer_*.
#pragma once

#include "common.cuh"
#include "sssp_warp.cuh"

namespace mtd
{
	struct BigParams
	{
		int W, H;               // full grid
		int y0, y1;             // rows owned by this rank (single GPU: 0, H)
		int tilesX, tilesY;     // tiles of the owned stripe
		uint32_t *dist;         // [H_local][W] rows y0..y1 (local row r = y - y0)
		const uint32_t *wpack;  // [H_local][W]  (wr - 255) | (wd - 255) << 16 ; wd of the last local row refers to row y1 (valid if y1 < H)
		const uint8_t *blocked; // [H_local][W]
		const uint32_t *sources; // device array of global tile indices
		int nsrc;
		uint32_t *list[2];      // tile work-lists
		uint32_t *count;        // [2]
		uint32_t *cursor;       // [1]
		uint32_t *stamp;        // [tiles] last round this tile was queued for
		uint32_t *total_active; // [1] running sum of processed tiles (statistics)
		int *status;
		// stripe neighbours (null on a single GPU or at the grid edge)
		const uint32_t *up_row;    // distance row y0-1 on the rank above (peer memory), [W]
		const uint32_t *down_row;  // distance row y1 on the rank below, [W]
		const uint8_t *up_blocked; // blocked row y0-1 / y1 (local copies made at upload)
		const uint8_t *down_blocked;
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
	__global__ void k_fill_u32(uint32_t *p, size_t n, uint32_t v)
	{
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
			p[i] = v;
	}

	template<int TS>
	__global__ void k_big_seed(const BigParams P, uint32_t extra_blocked_unused)
	{
		// one thread per source: distance 0 and queue its tile for round 0
		const int s = blockIdx.x * blockDim.x + threadIdx.x;
		if (s >= P.nsrc)
			return;
		const uint32_t src = P.sources[s];
		const int x = (int)(src % (uint32_t)P.W), y = (int)(src / (uint32_t)P.W);
		if (y < P.y0 || y >= P.y1)
			return;
		P.dist[(size_t)(y - P.y0) * P.W + x] = 0u;
		const uint32_t tile = (uint32_t)(((y - P.y0) / TS) * P.tilesX + x / TS);
		if (atomicMax(&P.stamp[tile], 1u) < 1u) // round r uses stamp r + 1
			P.list[0][atomicAdd(&P.count[0], 1u)] = tile;
	}

	__global__ void k_round_advance(const BigParams P, int round, uint32_t *host_flag)
	{
		// after round `round`: reset the consumed list and the cursor; publish the size of the next list
		P.count[round & 1] = 0u;
		*P.cursor = 0u;
		if (host_flag)
			*host_flag = P.count[(round + 1) & 1];
	}

	// ---- one round ------------------------------------------------------------------------------------
	template<int TS>
	struct TileSmem
	{
		static constexpr int RW = TS + 2;
		static constexpr int CELLS = RW * RW;
		static constexpr int NW = (CELLS + 31) / 32;
		static constexpr int QCAP = TS >= 64 ? 512 : 256;
		uint32_t dist[CELLS];
		uint16_t wr[CELLS];
		uint16_t wd[CELLS];
		uint32_t inq[NW];
		uint16_t q[QCAP];
		uint32_t edge_old[4][TS]; // top, bottom, left, right interior lines as loaded
	};

	__device__ __forceinline__ void activate_tile(const BigParams &P, int tx, int ty, int round)
	{
		if (tx < 0 || tx >= P.tilesX || ty < 0 || ty >= P.tilesY)
			return;
		const uint32_t tile = (uint32_t)(ty * P.tilesX + tx);
		const uint32_t want = (uint32_t)round + 2u; // stamp for round + 1
		if (atomicMax(&P.stamp[tile], want) < want)
			P.list[(round + 1) & 1][atomicAdd(&P.count[(round + 1) & 1], 1u)] = tile;
	}

	template<int TS, int WARPS>
	__global__ void __launch_bounds__(WARPS * 32) k_tile_round(const BigParams P, int round)
	{
		extern __shared__ __align__(16) unsigned char smem_raw[];
		using TSm = TileSmem<TS>;
		TSm &S = ((TSm *)smem_raw)[threadIdx.x >> 5];
		constexpr int RW = TSm::RW;
		const uint32_t lane = lane_id();
		const uint32_t lt = lanemask_lt();
		const int W = P.W;
		const int Hl = P.y1 - P.y0;
		const uint32_t *list = P.list[round & 1];
		const uint32_t count = P.count[round & 1];

		for (;;)
		{
			uint32_t item = 0;
			if (lane == 0)
				item = atomicAdd(P.cursor, 1u);
			item = __shfl_sync(FULL_MASK, item, 0);
			if (item >= count)
				break;
			const uint32_t tile = list[item];
			const int tx = (int)(tile % (uint32_t)P.tilesX), ty = (int)(tile / (uint32_t)P.tilesX);
			const int gx0 = tx * TS - 1;          // global x of region column 0
			const int ly0 = ty * TS - 1;          // local row of region row 0

			// ---- load region (interior + apron) ----
			for (int ry = 0; ry < RW; ry++)
			{
				const int ly = ly0 + ry; // local row; may be -1 or Hl (neighbour stripe or outside)
				const uint32_t *drow = nullptr;
				const uint32_t *wrow = nullptr;
				const uint8_t *brow = nullptr;
				if (ly >= 0 && ly < Hl)
				{
					drow = P.dist + (size_t)ly * W;
					wrow = P.wpack + (size_t)ly * W;
					brow = P.blocked + (size_t)ly * W;
				}
				else if (ly < 0 && P.up_row)
				{
					drow = P.up_row;
					wrow = P.up_wpack;
					brow = P.up_blocked;
				}
				else if (ly >= Hl && P.down_row)
				{
					drow = P.down_row;
					wrow = nullptr; // weights of the row below are never needed (edges are stored on the upper/left cell)
					brow = P.down_blocked;
				}
				for (int rx = (int)lane; rx < RW; rx += 32)
				{
					const int gx = gx0 + rx;
					uint32_t v = 0u, wp = 0u; // outside the grid: parked at 0 like a blocked tile
					if (drow && gx >= 0 && gx < W)
					{
						v = __ldcg(drow + gx); // L2: other SMs (and peers) update these concurrently
						if (brow[gx] && v != 0u)
							v = 0u; // blocked, not a source
						if (wrow)
							wp = wrow[gx];
					}
					S.dist[ry * RW + rx] = v;
					S.wr[ry * RW + rx] = (uint16_t)(wp & 0xFFFFu);
					S.wd[ry * RW + rx] = (uint16_t)(wp >> 16);
				}
			}
			for (int i = (int)lane; i < TSm::NW; i += 32)
				S.inq[i] = 0u;
			__syncwarp();
			// remember the interior border lines to detect changes
			for (int i = (int)lane; i < TS; i += 32)
			{
				S.edge_old[0][i] = S.dist[1 * RW + 1 + i];
				S.edge_old[1][i] = S.dist[TS * RW + 1 + i];
				S.edge_old[2][i] = S.dist[(1 + i) * RW + 1];
				S.edge_old[3][i] = S.dist[(1 + i) * RW + TS];
			}

			// ---- seeds: finite apron cells (improved neighbours) and every source inside the region.
			//      A stored 0 is either a parked (blocked / outside) cell or a source; only the source list says which,
			//      so sources are seeded explicitly -- also when they sit in the apron, because the first hop across a
			//      tile border is only ever relaxed by the tile that owns the destination cell. ----
			uint32_t nq = 0, dmin = MTD_INF;
			bool dropped0 = false;
			auto seed = [&](bool want, int ry, int rx, bool is_source)
			{
				uint32_t v = MTD_INF;
				const uint32_t c = (uint32_t)(ry * RW + rx);
				const uint32_t bit = 1u << (c & 31u);
				if (want)
				{
					v = S.dist[c];
					want = is_source ? v == 0u : (v != MTD_INF && v != 0u);
				}
				if (want)
					want = (atomicOr(&S.inq[c >> 5], bit) & bit) == 0; // dedup
				const uint32_t bal = __ballot_sync(FULL_MASK, want);
				if (want)
				{
					const uint32_t slot = nq + __popc(bal & lt);
					if (slot < (uint32_t)TSm::QCAP)
						S.q[slot] = pack_yx(ry, rx);
					dmin = min(dmin, v);
				}
				const uint32_t total = nq + (uint32_t)__popc(bal);
				if (total > (uint32_t)TSm::QCAP)
					dropped0 = true; // stays marked in inq; the engine re-collects it
				nq = min(total, (uint32_t)TSm::QCAP);
			};
			for (int i0 = 0; i0 < RW; i0 += 32)
			{
				const int i = i0 + (int)lane;
				const bool in = i >= 1 && i < RW - 1; // corners have no neighbour inside the tile
				const int ii = in ? i : 1;
				seed(in, 0, ii, false);
				seed(in, RW - 1, ii, false);
				seed(in, ii, 0, false);
				seed(in, ii, RW - 1, false);
			}
			for (int s0 = 0; s0 < P.nsrc; s0 += 32)
			{
				const int s = s0 + (int)lane;
				bool want = false;
				int ry = 0, rx = 0;
				if (s < P.nsrc)
				{
					const uint32_t src = P.sources[s];
					const int sx = (int)(src % (uint32_t)W), sy = (int)(src / (uint32_t)W) - P.y0;
					rx = sx - gx0;
					ry = sy - ly0;
					want = rx >= 0 && rx < RW && ry >= 0 && ry < RW;
					if (!want)
						rx = ry = 0;
				}
				seed(want, ry, rx, true);
			}
			dmin = warp_min_u32(dmin);
			__syncwarp();

			WarpTile<uint16_t, TSm::QCAP> t;
			t.dist = S.dist;
			t.wr = S.wr;
			t.wd = S.wd;
			t.inq = S.inq;
			t.q = S.q;
			t.W = RW;
			t.H = RW;
			const bool guard = (nq || dropped0) ? warp_relax(t, nq, dmin, dropped0) : false;
			if (guard && lane == 0)
				atomicOr(P.status, 1);
			__syncwarp();

			// ---- write back improved interior cells; 0 = parked/source: global already holds the right value ----
			for (int ry = 1; ry <= TS; ry++)
			{
				const int ly = ly0 + ry;
				if (ly >= Hl)
					break;
				uint32_t *drow = P.dist + (size_t)ly * W;
				for (int rx = 1 + (int)lane; rx <= TS; rx += 32)
				{
					const int gx = gx0 + rx;
					const uint32_t v = S.dist[ry * RW + rx];
					if (gx < W && v != 0u && v != MTD_INF)
						__stcg(drow + gx, v);
				}
			}
			// ---- activate neighbours across improved border lines ----
			bool ch[4] = { false, false, false, false };
			for (int i = (int)lane; i < TS; i += 32)
			{
				ch[0] |= S.dist[1 * RW + 1 + i] != S.edge_old[0][i];
				ch[1] |= S.dist[TS * RW + 1 + i] != S.edge_old[1][i];
				ch[2] |= S.dist[(1 + i) * RW + 1] != S.edge_old[2][i];
				ch[3] |= S.dist[(1 + i) * RW + TS] != S.edge_old[3][i];
			}
			const bool c0 = __any_sync(FULL_MASK, ch[0]), c1 = __any_sync(FULL_MASK, ch[1]);
			const bool c2 = __any_sync(FULL_MASK, ch[2]), c3 = __any_sync(FULL_MASK, ch[3]);
			if (lane == 0)
			{
				__threadfence(); // distances are visible before the neighbour can be scheduled
				if (c0)
				{
					if (ty > 0)
						activate_tile(P, tx, ty - 1, round);
					else if (P.peer_flag_up)
						P.peer_flag_up[tx] = 1u;
				}
				if (c1)
				{
					if (ty < P.tilesY - 1)
						activate_tile(P, tx, ty + 1, round);
					else if (P.peer_flag_down)
						P.peer_flag_down[tx] = 1u;
				}
				if (c2)
					activate_tile(P, tx - 1, ty, round);
				if (c3)
					activate_tile(P, tx + 1, ty, round);
				atomicAdd(P.total_active, 1u);
			}
			__syncwarp();
		}
	}

	// ---- next-hop tiles from final distances -----------------------------------------------------------
	// one thread per tile; neighbours come through L1/L2 (each distance is read by 5 threads of nearby warps)
	__global__ void k_directions(int W, int Hl, int y0, int H, const uint32_t *__restrict__ dist, const uint32_t *__restrict__ wpack, const uint32_t *__restrict__ up_row, const uint32_t *__restrict__ up_wpack, const uint32_t *__restrict__ down_row, uint32_t *__restrict__ tiles)
	{
		const size_t total = (size_t)W * Hl;
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
		{
			const int x = (int)(i % (size_t)W);
			const int ly = (int)(i / (size_t)W);
			const int gy = ly + y0;
			const uint32_t dv = dist[i];
			uint32_t best = MTD_INF, bestd = MTD_INF;
			if (dv != MTD_INF && dv != 0u)
			{
				const uint32_t wp = wpack[i];
				if (gy > 0)
				{
					const bool local = ly > 0;
					const uint32_t du = local ? dist[i - W] : (up_row ? up_row[x] : MTD_INF);
					const uint32_t wu = local ? wpack[i - W] : (up_wpack ? up_wpack[x] : 0u);
					if (du != MTD_INF && du + MIN_EDGE + (wu >> 16) == dv && du < bestd)
					{
						bestd = du;
						best = (uint32_t)((size_t)(gy - 1) * W + x);
					}
				}
				if (x > 0)
				{
					const uint32_t du = dist[i - 1];
					if (du != MTD_INF && du + MIN_EDGE + (wpack[i - 1] & 0xFFFFu) == dv && du < bestd)
					{
						bestd = du;
						best = (uint32_t)((size_t)gy * W + x - 1);
					}
				}
				if (x < W - 1)
				{
					const uint32_t du = dist[i + 1];
					if (du != MTD_INF && du + MIN_EDGE + (wp & 0xFFFFu) == dv && du < bestd)
					{
						bestd = du;
						best = (uint32_t)((size_t)gy * W + x + 1);
					}
				}
				if (gy < H - 1)
				{
					const bool local = ly < Hl - 1;
					const uint32_t du = local ? dist[i + W] : (down_row ? down_row[x] : MTD_INF);
					if (du != MTD_INF && du + MIN_EDGE + (wp >> 16) == dv && du < bestd)
					{
						bestd = du;
						best = (uint32_t)((size_t)(gy + 1) * W + x);
					}
				}
			}
			__stcs(tiles + i, best);
		}
	}
}
