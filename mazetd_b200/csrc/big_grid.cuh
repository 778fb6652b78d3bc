// big_grid.cuh -- distance / direction maps for grids that do not fit shared memory (4096^2 ... 16384^2).
//
// Directions::update (sources/map/paths.cpp:24-78) as tile-local wavefront relaxation, all layers of one
// Waypoints::update in flight together (a work item is one (layer, tile) pair):
//   k_edge_weights   Grid::neighborDistance of every right/down edge, once per grid (sources/map/grid.cpp:141-145)
//   k_big_init       distances := m, blocked tiles PARKED at 0 (paths.cpp:40-43, 65-67), work-list cleared
//   k_big_seed       sources := 0 (paths.cpp:48-49), their tiles queued for round 0
//   k_tile_rounds    (tile_warp.cuh) persistent cooperative kernel running ALL global rounds: every active 32x32 tile is
//                    staged with a 1-cell apron in shared memory by one warp (16-byte cp.async.cg rows, L2-coherent),
//                    relaxed to local convergence by a warp-synchronous bucketed wavefront, written back with 16-byte
//                    stores, and the neighbour tiles it can still improve are queued for the next round; rounds are
//                    separated by one grid-wide barrier, the three work-lists rotate
//   k_big_finalize   tiles[v] from final distances (closed form of the heap order, see sssp_warp.cuh); parked blocked
//                    tiles read m again
// The fixpoint is unique, so the result is bit-identical to the reference whatever the schedule.
//
// Frame layout in shared memory: (TS + 2) rows of PITCH = TS + 8 words; grid column gx0 + rx sits at word rx + 3, so the
// 64 interior words of a row start 16-byte aligned. The apron is read once (each finite apron label relaxes its one
// interior neighbour) and then parked at 0 like blocked cells: the four neighbours of an interior cell are p-1, p+1,
// p-PITCH, p+PITCH with no bounds or flag test, and nothing is ever relaxed outward.
//
// Row-striped multi-GPU (one rank per stripe): each rank runs the same kernels on rows [y0, y1) of the grid; the apron rows
// just outside a stripe are read from LOCAL copies of the neighbour stripes' boundary rows (up_row / down_row), which the
// host protocol (mazetd_b200/striped.py) refreshes between relaxation sweeps and k_stripe_merge folds in.
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
		uint32_t *list[3];      // work-lists of items = layer * ntiles + tile; round r consumes list[r % 3], fills list[(r + 1) % 3]
		uint32_t *count;        // [3]
		uint32_t *cursor;       // [3]
		uint32_t *stamp;        // [nlayers * ntiles] last round this item was queued for
		uint32_t *key;          // [nlayers * ntiles] smallest label offered to the item since its last visit (MTD_INF: nothing)
		uint32_t *minkey;       // [3] smallest key on each work-list
		uint32_t tile_delta;    // a round visits the items whose key is below (smallest key on the list) + tile_delta; the rest wait
		uint32_t *total_active; // [1] running sum of processed items (statistics)
		int *status;
		int aligned16;          // W % 4 == 0 and every base pointer 16-byte aligned: rows move as 16-byte vectors
		// stripe neighbours (null on a single GPU or at the grid edge); nlayers == 1 when striped
		const uint32_t *up_row;    // local copy of distance row y0-1 (owned by the rank above), [W]
		const uint32_t *down_row;  // local copy of distance row y1 (owned by the rank below), [W]
		const uint32_t *up_wpack;  // wpack row y0-1 (for wd of the edge (y0-1) -> y0), local copy
		// peer-coupled stripes (ctl0 != null): the stripes of all ranks advance round by round inside ONE kernel run. A visit
		// of a border tile stores its boundary row straight into the neighbour rank's halo copy (up_row / down_row above then
		// point into this rank's peer block, written by the neighbours over NVLink) and posts the offered key in the
		// neighbour's act array. Between rounds every rank stores a phase tag (then its queue length and smallest key) into
		// ITS slot of EVERY rank's block and waits on its own block until all slots carry the tag: one NVLink store latency
		// per phase, no remote polling, no remote atomics.
		uint32_t *peer_up_halo;    // the rank above: its copy of my first row [W]           (null: no rank above)
		uint32_t *peer_up_act;     // the rank above: keys offered to its last tile row [tilesX]
		uint32_t *peer_down_halo;  // the rank below: its copy of my last row [W]
		uint32_t *peer_down_act;   // the rank below: keys offered to its first tile row [tilesX]
		uint32_t *act_up;          // mine: keys the rank above offered to my first tile row (MTD_INF: none)
		uint32_t *act_down;        // mine: keys the rank below offered to my last tile row
		uint32_t *ctl0;            // non-null = peer-coupled mode (rank 0's block)
		uint32_t *ctl;             // my block (PEER_* words below)
		uint32_t *peers[16];       // every rank's block, mine included
		int world, rank;
		int soft_grid;             // the grid is smaller than the GPU and was launched plainly (several stripes share one GPU in
		                           // tests): CTAs meet on a counter (count[11]) instead of grid.sync()
	};

	// words at the head of a peer block
	constexpr int PEER_MAX_WORLD = 16;
	constexpr int PEER_GPENDING = 12; // local: work items queued on all ranks for the coming round
	constexpr int PEER_EPOCH = 13;    // local: phases this rank has been through since the block was created
	constexpr int PEER_SLOT_A = 16;   // [16] phase tag of rank j ("everything I wrote into your block this round has landed")
	constexpr int PEER_SLOT_B = 32;   // [16] x 64 bit: smallest key << 32 | queue length (24 bits) << 8 | phase tag (8 bits) of rank j
	// peer-coupled DATAFLOW kernel (k_tile_async with ctl0 != null): no rounds, so instead of the two phases above
	constexpr int PEER_AREQ = 64;     // written by rank 0 into every block: number of the termination wave it is collecting
	constexpr int PEER_ASTOP = 65;    // written by rank 0 into every block: the run is over
	constexpr int PEER_AWM = 80;      // [16] smallest pending key of rank j (written by rank j into every block): the priority window is global
	constexpr int PEER_AREPLY = 96;   // [16] x 64 bit, read in rank 0's block: wave tag (8) | idle (1) | border offers made (27) | border offers consumed (28) of rank j
	constexpr int PEER_CTL_WORDS = 128;

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

	// four tiles per thread (n % 4 == 0, 16-byte aligned bases)
	__global__ void k_big_init4(uint32_t *__restrict__ dist, size_t layer_stride, int nlayers, size_t n4, const uint8_t *__restrict__ blocked)
	{
		for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (size_t)gridDim.x * blockDim.x)
		{
			const uchar4 b = reinterpret_cast<const uchar4 *>(blocked)[q];
			const uint4 v = make_uint4(b.x ? 0u : MTD_INF, b.y ? 0u : MTD_INF, b.z ? 0u : MTD_INF, b.w ? 0u : MTD_INF);
			for (int l = 0; l < nlayers; l++)
				__stcg(reinterpret_cast<uint4 *>(dist + (size_t)l * layer_stride) + q, v);
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
		P.key[item] = 0u;
		P.minkey[0] = 0u;
		if (atomicMax(&P.stamp[item], 1u) < 1u) // round r uses stamp r + 1
			P.list[0][atomicAdd(&P.count[0], 1u)] = item;
	}

#ifdef MTD_TIMING
	// developer probe (never in the shipped build): cycles per phase of a tile visit, summed over all visits
	__device__ unsigned long long g_big_timing[16];
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

	// Queue item (layer, tx, ty) for round + 1 because a label `key` is on offer to one of its cells.
	__device__ __forceinline__ void activate_item(const BigParams &P, int layer, int tx, int ty, int round, uint32_t key)
	{
		if (tx < 0 || tx >= P.tilesX || ty < 0 || ty >= P.tilesY)
			return;
		const uint32_t item = (uint32_t)layer * (uint32_t)(P.tilesX * P.tilesY) + (uint32_t)(ty * P.tilesX + tx);
		const uint32_t want = (uint32_t)round + 2u; // stamp for round + 1
		const int slot = (round + 1) % 3;
		atomicMin(&P.key[item], key);
		atomicMin(&P.minkey[slot], key);
		if (atomicMax(&P.stamp[item], want) < want)
			P.list[slot][atomicAdd(&P.count[slot], 1u)] = item;
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
			activate_item(P, 0, x / TS, 0, -1, 0u);
		else if (y == P.y1)
			activate_item(P, 0, x / TS, P.tilesY - 1, -1, 0u);
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
					activate_item(P, 0, x / TS, ty, round - 1, n);
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

	// The same pass, four tiles of one row per thread (W % 4 == 0, 16-byte aligned bases): the own row, the rows above and
	// below and the two weight rows move as 16-byte vectors, so the kernel issues a quarter of the load instructions and
	// streams at HBM speed instead of queueing in the load/store unit.
	// The next-hop pass of k_big_finalize for widths that are a multiple of 4, four tiles of a row per thread, with the labels and
	// weights of a TW x TH block of cells (plus one row above / below and one quad left / right) staged in shared memory by
	// 16-byte loads: every label is read from global memory once instead of three times (own row, as "down" of the row
	// above, as "up" of the row below), the loads of a thread are independent of each other, and a quad with parked cells
	// goes back as one 16-byte store. HBM-bound streaming: 4 B labels + 4 B weights +
	// 1 B blocked in, 4 B next hops + (most quads) 4 B labels out per tile.
	constexpr int FIN_TW = 256, FIN_TH = 16;
	__global__ void __launch_bounds__(256) k_big_finalize_tiled(int W, int Hl, int y0, int H, uint32_t *__restrict__ dist_base, size_t layer_stride, const uint32_t *__restrict__ wpack, const uint8_t *__restrict__ blocked, const uint32_t *__restrict__ sources_base, int nsrc, const uint32_t *__restrict__ up_row, const uint32_t *__restrict__ up_wpack, const uint32_t *__restrict__ down_row, uint32_t *__restrict__ tiles_base, size_t tiles_stride, int ntx)
	{
		constexpr int QD = FIN_TW / 4 + 2, QW = FIN_TW / 4 + 1;
		__shared__ uint4 sd[FIN_TH + 2][QD]; // labels: row r = local row ly0 - 1 + r, column c = the quad at x0 - 4 + 4 c
		__shared__ uint4 sw[FIN_TH + 1][QW]; // packed weights: row r = local row ly0 - 1 + r, column c = the quad at x0 - 4 + 4 c
		__shared__ uint32_t filter[64];
		const int layer = blockIdx.y;
		uint32_t *dist = dist_base + (size_t)layer * layer_stride;
		uint32_t *tiles = tiles_base ? tiles_base + (size_t)layer * tiles_stride : nullptr;
		const uint32_t *sources = sources_base + (size_t)layer * nsrc;
		const int x0 = (int)(blockIdx.x % (unsigned)ntx) * FIN_TW, ly0 = (int)(blockIdx.x / (unsigned)ntx) * FIN_TH;
		const uint4 inf4 = make_uint4(MTD_INF, MTD_INF, MTD_INF, MTD_INF);
		const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
		for (int i = threadIdx.x; i < 64; i += blockDim.x)
			filter[i] = 0u;
		for (int idx = threadIdx.x; idx < (FIN_TH + 2) * QD; idx += blockDim.x)
		{
			const int r = idx / QD, c = idx - r * QD;
			const int ly = ly0 - 1 + r, x = x0 - 4 + 4 * c, gy = ly + y0;
			uint4 v = inf4; // outside the grid: no predecessor there
			if (x >= 0 && x < W && gy >= 0 && gy < H)
			{
				if (ly >= 0 && ly < Hl)
					v = *reinterpret_cast<const uint4 *>(dist + (size_t)ly * W + x);
				else if (ly == -1 && up_row)
					v = *reinterpret_cast<const uint4 *>(up_row + x);
				else if (ly == Hl && down_row)
					v = *reinterpret_cast<const uint4 *>(down_row + x);
			}
			sd[r][c] = v;
		}
		for (int idx = threadIdx.x; idx < (FIN_TH + 1) * QW; idx += blockDim.x)
		{
			const int r = idx / QW, c = idx - r * QW;
			const int ly = ly0 - 1 + r, x = x0 - 4 + 4 * c;
			uint4 v = zero4;
			if (x >= 0 && x < W && ly + y0 >= 0)
			{
				if (ly >= 0 && ly < Hl)
					v = *reinterpret_cast<const uint4 *>(wpack + (size_t)ly * W + x);
				else if (ly == -1 && up_row && up_wpack)
					v = *reinterpret_cast<const uint4 *>(up_wpack + x);
			}
			sw[r][c] = v;
		}
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
		for (int idx = threadIdx.x; idx < FIN_TH * (FIN_TW / 4); idx += blockDim.x)
		{
			const int r = idx / (FIN_TW / 4), c = idx - r * (FIN_TW / 4);
			const int ly = ly0 + r, x = x0 + 4 * c;
			if (ly >= Hl || x >= W)
				continue;
			const size_t i = (size_t)ly * W + x;
			const uint32_t g = (uint32_t)((size_t)(ly + y0) * W + x);
			const uint4 dq = sd[r + 1][c + 1];
			uint32_t dv[4] = { dq.x, dq.y, dq.z, dq.w };
			uint32_t best[4] = { MTD_INF, MTD_INF, MTD_INF, MTD_INF };
			bool any_zero = false, any_live = false;
#pragma unroll
			for (int j = 0; j < 4; j++)
			{
				any_zero |= dv[j] == 0u;
				any_live |= dv[j] != 0u && dv[j] != MTD_INF;
			}
			if (any_live && tiles)
			{
				const uint4 du4 = sd[r][c + 1], dd4 = sd[r + 2][c + 1], wu4 = sw[r][c + 1], wq = sw[r + 1][c + 1];
				const uint32_t wp[4] = { wq.x, wq.y, wq.z, wq.w };
				const uint32_t du[4] = { du4.x, du4.y, du4.z, du4.w };
				const uint32_t dd[4] = { dd4.x, dd4.y, dd4.z, dd4.w };
				const uint32_t wu[4] = { wu4.x, wu4.y, wu4.z, wu4.w };
				const uint32_t dleft = sd[r + 1][c].w, wleft = sw[r + 1][c].w & 0xFFFFu, dright = sd[r + 1][c + 2].x;
#pragma unroll
				for (int j = 0; j < 4; j++)
				{
					const uint32_t v = dv[j];
					if (v == 0u || v == MTD_INF)
						continue;
					uint32_t bestd = MTD_INF;
					auto cand = [&](uint32_t dn, uint32_t w, uint32_t ug)
					{
						if (dn != MTD_INF && dn + MIN_EDGE + w == v && dn < bestd && (dn != 0u || is_source(ug)))
						{
							bestd = dn;
							best[j] = ug;
						}
					};
					// ascending tile index, so equal labels resolve to the smaller index: up, left, right, down
					cand(du[j], wu[j] >> 16, g + (uint32_t)j - (uint32_t)W);
					cand(j > 0 ? dv[j - 1] : dleft, j > 0 ? wp[j - 1] & 0xFFFFu : wleft, g + (uint32_t)j - 1u);
					cand(j < 3 ? dv[j + 1] : dright, wp[j] & 0xFFFFu, g + (uint32_t)j + 1u);
					cand(dd[j], wp[j] >> 16, g + (uint32_t)j + (uint32_t)W);
				}
			}
			if (any_zero)
			{
				// source (keeps 0, no next hop: paths.cpp:49) or parked blocked tile (never relaxed: paths.cpp:65-67). Blocks
				// that read these cells as neighbours meanwhile treat 0 (of a tile that is no source) and m alike.
				const uchar4 bq = *reinterpret_cast<const uchar4 *>(blocked + i);
				const uint8_t b[4] = { bq.x, bq.y, bq.z, bq.w };
				bool changed = false;
#pragma unroll
				for (int j = 0; j < 4; j++)
					if (dv[j] == 0u && b[j] && !is_source(g + (uint32_t)j))
					{
						dv[j] = MTD_INF;
						changed = true;
					}
				if (changed)
					*reinterpret_cast<uint4 *>(dist + i) = make_uint4(dv[0], dv[1], dv[2], dv[3]);
			}
			if (tiles)
				__stcs(reinterpret_cast<uint4 *>(tiles + i), make_uint4(best[0], best[1], best[2], best[3]));
		}
	}

}

#include "tile_warp.cuh"
