// small_maps.cuh -- kernels for maps that fit one CTA's shared memory (the game's 101x101 map and batches of
// them): the procedural grid generator and the fused Waypoints::update.
//
//   k_mapgen_grid      newProcedural + newGrid + spawn picking      (sources/map/procedural.cpp:13-70,
//                                                                    sources/map/grid.cpp:12-198,
//                                                                    sources/map/paths.cpp:173-196)
//   k_waypoints_small  Waypoints::update: n Dijkstra layers, next-hop tiles, full-path walks, aggregates
//                                                                   (sources/map/paths.cpp:24-139)
// One CTA owns one map at a time and loops over maps (persistent grid sized from the SM count).
#pragma once

#include "common.cuh"
#include "noise.cuh"
#include "sssp_warp.cuh"

namespace mtd
{
	enum : int
	{
		ST_GUARD = 1,      // a distance reached the uint32 guard band
		ST_NO_VALID = 2,   // generated map has no valid tile (reference would spin forever in findSpawnPoint)
		ST_PATH_TRUNC = 4, // a full path exceeded fullpath_cap (length is still exact)
	};

	// =============================================================================================
	// procedural grid generator
	// =============================================================================================
	struct MapgenGridParams
	{
		int W, H, ox, oy, radius, nwp;
		uint64_t s1, s2;
		int first, nmaps;
		int16_t *elev;      // [nmaps][T]
		uint16_t *flags;    // [nmaps][T]
		uint32_t *wp_tiles; // [nmaps][nwp]
		int *status;
	};

	// shared memory: parent u32[T] | elev i16[T] | flags u16[T] | grad float2[128]
	__host__ __device__ inline size_t mapgen_grid_smem(int T)
	{
		return (size_t)T * 4 + (size_t)((T + 1) & ~1) * 2 * 2 + 128 * sizeof(float2) + 64;
	}

	__device__ __forceinline__ uint32_t uf_find(uint32_t *parent, uint32_t i)
	{
		// path-halving-free find: parents only ever decrease, so a plain chase terminates
		uint32_t p = parent[i];
		while (p != i)
		{
			i = p;
			p = parent[i];
		}
		return i;
	}

	__device__ __forceinline__ void uf_union(uint32_t *parent, uint32_t a, uint32_t b)
	{
		// lock-free union by minimum index (root = smallest tile index of the component)
		for (;;)
		{
			a = uf_find(parent, a);
			b = uf_find(parent, b);
			if (a == b)
				return;
			if (a > b)
			{
				const uint32_t t = a;
				a = b;
				b = t;
			}
			const uint32_t old = atomicMin(&parent[b], a);
			if (old == b)
				return;
			b = old; // somebody re-parented b meanwhile: merge that root too
		}
	}

	__global__ void __launch_bounds__(512, 1) k_mapgen_grid(const MapgenGridParams P)
	{
		extern __shared__ __align__(16) unsigned char smem_raw[];
		const int W = P.W, H = P.H, T = W * H;
		const int Tp = (T + 1) & ~1;
		uint32_t *parent = (uint32_t *)smem_raw;
		int16_t *elev = (int16_t *)(parent + T);
		uint16_t *flags = (uint16_t *)(elev + Tp);
		float2 *grad = (float2 *)(((uintptr_t)(flags + Tp) + 15) & ~(uintptr_t)15);
		__shared__ noise::ElevationParams ep;
		__shared__ Rng rng;
		__shared__ unsigned long long best_key;

		noise::fill_gradients(grad);

		for (int mapi = blockIdx.x; mapi < P.nmaps; mapi += gridDim.x)
		{
			__syncthreads();
			if (threadIdx.x == 0)
			{
				// RNG draw order = member initialisation order of ProceduralImpl (procedural.cpp:13,24,33,43,52,61)
				Rng g;
				g.s0 = P.s1 + (uint64_t)(P.first + mapi);
				g.s1 = P.s2;
				const float r1 = g.chance();
				const float r2 = g.chance();
				ep.off_x = __fmul_rn(r1, 1000.0f);
				ep.off_y = __fmul_rn(r2, 1000.0f);
				ep.seed = (int)(uint32_t)g.next();
				g.next(); // water, grass, dirt, roughness seeds: drawn to keep the stream aligned
				g.next();
				g.next();
				g.next();
				ep.bounding = noise::fractal_bounding(4, 0.55f);
				rng = g;
				best_key = 0ull;
			}
			__syncthreads();

			// ---- elevations: grid.cpp:153-159 ----
			const noise::ElevationParams lp = ep;
			for (int i = threadIdx.x; i < T; i += blockDim.x)
			{
				const int x = i % W, y = i / W;
				float e = noise::elevation(grad, lp, (float)(x - P.ox), (float)(y - P.oy));
				e = e > -100.0f ? e : -100.0f;
				e = e < 100.0f ? e : 100.0f;
				elev[i] = rtos16(e);
			}
			__syncthreads();

			// ---- biome flags, slope, radius: grid.cpp:160-183 (thresholds are rtos16(-7), (1), (6), (0.35)) ----
			for (int i = threadIdx.x; i < T; i += blockDim.x)
			{
				const int x = i % W, y = i / W;
				const int16_t me = elev[i];
				uint16_t f = me < (int16_t)-1785 ? F_WATER : me < (int16_t)255 ? F_SUN : me < (int16_t)1530 ? F_WIND : F_SNOW;
				int16_t r = 0; // maxSlope, grid.cpp:12-27 (int16 wrap semantics kept)
				if (x > 0)
				{
					const int16_t d = (int16_t)(me - elev[i - 1]);
					const int16_t a = d < 0 ? (int16_t)-d : d;
					r = r > a ? r : a;
				}
				if (x < W - 1)
				{
					const int16_t d = (int16_t)(me - elev[i + 1]);
					const int16_t a = d < 0 ? (int16_t)-d : d;
					r = r > a ? r : a;
				}
				if (y > 0)
				{
					const int16_t d = (int16_t)(me - elev[i - W]);
					const int16_t a = d < 0 ? (int16_t)-d : d;
					r = r > a ? r : a;
				}
				if (y < H - 1)
				{
					const int16_t d = (int16_t)(me - elev[i + W]);
					const int16_t a = d < 0 ? (int16_t)-d : d;
					r = r > a ? r : a;
				}
				if (r > (int16_t)89)
					f = F_INVALID;
				if (P.radius >= 0)
				{
					const float px = (float)(x - P.ox), py = (float)(y - P.oy);
					const float l2 = __fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py));
					if (l2 > (float)(P.radius * P.radius))
						f = F_INVALID;
				}
				flags[i] = f;
				parent[i] = (uint32_t)i;
			}
			__syncthreads();

			// ---- fillUnreachable: grid.cpp:29-78 as union-find (root = lowest tile index = discovery order) ----
			for (int i = threadIdx.x; i < T; i += blockDim.x)
			{
				if (flags[i] == F_INVALID)
					continue;
				const int x = i % W, y = i / W;
				if (x < W - 1 && flags[i + 1] != F_INVALID)
					uf_union(parent, (uint32_t)i, (uint32_t)i + 1);
				if (y < H - 1 && flags[i + W] != F_INVALID)
					uf_union(parent, (uint32_t)i, (uint32_t)(i + W));
			}
			__syncthreads();
			// flatten, then count members in the upper 16 bits of the root's word (T < 65536)
			for (int i = threadIdx.x; i < T; i += blockDim.x)
				if (flags[i] != F_INVALID)
				{
					const uint32_t r = uf_find(parent, (uint32_t)i);
					if (r != (uint32_t)i)
						parent[i] = r; // benign: only shortens chains, roots are untouched
				}
			__syncthreads();
			for (int i = threadIdx.x; i < T; i += blockDim.x)
				if (flags[i] != F_INVALID)
					atomicAdd(&parent[parent[i] & 0xFFFFu], 1u << 16);
			__syncthreads();
			// largest count; ties -> first discovered = smallest root (grid.cpp:68-72 uses strict '>')
			{
				unsigned long long key = 0ull;
				for (int i = threadIdx.x; i < T; i += blockDim.x)
					if (flags[i] != F_INVALID && (parent[i] & 0xFFFFu) == (uint32_t)i)
					{
						const unsigned long long k = ((unsigned long long)(parent[i] >> 16) << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)i);
						key = k > key ? k : key;
					}
				for (int o = 16; o > 0; o >>= 1)
				{
					const unsigned long long other = __shfl_xor_sync(FULL_MASK, key, o);
					key = other > key ? other : key;
				}
				if (lane_id() == 0 && key)
					atomicMax(&best_key, key);
			}
			__syncthreads();
			const bool any_valid = best_key != 0ull;
			const uint32_t best_root = 0xFFFFFFFFu - (uint32_t)(best_key & 0xFFFFFFFFull);
			for (int i = threadIdx.x; i < T; i += blockDim.x)
				if (flags[i] == F_INVALID || !any_valid || (parent[i] & 0xFFFFu) != best_root)
					flags[i] = F_INVALID;
			__syncthreads();

			// ---- findBorder + Mana: grid.cpp:80-97, 186-196 (border bits only land on invalid tiles, so in-place is safe) ----
			for (int i = threadIdx.x; i < T; i += blockDim.x)
			{
				const int x = i % W, y = i / W;
				uint16_t f = flags[i];
				if (f & F_INVALID)
				{
					bool b = false;
					b |= x > 0 && !(flags[i - 1] & F_INVALID);
					b |= x < W - 1 && !(flags[i + 1] & F_INVALID);
					b |= y > 0 && !(flags[i - W] & F_INVALID);
					b |= y < H - 1 && !(flags[i + W] & F_INVALID);
					if (b)
						f |= F_BORDER;
				}
				else
					f |= F_MANA;
				flags[i] = f;
			}
			__syncthreads();

			// ---- newWaypoints spawn picking: paths.cpp:173-183, 190-196 (inherently sequential in the RNG stream) ----
			if (threadIdx.x == 0)
			{
				Rng g = rng;
				for (int k = 0; k < P.nwp; k++)
				{
					uint32_t t = MTD_INF;
					if (any_valid)
					{
						for (;;)
						{
							t = g.range_u32(0u, (uint32_t)T);
							if (!(flags[t] & (F_INVALID | F_WAYPOINT)))
								break;
						}
						flags[t] |= F_WAYPOINT;
					}
					P.wp_tiles[(size_t)mapi * P.nwp + k] = t;
				}
				if (!any_valid)
					atomicOr(P.status, ST_NO_VALID);
			}
			__syncthreads();

			// ---- write out (coalesced 4-byte words; T even-padded arrays are 4-byte aligned per map only if T is even,
			//      so use 2-byte stores: 40 KB per map, negligible next to the layers) ----
			int16_t *ge = P.elev + (size_t)mapi * T;
			uint16_t *gf = P.flags + (size_t)mapi * T;
			for (int i = threadIdx.x; i < T; i += blockDim.x)
			{
				ge[i] = elev[i];
				gf[i] = flags[i];
			}
		}
	}

	// =============================================================================================
	// Waypoints::update for small maps
	// =============================================================================================
	constexpr int SMALL_QCAP = 1024;
	constexpr int SMALL_MAX_RESIDENT = 4; // layers concurrently in shared memory (one warp each)
	constexpr int SMALL_THREADS = 512;
	constexpr int MTD_MAXSRC_SMALL = 32;

	struct WaypointsSmallParams
	{
		int W, H;
		int n;       // layers per map
		int nmaps;
		int nsrc;    // sources per layer: 1 = reference; > 1 = multi-source extension (then n must be 1)
		const int16_t *elev;    // [nmaps][T]
		const uint16_t *flags;  // [nmaps][T]
		const uint32_t *wp;     // [nmaps][n] (or [nsrc] when nsrc > 1)
		uint32_t extra_blocked; // MTD_M = none
		uint32_t *tiles;        // [nmaps][n][T] or null
		uint32_t *dist;         // [nmaps][n][T]; required when n > SMALL_MAX_RESIDENT and paths are wanted
		uint32_t *fullpath;     // [nmaps][n][cap] or null
		uint32_t cap;
		uint32_t *fullpath_len;  // [nmaps][n] or null: null skips walks + aggregates
		uint32_t *full_distance; // [nmaps][n]
		uint32_t *agg;           // [nmaps][3]
		int *status;
	};

	// shared: dist u32[R][T] | wr u16[T] | wd u16[T] | inq u32[R][NW] | q u16[R][QCAP] | D u32[64] | fd u32[8] ...
	__host__ __device__ inline size_t waypoints_small_smem(int T, int resident)
	{
		const size_t Tp = (size_t)(T + 1) & ~(size_t)1;
		const size_t nw = (size_t)(T + 31) >> 5;
		return (size_t)resident * T * 4 + Tp * 2 * 2 + (size_t)resident * nw * 4 + (size_t)resident * SMALL_QCAP * 2 + 64 * 4 + 64;
	}

	__global__ void __launch_bounds__(SMALL_THREADS, 1) k_waypoints_small(const WaypointsSmallParams P)
	{
		extern __shared__ __align__(16) unsigned char smem_raw[];
		const int W = P.W, H = P.H, T = W * H, n = P.n;
		const int Tp = (T + 1) & ~1;
		const int NW = (T + 31) >> 5;
		const int R = n < SMALL_MAX_RESIDENT ? n : SMALL_MAX_RESIDENT;
		uint32_t *sdist = (uint32_t *)smem_raw;
		uint16_t *wr = (uint16_t *)(sdist + (size_t)R * T);
		uint16_t *wd = wr + Tp;
		uint32_t *inq = (uint32_t *)(wd + Tp);
		uint16_t *q = (uint16_t *)(inq + (size_t)R * NW);
		uint32_t *Dtab = (uint32_t *)(q + (size_t)R * SMALL_QCAP);
		__shared__ uint32_t s_fd[FIND_MAXN];
		__shared__ uint32_t s_wp[MTD_MAXSRC_SMALL];
		__shared__ int s_guard;

		const int warp = threadIdx.x >> 5;
		const uint32_t lane = lane_id();
		const bool resident_all = n <= SMALL_MAX_RESIDENT;
		const int nsrc_layer = P.nsrc > 1 ? P.nsrc : 1;

		for (int mapi = blockIdx.x; mapi < P.nmaps; mapi += gridDim.x)
		{
			const int16_t *ge = P.elev + (size_t)mapi * T;
			const uint16_t *gf = P.flags + (size_t)mapi * T;
			const uint32_t *gwp = P.wp + (size_t)mapi * (P.nsrc > 1 ? P.nsrc : n);
			uint32_t *gtiles = P.tiles ? P.tiles + (size_t)mapi * n * T : nullptr;
			uint32_t *gdist = P.dist ? P.dist + (size_t)mapi * n * T : nullptr;

			__syncthreads();
			if (threadIdx.x == 0)
				s_guard = 0;
			for (int i = threadIdx.x; i < (P.nsrc > 1 ? P.nsrc : n) && i < MTD_MAXSRC_SMALL; i += blockDim.x)
				s_wp[i] = gwp[i];

			// ---- edge weights: Grid::neighborDistance of the right and down edge of every tile (grid.cpp:141-145) ----
			for (int i = threadIdx.x; i < T; i += blockDim.x)
			{
				const int x = i % W, y = i / W;
				const int16_t e = ge[i];
				wr[i] = x < W - 1 ? (uint16_t)(edge_weight(e, ge[i + 1]) - MIN_EDGE) : (uint16_t)0;
				wd[i] = y < H - 1 ? (uint16_t)(edge_weight(e, ge[i + W]) - MIN_EDGE) : (uint16_t)0;
			}
			__syncthreads();

			for (int l0 = 0; l0 < n; l0 += R)
			{
				const int lb = (n - l0) < R ? (n - l0) : R; // layers in this batch
				// ---- init: paths.cpp:40-43; blocked tiles are parked at 0 during the run ----
				for (int i = threadIdx.x; i < T; i += blockDim.x)
				{
					const bool blocked = (gf[i] & F_BLOCKING) != 0 || (uint32_t)i == P.extra_blocked;
					const uint32_t v = blocked ? 0u : MTD_INF;
					for (int r = 0; r < lb; r++)
						sdist[(size_t)r * T + i] = v;
				}
				for (int i = threadIdx.x; i < lb * NW; i += blockDim.x)
					inq[i] = 0u;
				__syncthreads();

				// ---- one warp per layer: wavefront relaxation ----
				if (warp < lb)
				{
					WarpTile<uint16_t, SMALL_QCAP> t;
					t.dist = sdist + (size_t)warp * T;
					t.wr = wr;
					t.wd = wd;
					t.inq = inq + (size_t)warp * NW;
					t.q = q + (size_t)warp * SMALL_QCAP;
					t.W = W;
					t.H = H;
					uint32_t nseed = 0;
					if (lane == 0)
					{
						for (int s = 0; s < nsrc_layer; s++)
						{
							const uint32_t src = P.nsrc > 1 ? s_wp[s] : s_wp[l0 + warp];
							const uint32_t bit = 1u << (src & 31u);
							t.dist[src] = 0u; // paths.cpp:48-49
							if (!(t.inq[src >> 5] & bit))
							{
								t.inq[src >> 5] |= bit;
								t.q[nseed++] = pack_yx((int)(src / (uint32_t)W), (int)(src % (uint32_t)W));
							}
						}
					}
					nseed = __shfl_sync(FULL_MASK, nseed, 0);
					__syncwarp();
					if (warp_relax(t, nseed, 0u) && lane == 0)
						s_guard = 1;
				}
				__syncthreads();

				// ---- blocked, non-source tiles read "unreachable" (they were never relaxed: paths.cpp:65-67) ----
				for (int i = threadIdx.x; i < T; i += blockDim.x)
				{
					const bool blocked = (gf[i] & F_BLOCKING) != 0 || (uint32_t)i == P.extra_blocked;
					if (!blocked)
						continue;
					for (int r = 0; r < lb; r++)
					{
						bool is_src = false;
						for (int s = 0; s < nsrc_layer; s++)
							is_src |= (P.nsrc > 1 ? s_wp[s] : s_wp[l0 + r]) == (uint32_t)i;
						if (!is_src)
							sdist[(size_t)r * T + i] = MTD_INF;
					}
				}
				__syncthreads();

				// ---- next-hop tiles + write-out of this batch (coalesced, streaming stores) ----
				for (int r = 0; r < lb; r++)
				{
					const uint32_t *dl = sdist + (size_t)r * T;
					uint32_t *od = gdist ? gdist + (size_t)(l0 + r) * T : nullptr;
					uint32_t *ot = gtiles ? gtiles + (size_t)(l0 + r) * T : nullptr;
					if (od)
						for (int i = threadIdx.x; i < T; i += blockDim.x)
							st_stream_u32(od + i, dl[i]);
					if (ot)
						for (int i = threadIdx.x; i < T; i += blockDim.x)
							st_stream_u32(ot + i, next_hop(dl, wr, wd, W, H, i % W, i / W));
				}
				__syncthreads();
			}

			if (s_guard && threadIdx.x == 0)
				atomicOr(P.status, ST_GUARD);
			if (!P.fullpath_len || P.nsrc > 1)
				continue;

			// ---- inter-waypoint table D[a][b] = dist_b[wp_a] ----
			for (int i = threadIdx.x; i < n * n; i += blockDim.x)
			{
				const int a = i / n, b = i % n;
				const uint32_t *dl = resident_all ? sdist + (size_t)b * T : gdist + (size_t)b * T;
				Dtab[a * FIND_MAXN + b] = dl[s_wp[a]];
			}
			__syncthreads();

			// ---- full-path walk: paths.cpp:87-108. One lane per spawner (a dependent pointer chase). ----
			if (warp < n && lane == 0)
			{
				const uint32_t *dbase = resident_all ? sdist : gdist;
				uint32_t *path = P.fullpath ? P.fullpath + ((size_t)mapi * n + warp) * P.cap : nullptr;
				uint32_t bits = 0, prev = MTD_INF, distance = 0, len = 0;
				uint32_t tile = s_wp[warp];
				while (tile != MTD_INF)
				{
					for (int k = 0; k < n; k++)
						if (s_wp[k] == tile)
							bits |= 1u << k;
					if (path && len < P.cap)
						path[len] = tile;
					len++;
					if (prev != MTD_INF)
					{
						// Grid::neighborDistance(prev, tile) for adjacent tiles = stored edge weight
						const uint32_t lo = prev < tile ? prev : tile, hi = prev < tile ? tile : prev;
						distance += MIN_EDGE + (hi - lo == 1u ? wr[lo] : wd[lo]);
					}
					prev = tile;
					const FindResult fr = find_best(n, bits, Dtab, [&](uint32_t i) { return dbase[(size_t)i * T + tile]; });
					tile = fr.layer == MTD_INF ? MTD_INF : next_hop(dbase + (size_t)fr.layer * T, wr, wd, W, H, (int)(tile % (uint32_t)W), (int)(tile / (uint32_t)W));
					if (len > (uint32_t)T * (uint32_t)n + 8u)
						break; // cannot happen on consistent layers; guards against a hang on corrupted input
				}
				if (path)
					for (uint32_t i = len; i < P.cap; i++)
						path[i] = MTD_INF;
				if (P.fullpath && len > P.cap)
					atomicOr(P.status, ST_PATH_TRUNC);
				P.fullpath_len[(size_t)mapi * n + warp] = len;
				P.full_distance[(size_t)mapi * n + warp] = distance;
				s_fd[warp] = distance;
			}
			__syncthreads();

			// ---- aggregates: paths.cpp:122-138 ----
			if (threadIdx.x == 0)
			{
				uint32_t sum = 0;
				for (int i = 0; i < n; i++)
					sum += s_fd[i];
				const float avg = __fdiv_rn(stor((int32_t)sum), (float)(unsigned long long)n);
				uint32_t best = 0;
				for (int i = 0; i < n; i++)
					if (s_fd[i] < s_fd[best])
						best = (uint32_t)i;
				uint32_t *ag = P.agg + (size_t)mapi * 3;
				ag[0] = (uint32_t)__float2ll_rz(avg);
				ag[1] = (uint32_t)__float2ll_rz(stor((int32_t)s_fd[best]));
				ag[2] = best;
			}
		}
	}
}
