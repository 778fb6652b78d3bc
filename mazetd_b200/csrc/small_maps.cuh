// small_maps.cuh -- the procedural grid generator for maps that fit one CTA's shared memory (the game's 101x101
// map and batches of them):
//
//   k_mapgen_grid      newProcedural + newGrid + spawn picking      (sources/map/procedural.cpp:13-70,
//                                                                    sources/map/grid.cpp:12-198,
//                                                                    sources/map/paths.cpp:173-196)
// One CTA owns one map at a time and loops over maps (persistent grid sized from the SM count). The fused
// Waypoints::update that follows it is in waypoints_small.cuh.
#pragma once

#include "common.cuh"
#include "noise.cuh"
#include "sssp_warp.cuh"
#include "waypoints_small.cuh"

namespace mtd
{
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
		const int16_t *elev_in; // non-null: caller's elevations [nmaps][T] replace the noise (newGrid stages after grid.cpp:159)
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

		const uint32_t magicW = W > 1 ? (uint32_t)((0x100000000ull + (uint32_t)W - 1u) / (uint32_t)W) : 0u; // y = umulhi(i, magicW) for i < 65536, W <= 256

		noise::fill_gradients(grad);
#ifdef MTD_TIMING
		long long phase_last__ = clock64();
#define MTD_GPHASE(k) do { if (blockIdx.x == 0 && threadIdx.x == 0) { const long long now__ = clock64(); g_phase[8 + k] += (unsigned long long)(now__ - phase_last__); phase_last__ = now__; } } while (0)
#else
#define MTD_GPHASE(k)
#endif

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
			MTD_GPHASE(0); // rng

			// ---- elevations: grid.cpp:153-159 ----
			const noise::ElevationParams lp = ep;
			if (P.elev_in)
			{
				const int16_t *src = P.elev_in + (size_t)mapi * T;
				for (int i = threadIdx.x; i < T; i += blockDim.x)
					elev[i] = src[i];
			}
			else
				for (int i = threadIdx.x; i < T; i += blockDim.x)
				{
					const int y = (int)div_w((uint32_t)i, magicW), x = i - y * W;
					float e = noise::elevation(grad, lp, (float)(x - P.ox), (float)(y - P.oy));
					e = e > -100.0f ? e : -100.0f;
					e = e < 100.0f ? e : 100.0f;
					elev[i] = rtos16(e);
				}
			__syncthreads();
			MTD_GPHASE(1); // noise

			// ---- biome flags, slope, radius: grid.cpp:160-183 (thresholds are rtos16(-7), (1), (6), (0.35)) ----
			for (int i = threadIdx.x; i < T; i += blockDim.x)
			{
				const int y = (int)div_w((uint32_t)i, magicW), x = i - y * W;
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
			MTD_GPHASE(2); // flags

			// ---- fillUnreachable: grid.cpp:29-78 as union-find (root = lowest tile index = discovery order) ----
			// (1) horizontal runs by warp ballot: every valid tile points at the start of its run inside its 32-tile chunk
			//     (or at the last tile of the previous chunk when the run continues there): chains are <= ceil(W/32)+1 deep
			//     and no atomics are needed; (2) one union per pair of vertically overlapping runs, not per tile.
			for (int base = 0; base < T; base += (int)blockDim.x)
			{
				const int i = base + (int)threadIdx.x;
				bool valid = false, linked = false;
				if (i < T)
				{
					valid = flags[i] != F_INVALID;
					const int y = (int)div_w((uint32_t)i, magicW), x = i - y * W;
					linked = valid && x > 0 && flags[i - 1] != F_INVALID;
				}
				const uint32_t lm = __ballot_sync(FULL_MASK, linked);
				if (valid)
				{
					const uint32_t lane = lane_id();
					const uint32_t stops = ~lm & (0xFFFFFFFFu >> (31u - lane)); // lanes <= mine that start a run
					parent[i] = stops ? (uint32_t)(i - (int)lane + (31 - __clz((int)stops))) : (uint32_t)(i - (int)lane - 1);
				}
			}
			__syncthreads();
			MTD_GPHASE(3); // runs
			for (int i = threadIdx.x; i < T - W; i += blockDim.x)
			{
				if (flags[i] == F_INVALID || flags[i + W] == F_INVALID)
					continue;
				const int y = (int)div_w((uint32_t)i, magicW), x = i - y * W;
				// the pair to the left already joined the same two runs
				if (x > 0 && flags[i - 1] != F_INVALID && flags[i + W - 1] != F_INVALID)
					continue;
				uf_union(parent, (uint32_t)i, (uint32_t)(i + W));
			}
			__syncthreads();
			MTD_GPHASE(4); // unions
			// flatten, then count members in the upper 16 bits of the root's word (T < 65536)
			for (int i = threadIdx.x; i < T; i += blockDim.x)
				if (flags[i] != F_INVALID)
				{
					const uint32_t r = uf_find(parent, (uint32_t)i);
					if (r != (uint32_t)i)
						parent[i] = r; // benign: only shortens chains, roots are untouched
				}
			__syncthreads();
			MTD_GPHASE(5); // flatten
			for (int base = 0; base < T; base += (int)blockDim.x)
			{
				const int i = base + (int)threadIdx.x;
				const bool valid = i < T && flags[i] != F_INVALID;
				const uint32_t r = valid ? (parent[i] & 0xFFFFu) : 0xFFFFFFFFu;
				// one atomic per distinct root per warp
				const uint32_t peers = __match_any_sync(FULL_MASK, r);
				if (valid && (uint32_t)(__ffs((int)peers) - 1) == lane_id())
					atomicAdd(&parent[r], (uint32_t)__popc(peers) << 16);
			}
			__syncthreads();
			MTD_GPHASE(6); // count
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
			MTD_GPHASE(7); // best
			const bool any_valid = best_key != 0ull;
			const uint32_t best_root = 0xFFFFFFFFu - (uint32_t)(best_key & 0xFFFFFFFFull);
			for (int i = threadIdx.x; i < T; i += blockDim.x)
				if (flags[i] == F_INVALID || !any_valid || (parent[i] & 0xFFFFu) != best_root)
					flags[i] = F_INVALID;
			__syncthreads();
			MTD_GPHASE(8); // apply

			// ---- findBorder + Mana: grid.cpp:80-97, 186-196 (border bits only land on invalid tiles, so in-place is safe) ----
			for (int i = threadIdx.x; i < T; i += blockDim.x)
			{
				const int y = (int)div_w((uint32_t)i, magicW), x = i - y * W;
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
			MTD_GPHASE(9); // border

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
				if (!any_valid && P.nwp > 0) // without spawn picking an empty map is a legitimate newGrid result
					atomicOr(P.status, ST_NO_VALID);
			}
			__syncthreads();
			MTD_GPHASE(10); // spawn picking

			// ---- write out (coalesced 4-byte words; T even-padded arrays are 4-byte aligned per map only if T is even,
			//      so use 2-byte stores: 40 KB per map, negligible next to the layers) ----
			int16_t *ge = P.elev + (size_t)mapi * T;
			uint16_t *gf = P.flags + (size_t)mapi * T;
			for (int i = threadIdx.x; i < T; i += blockDim.x)
			{
				ge[i] = elev[i];
				gf[i] = flags[i];
			}
			MTD_GPHASE(11); // write-out
		}
#ifdef MTD_TIMING
		if (blockIdx.x == 0 && threadIdx.x == 0)
		{
			printf("MAPGEN rng=%llu noise=%llu flags=%llu runs=%llu unions=%llu flatten=%llu count=%llu best=%llu apply=%llu border=%llu spawn=%llu writeout=%llu (cycles, CTA 0)\n", g_phase[8], g_phase[9], g_phase[10], g_phase[11], g_phase[12], g_phase[13], g_phase[14], g_phase[15], g_phase[16], g_phase[17], g_phase[18], g_phase[19]);
			for (int k = 8; k < 24; k++)
				g_phase[k] = 0;
		}
#endif
	}
}
