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
#include "mapgen_steps.cuh"
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

	// (512, 2): two CTAs per SM hide each other's latency in the label-propagation passes; 64 registers is what that allows
	__global__ void __launch_bounds__(512, 2) k_mapgen_grid(const MapgenGridParams P)
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
		MapgenShape S;
		S.W = W;
		S.H = H;
		S.ox = P.ox;
		S.oy = P.oy;
		S.radius = P.radius;
		S.magicW = div_w_magic(W);
		const int tid = (int)threadIdx.x, nt = (int)blockDim.x;
		auto sync = [] { __syncthreads(); };

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
			if (tid == 0)
			{
				Rng g;
				noise::ElevationParams e;
				mapgen_draws(P.s1, P.s2, P.first + mapi, e, g);
				ep = e;
				rng = g;
			}
			__syncthreads();
			MTD_GPHASE(0); // rng

			// ---- elevations: grid.cpp:153-159 (or the caller's, newGrid after grid.cpp:159) ----
			if (P.elev_in)
			{
				const int16_t *src = P.elev_in + (size_t)mapi * T;
				for (int i = tid; i < T; i += nt)
					elev[i] = src[i];
			}
			else
			{
				const noise::ElevationParams lp = ep;
				mapgen_elevations(grad, lp, S, elev, tid, nt);
			}
			__syncthreads();
			MTD_GPHASE(1); // noise

			const bool any_valid = mapgen_flags(S, elev, flags, parent, &best_key, tid, nt, sync);
			MTD_GPHASE(2); // flags, components, border

			if (tid == 0)
				mapgen_pick_spawns(rng, T, flags, P.nwp, any_valid, P.wp_tiles + (size_t)mapi * P.nwp, P.status);
			__syncthreads();
			MTD_GPHASE(3); // spawn picking

			// ---- write out (2-byte stores: a map's arrays are 4-byte aligned only if T is even; 40 KB per map, negligible
			//      next to the layers) ----
			int16_t *ge = P.elev + (size_t)mapi * T;
			uint16_t *gf = P.flags + (size_t)mapi * T;
			for (int i = tid; i < T; i += nt)
			{
				ge[i] = elev[i];
				gf[i] = flags[i];
			}
			MTD_GPHASE(4); // write-out
		}
#ifdef MTD_TIMING
		if (blockIdx.x == 0 && threadIdx.x == 0)
		{
			printf("MAPGEN rng=%llu noise=%llu flags+components+border=%llu spawn=%llu writeout=%llu (cycles, CTA 0)\n", g_phase[8], g_phase[9], g_phase[10], g_phase[11], g_phase[12]);
			for (int k = 8; k < 24; k++)
				g_phase[k] = 0;
		}
#endif
	}
}
