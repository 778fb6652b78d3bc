// towers.cuh -- the two steps that follow steering in every game tick:
//   * "spatial rebuild": the monster index of sources/game/spatial.cpp:24-34 (cage's BVH over MovementComponent::position()
//     of every monster, y flattened) as a UNIFORM GRID: one bin per map tile, counting sort of the resident monster SoA
//     (k_spatial_bin -> k_spatial_scan -> k_spatial_scatter). Positions are lerps between tile centres, so a monster always
//     lies in the unit square of a tile; a row of bins is one contiguous run of the sorted arrays.
//   * "attacks": sources/game/attacks.cpp:132-160 for every tower, one warp per tower: firing delay, applyMods (:40-71),
//     sphere query around the tower + class filter (:73-92), mana (:95-105), target picker (:171-216), optional splash
//     query around the target (:94-98), damage-over-time added to the hit monsters (:120-130, atomics: sums commute).
// What is restated is the SET a query returns (sqrt(dx*dx + dz*dz) <= radius, geometry.cpp:82-86, 471-474); the order in
// which cage's BVH enumerates it is not observable, so ties of a picker's value go to the lowest monster index (the oracle
// does the same). TargetingEnum::Random depends on that order and on the global RNG: unpinned, see mtd_towers_upload.
#pragma once

#include "common.cuh"
#include "monsters.cuh"

namespace mtd
{
	struct TowersView
	{
		int n;
		const uint32_t *tile, *damage, *overTime, *firingPeriod, *manaCost;
		const float *firingRange, *splashRadius;
		const uint8_t *invalidClasses, *element, *bonus, *targeting;
		uint32_t *firingDelay, *mana;
		const unsigned long long *random_words; // may be null
		int32_t *target;                        // picked monster or -1
		uint32_t *hits;                         // monsters that received the dot
	};

	struct SpatialView
	{
		int W, H, ox, oy;
		int n;                 // monsters
		float *pos;            // [n][3] MovementComponent::position() at this tick
		uint32_t *bin_count;   // [W*H + 1]
		uint32_t *bin_start;   // [W*H + 1] exclusive scan of bin_count
		uint32_t *bin_of;      // [n] bin of monster i (MTD_INF: dead)
		uint32_t *rank;        // [n] arrival order inside the bin
		uint32_t *sorted_idx;  // [n] monster indices, bins in row-major order
		float2 *sorted_xz;     // [n] their flattened positions
	};

	__device__ __forceinline__ void monster_position(const LayersView &L, const MonstersView &M, int i, uint32_t gameTime, float &x, float &y, float &z)
	{
		// game.cpp:15-21, same arithmetic as k_monsters_positions
		const uint32_t ts = M.tileStart[i], te = M.tileEnd[i];
		const float num = __ll2float_rn((long long)gameTime - (long long)M.timeStart[i]);
		const float den = __ll2float_rn((long long)M.timeEnd[i] - (long long)M.timeStart[i]);
		float fac = __fdiv_rn(num, den);
		fac = fac > 0.0f ? fac : 0.0f;
		fac = fac < 1.0f ? fac : 1.0f;
		const float ax = (float)((int)(ts % (uint32_t)L.W) - L.ox), az = (float)((int)(ts / (uint32_t)L.W) - L.oy);
		const float bx = (float)((int)(te % (uint32_t)L.W) - L.ox), bz = (float)((int)(te / (uint32_t)L.W) - L.oy);
		const float ay = stor(L.elev[ts]), by = stor(L.elev[te]);
		x = __fadd_rn(ax, __fmul_rn(__fsub_rn(bx, ax), fac));
		y = __fadd_rn(ay, __fmul_rn(__fsub_rn(by, ay), fac));
		z = __fadd_rn(az, __fmul_rn(__fsub_rn(bz, az), fac));
	}

	__device__ __forceinline__ int bin_coord(float v, int off, int n)
	{
		float t = v + (float)off + 0.5f;
		t = fminf(fmaxf(t, -1.0f), (float)n); // also +-inf radii
		const int b = (int)floorf(t);
		return b < 0 ? 0 : (b >= n ? n - 1 : b);
	}

	// spatial.cpp:30: position of every live monster, its bin, its arrival order in the bin
	__global__ void k_spatial_bin(LayersView L, MonstersView M, SpatialView S, uint32_t gameTime)
	{
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= M.n)
			return;
		float x, y, z;
		monster_position(L, M, i, gameTime, x, y, z);
		S.pos[(size_t)i * 3 + 0] = x;
		S.pos[(size_t)i * 3 + 1] = y;
		S.pos[(size_t)i * 3 + 2] = z;
		if (!M.alive[i])
		{
			S.bin_of[i] = MTD_INF;
			return;
		}
		const uint32_t bin = (uint32_t)(bin_coord(z, S.oy, S.H) * S.W + bin_coord(x, S.ox, S.W));
		S.bin_of[i] = bin;
		S.rank[i] = atomicAdd(S.bin_count + bin, 1u);
	}

	// exclusive scan of the bin counts by one CTA (the game map has 10 201 bins)
	__global__ void k_spatial_scan(SpatialView S)
	{
		__shared__ uint32_t warp_sums[32];
		__shared__ uint32_t carry;
		const int nb = S.W * S.H + 1;
		if (threadIdx.x == 0)
			carry = 0;
		__syncthreads();
		for (int base = 0; base < nb; base += (int)blockDim.x)
		{
			const int i = base + (int)threadIdx.x;
			const uint32_t v = i < nb - 1 ? S.bin_count[i] : 0u;
			uint32_t inc = v;
			for (int o = 1; o < 32; o <<= 1)
			{
				const uint32_t t = __shfl_up_sync(FULL_MASK, inc, o);
				if ((int)(threadIdx.x & 31) >= o)
					inc += t;
			}
			if ((threadIdx.x & 31) == 31)
				warp_sums[threadIdx.x >> 5] = inc;
			__syncthreads();
			if (threadIdx.x < 32)
			{
				uint32_t w = threadIdx.x < (blockDim.x >> 5) ? warp_sums[threadIdx.x] : 0u;
				for (int o = 1; o < 32; o <<= 1)
				{
					const uint32_t t = __shfl_up_sync(FULL_MASK, w, o);
					if ((int)threadIdx.x >= o)
						w += t;
				}
				warp_sums[threadIdx.x] = w;
			}
			__syncthreads();
			const uint32_t before = carry + ((threadIdx.x >> 5) ? warp_sums[(threadIdx.x >> 5) - 1] : 0u) + inc - v;
			if (i < nb)
				S.bin_start[i] = before;
			__syncthreads();
			if (threadIdx.x == blockDim.x - 1)
				carry = before + v;
			__syncthreads();
		}
	}

	__global__ void k_spatial_scatter(SpatialView S)
	{
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= S.n)
			return;
		const uint32_t bin = S.bin_of[i];
		if (bin == MTD_INF)
			return;
		const uint32_t j = S.bin_start[bin] + S.rank[i];
		S.sorted_idx[j] = (uint32_t)i;
		S.sorted_xz[j] = make_float2(S.pos[(size_t)i * 3 + 0], S.pos[(size_t)i * 3 + 2]);
	}

	// distanceSquared on the y-flattened positions: ((0 + dx*dx) + 0) + dz*dz (cage-core/math.h:919-938)
	__device__ __forceinline__ float flat_dist2(float ax, float az, float bx, float bz)
	{
		const float dx = __fsub_rn(ax, bx), dz = __fsub_rn(az, bz);
		return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dz, dz));
	}

	// Runs `f(j)` for every sorted entry j whose bin can hold a point within `r` of (cx, cz); lanes stride over the runs.
	template<class F>
	__device__ __forceinline__ void for_candidates(const SpatialView &S, float cx, float cz, float r, uint32_t lane, F f)
	{
		const int bx0 = bin_coord(cx - r, S.ox, S.W), bx1 = bin_coord(cx + r, S.ox, S.W);
		const int by0 = bin_coord(cz - r, S.oy, S.H), by1 = bin_coord(cz + r, S.oy, S.H);
		for (int by = by0; by <= by1; by++)
		{
			const uint32_t a = S.bin_start[by * S.W + bx0], b = S.bin_start[by * S.W + bx1 + 1];
			for (uint32_t j = a + lane; j < b; j += 32u)
				f(j);
		}
	}

	struct TargetStats
	{
		const uint8_t *monsterClass;   // [n] MonsterClassFlags (null: none)
		const uint32_t *timeToArrive;  // [n]
		const long long *life;         // [n]
		uint32_t *dot_duration, *dot_damage; // [5][n]
	};

	// (value, index) lexicographic minimum over the warp
	__device__ __forceinline__ void warp_best(float &v, uint32_t &i)
	{
		for (int o = 16; o > 0; o >>= 1)
		{
			const float ov = __shfl_xor_sync(FULL_MASK, v, o);
			const uint32_t oi = __shfl_xor_sync(FULL_MASK, i, o);
			if (oi != MTD_INF && (i == MTD_INF || ov < v || (ov == v && oi < i)))
			{
				v = ov;
				i = oi;
			}
		}
	}

	// attacks.cpp:132-160, one warp per tower
	__global__ void k_towers_attack(TowersView Tw, SpatialView S, TargetStats G, int Wgrid, int ox, int oy)
	{
		const int t = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
		const uint32_t lane = lane_id();
		if (t >= Tw.n)
			return;
		if (lane == 0)
		{
			Tw.target[t] = -1;
			Tw.hits[t] = 0u;
		}
		const uint32_t delay = Tw.firingDelay[t];
		if (delay > 0u) // :136-140
		{
			if (lane == 0)
				Tw.firingDelay[t] = delay - 1u;
			return;
		}
		// applyMods, :40-71
		uint32_t damage = Tw.damage[t], overTime = Tw.overTime[t], firingPeriod = Tw.firingPeriod[t], manaCost = Tw.manaCost[t];
		float firingRange = Tw.firingRange[t], splashRadius = Tw.splashRadius[t];
		const int element = Tw.element[t];
		if (element != 0)
		{
			overTime = (uint32_t)__float2int_rz(__fmul_rn(__fsqrt_rn((float)damage), 20.0f)); // numeric_cast<uint32>(sqrt(damage) * 20)
			damage *= 3u;
		}
		switch (Tw.bonus[t])
		{
		case 1: damage *= 2u; manaCost *= 2u; break;
		case 2: firingPeriod /= 2u; break;
		case 3: firingRange = __fadd_rn(firingRange, 4.0f); break;
		case 4: splashRadius = __fadd_rn(splashRadius, 2.0f); manaCost *= 3u; break;
		case 5: overTime /= 5u; break;
		case 6: manaCost /= 3u; break;
		default: break;
		}
		const uint32_t tile = Tw.tile[t];
		const float mx = (float)((int)(tile % (uint32_t)Wgrid) - ox), mz = (float)((int)(tile / (uint32_t)Wgrid) - oy);
		const uint8_t invalid = Tw.invalidClasses[t];
		const int targeting = Tw.targeting[t];
		const size_t n = (size_t)S.n;

		// findMonstersForTower + pickTargetMonster (ties: lowest monster index)
		float bv = 0.0f;
		uint32_t bi = MTD_INF, count = 0;
		for_candidates(S, mx, mz, firingRange, lane, [&](uint32_t j)
					   {
						   const float2 p = S.sorted_xz[j];
						   const float d2 = flat_dist2(p.x, p.y, mx, mz);
						   if (!(__fsqrt_rn(d2) <= firingRange))
							   return;
						   const uint32_t i = S.sorted_idx[j];
						   if (G.monsterClass && (G.monsterClass[i] & invalid))
							   return;
						   count++;
						   float v;
						   switch (targeting)
						   {
						   case 1: v = __uint2float_rn(G.timeToArrive[i]); break;
						   case 2: v = __int2float_rn(-(int)G.timeToArrive[i]); break;
						   case 3: v = d2; break;
						   case 4: v = -d2; break;
						   case 5: v = __ll2float_rn(-G.life[i]); break;
						   case 6: v = __ll2float_rn(G.life[i]); break;
						   default: v = 0.0f; break;
						   }
						   if (bi == MTD_INF || v < bv || (v == bv && i < bi))
						   {
							   bv = v;
							   bi = i;
						   }
					   });
		count = __reduce_add_sync(FULL_MASK, count);
		if (count == 0u)
			return; // :152-153
		if (manaCost > 0u) // consumeMana, :95-105
		{
			const uint32_t mana = Tw.mana[t];
			if (mana < manaCost)
				return;
			__syncwarp();
			if (lane == 0)
				Tw.mana[t] = mana - manaCost;
		}
		warp_best(bv, bi);
		if (targeting == 0)
		{
			// Random (unpinned): candidate number random_words[t] % count in ascending monster index, found by bisection
			const unsigned long long want = (Tw.random_words ? Tw.random_words[t] : 0ull) % (unsigned long long)count;
			uint32_t lo = 0, hi = (uint32_t)S.n - 1u;
			while (lo < hi)
			{
				const uint32_t mid = lo + (hi - lo) / 2u;
				uint32_t c = 0;
				for_candidates(S, mx, mz, firingRange, lane, [&](uint32_t j)
							   {
								   const float2 p = S.sorted_xz[j];
								   if (!(__fsqrt_rn(flat_dist2(p.x, p.y, mx, mz)) <= firingRange))
									   return;
								   const uint32_t i = S.sorted_idx[j];
								   if (G.monsterClass && (G.monsterClass[i] & invalid))
									   return;
								   c += i <= mid ? 1u : 0u;
							   });
				c = __reduce_add_sync(FULL_MASK, c);
				if ((unsigned long long)c > want)
					hi = mid;
				else
					lo = mid + 1u;
			}
			bi = lo;
		}
		// damageMonsters on the target, or on the splash set around it (:157-159, :120-130)
		uint32_t hits = 0;
		if (splashRadius > 0.0f)
		{
			const float tx = S.pos[(size_t)bi * 3 + 0], tz = S.pos[(size_t)bi * 3 + 2];
			for_candidates(S, tx, tz, splashRadius, lane, [&](uint32_t j)
						   {
							   const float2 p = S.sorted_xz[j];
							   if (!(__fsqrt_rn(flat_dist2(p.x, p.y, tx, tz)) <= splashRadius))
								   return;
							   const uint32_t i = S.sorted_idx[j];
							   if (G.monsterClass && (G.monsterClass[i] & invalid))
								   return;
							   atomicAdd(G.dot_damage + (size_t)element * n + i, damage);
							   atomicAdd(G.dot_duration + (size_t)element * n + i, overTime);
							   hits++;
						   });
			hits = __reduce_add_sync(FULL_MASK, hits);
		}
		else
		{
			if (lane == 0)
			{
				atomicAdd(G.dot_damage + (size_t)element * n + bi, damage);
				atomicAdd(G.dot_duration + (size_t)element * n + bi, overTime);
			}
			hits = 1u;
		}
		if (lane == 0)
		{
			Tw.target[t] = (int32_t)bi;
			Tw.hits[t] = hits;
			Tw.firingDelay[t] = firingPeriod; // += on a delay of 0 (:161)
		}
	}
}
