// monsters.cuh -- per-tick monster steering on SoA state (sources/game/monsters.cpp:55-99), the position lerp
// (sources/game/game.cpp:15-21), spawn state (sources/game/monsterSpawning.cpp:205-216), batched Waypoints::find
// (sources/map/paths.cpp:141-169) and the reachability scan of placingBuildingBlocksMonsters
// (sources/player/buildingsPlacement.cpp:30-41). One thread per monster / query; all gathers go to the resident
// distance layers in HBM/L2.
#pragma once

#include "common.cuh"
#include "sssp_warp.cuh"
#include "walk.cuh"

namespace mtd
{
	struct LayersView
	{
		int W, H, ox, oy, n;
		const int16_t *elev;
		const uint32_t *tiles; // [n][T]
		const uint32_t *dist;  // [n][T]
		const uint32_t *wp;    // [n]
		const uint32_t *Dtab;  // [8][8] inter-waypoint table D[a][b] = dist_b[wp_a]
	};

	struct MonstersView
	{
		int n;
		uint32_t *tile, *tileStart, *tileEnd, *timeStart, *timeEnd, *bits;
		const uint32_t *timeToArrive;
		const uint8_t *dotflags;
		const uint32_t *damage;
		uint8_t *alive;
		int *status; // ST_MONSTER_STUCK is raised here
	};

	__global__ void k_build_dtab(LayersView L, uint32_t *Dtab)
	{
		const int i = threadIdx.x;
		const size_t T = (size_t)L.W * L.H;
		if (i < L.n * L.n)
		{
			const int a = i / L.n, b = i % L.n;
			Dtab[a * FIND_MAXN + b] = L.dist[(size_t)b * T + L.wp[a]];
		}
	}

	__device__ __forceinline__ void device_find(const LayersView &L, uint32_t pos, uint32_t bits, uint32_t &tile, uint32_t &distance)
	{
		const size_t T = (size_t)L.W * L.H;
		const FindResult fr = find_best(L.n, bits, L.Dtab, [&](uint32_t i) { return L.dist[(size_t)i * T + pos]; });
		distance = fr.distance;
		tile = fr.layer == MTD_INF ? MTD_INF : L.tiles[(size_t)fr.layer * T + pos];
	}

	__global__ void k_find_batch(LayersView L, const uint32_t *pos, const uint32_t *bits, int nq, uint32_t *tile_out, uint32_t *dist_out)
	{
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= nq)
			return;
		uint32_t t, d;
		device_find(L, pos[i], bits[i], t, d);
		tile_out[i] = t;
		dist_out[i] = d;
	}

	// monsters.cpp:60-98, statement by statement
	__global__ void k_monsters_tick(LayersView L, MonstersView M, uint32_t gameTime, unsigned long long *damage_sum, uint32_t *alive_count)
	{
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		bool alive = false;
		if (i < M.n)
		{
			alive = M.alive[i] != 0;
			uint32_t timeEnd = M.timeEnd[i];
			if (alive && gameTime >= timeEnd) // :64
			{
				const uint32_t tile = M.tile[i];
				uint32_t bits = M.bits[i];
				if (tile >= (uint32_t)(L.W * L.H))
				{
					// not a tile of this grid (the reference would read out of bounds): leave the monster alone and report
					atomicOr(M.status, 32); // ST_MONSTER_STUCK
					goto count;
				}
				for (int k = 0; k < L.n; k++) // :67-69
					if (L.wp[k] == tile)
						bits |= 1u << k;
				M.bits[i] = bits;
				if (__popc(bits) == L.n) // :70-77
				{
					if (M.damage)
						atomicAdd(damage_sum, (unsigned long long)M.damage[i]);
					M.alive[i] = 0;
					alive = false;
				}
				else
				{
					uint32_t go_tile, go_dist;
					device_find(L, tile, bits, go_tile, go_dist); // :79
					if (go_tile >= (uint32_t)(L.W * L.H) || go_dist == 0u)
					{
						// CAGE_ASSERT(go.tile != m) / a division by zero in the reference (:80, :88): reachable through the
						// ABI (arbitrary uploaded tiles, a tile blocked by mtd_grid_set_flags): state untouched, reported
						atomicOr(M.status, 32); // ST_MONSTER_STUCK
						goto count;
					}
					M.tileStart[i] = tile;                          // :82-84
					M.tileEnd[i] = go_tile;
					M.tile[i] = go_tile;
					M.timeStart[i] = gameTime;
					const uint32_t tta = M.timeToArrive[i];
					if (tta > gameTime) // :85-89 (uint64 arithmetic through distNext)
					{
						const unsigned long long distNext = edge_weight(L.elev[tile], L.elev[go_tile]);
						timeEnd = (uint32_t)((unsigned long long)gameTime + (unsigned long long)(tta - gameTime) * distNext / (unsigned long long)go_dist);
					}
					timeEnd = max(timeEnd, gameTime + 1u); // :90
					const uint8_t df = M.dotflags ? M.dotflags[i] : (uint8_t)0;
					const bool fire = df & 1, water = df & 2;
					if (water && !fire) // :92-93
						timeEnd = gameTime + (timeEnd - gameTime) * 2u;
					if (fire && !water) // :94-95
						timeEnd = gameTime + (timeEnd - gameTime) / 2u;
					timeEnd = max(timeEnd, gameTime + 1u); // :96
					M.timeEnd[i] = timeEnd;
				}
			}
		}
	count:
		const uint32_t bal = __ballot_sync(FULL_MASK, alive);
		if (lane_id() == 0 && bal)
			atomicAdd(alive_count, (uint32_t)__popc(bal));
	}

	// ---- "damage monsters" (sources/game/monsters.cpp:101-183), integer state only -------------------------------
	// One thread per monster over SoA: dot arrays are [5][n] in DamageTypeEnum order (game.h:18-27). The RNG-driven visual
	// effects (:140-149) and the ghost entity (:176) are not state and stay on the host. Also refreshes the fire / water bits
	// the steering tick reads (MonsterComponent::affected, monsters.cpp:13-16, 92-95).
	struct DotsView
	{
		int n;
		uint32_t *duration, *damage; // [5][n]
		const uint8_t *resist;       // [n] DamageTypeFlags bits 0..4, or null
		long long *life;             // [n] sint64 (game.h:162)
		const uint32_t *money;       // [n] or null
		uint8_t *alive;              // [n]
		uint8_t *dotflags;           // [n] bit0 = affected(Fire), bit1 = affected(Water)
	};

	__global__ void k_monsters_dots(DotsView D, unsigned long long *money_sum, uint32_t *killed)
	{
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		const bool live = i < D.n && D.alive[i];
		unsigned long long my_money = 0;
		bool died = false;
		if (live)
		{
			const size_t n = (size_t)D.n;
			uint32_t du[5], da[5];
#pragma unroll
			for (int t = 0; t < 5; t++)
			{
				du[t] = D.duration[t * n + i];
				da[t] = D.damage[t * n + i];
			}
			// dotsEliminateOposing<Fire, Water>, <Poison, Magic> (monsters.cpp:101-114)
			if (du[1] > 0 && du[2] > 0)
			{
				da[1] = da[2] = 0;
				du[1] = du[2] = 30;
			}
			if (du[3] > 0 && du[4] > 0)
			{
				da[3] = da[4] = 0;
				du[3] = du[4] = 30;
			}
			const uint32_t resist = D.resist ? D.resist[i] : 0u;
			long long life = D.life[i];
#pragma unroll
			for (int t = 0; t < 5; t++) // applyDot<Physical, Fire, Water, Poison, Magic> in this order (monsters.cpp:166-170)
			{
				const bool super = da[t == 0 ? 3 : 4] > 0; // read NOW: an earlier applyDot may have drained it
				uint32_t d = da[t];
				if (du[t] > 1)
					d /= du[t];
				if (du[t] > 0)
					du[t]--;
				if (d != 0)
				{
					da[t] -= d;
					if (super)
						d *= 2u;
					if (resist & (1u << t))
						d /= 4u;
					life -= (long long)d;
				}
			}
#pragma unroll
			for (int t = 0; t < 5; t++)
			{
				D.duration[t * n + i] = du[t];
				D.damage[t * n + i] = da[t];
			}
			D.life[i] = life;
			D.dotflags[i] = (uint8_t)((da[1] > 0 ? 1u : 0u) | (da[2] > 0 ? 2u : 0u));
			if (life <= 0) // monsters.cpp:172-178
			{
				died = true;
				my_money = D.money ? D.money[i] : 0u;
				D.alive[i] = 0;
			}
		}
		// one atomic per warp
		const uint32_t dead = __ballot_sync(FULL_MASK, died);
		if (dead)
		{
			for (int o = 16; o > 0; o >>= 1)
				my_money += __shfl_down_sync(FULL_MASK, my_money, o);
			if ((threadIdx.x & 31) == 0)
			{
				atomicAdd(money_sum, my_money);
				atomicAdd(killed, (uint32_t)__popc(dead));
			}
		}
	}

	__global__ void k_dots_flags(DotsView D)
	{
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i < D.n)
			D.dotflags[i] = (uint8_t)((D.damage[(size_t)1 * D.n + i] > 0 ? 1u : 0u) | (D.damage[(size_t)2 * D.n + i] > 0 ? 2u : 0u));
	}

	// game.cpp:15-21: saturate(Real(gameTime - (sint64)timeStart) / Real(timeEnd - (sint64)timeStart)), interpolate
	__global__ void k_monsters_positions(LayersView L, MonstersView M, uint32_t gameTime, float *out3n)
	{
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= M.n)
			return;
		const uint32_t ts = M.tileStart[i], te = M.tileEnd[i];
		const float num = __ll2float_rn((long long)gameTime - (long long)M.timeStart[i]);
		const float den = __ll2float_rn((long long)M.timeEnd[i] - (long long)M.timeStart[i]);
		float fac = __fdiv_rn(num, den);
		fac = fac > 0.0f ? fac : 0.0f; // max(v, 0) = v > 0 ? v : 0 ; NaN -> 0 (cage-core/math.h:1021-1032)
		fac = fac < 1.0f ? fac : 1.0f;
		const float ax = (float)((int)(ts % (uint32_t)L.W) - L.ox), az = (float)((int)(ts / (uint32_t)L.W) - L.oy);
		const float bx = (float)((int)(te % (uint32_t)L.W) - L.ox), bz = (float)((int)(te / (uint32_t)L.W) - L.oy);
		const float ay = stor(L.elev[ts]), by = stor(L.elev[te]);
		out3n[(size_t)i * 3 + 0] = __fadd_rn(ax, __fmul_rn(__fsub_rn(bx, ax), fac));
		out3n[(size_t)i * 3 + 1] = __fadd_rn(ay, __fmul_rn(__fsub_rn(by, ay), fac));
		out3n[(size_t)i * 3 + 2] = __fadd_rn(az, __fmul_rn(__fsub_rn(bz, az), fac));
	}

	// monsterSpawning.cpp:205-216
	__global__ void k_monster_spawn(LayersView L, uint32_t spawner, uint32_t gameTime, float speed, uint32_t *out3)
	{
		const uint32_t position = L.wp[spawner];
		const uint32_t bits = 1u << spawner;
		uint32_t t, d;
		device_find(L, position, bits, t, d);
		out3[0] = position;
		out3[1] = bits;
		out3[2] = gameTime + (uint32_t)__float2ll_rz(__fdiv_rn(__fmul_rn(30.0f, stor((int32_t)d)), speed));
	}

	// Full-path walks + aggregates of Waypoints::update (sources/map/paths.cpp:87-139) on layers resident in global
	// memory (grids too large for the one-CTA kernel). One warp, lane s walks spawner s (walk.cuh). Launch: <<<1, 32>>>.
	__global__ void k_walk_global(LayersView L, uint32_t *fullpath, uint32_t cap, uint32_t *fullpath_len, uint32_t *full_distance, uint32_t *agg, int *status)
	{
		__shared__ uint32_t s_fd[FIND_MAXN], s_wp[FIND_MAXN], s_D[FIND_MAXN * FIND_MAXN];
		const uint32_t lane = lane_id();
		const int n = L.n;
		const size_t T = (size_t)L.W * L.H;
		GlobalLayersAcc acc;
		acc.gdist = L.dist;
		acc.gtiles = L.tiles;
		acc.gelev = L.elev;
		acc.T = T;
		if ((int)lane < n)
			s_wp[lane] = L.wp[lane];
		for (int i = (int)lane; i < FIND_MAXN * FIND_MAXN; i += 32)
			s_D[i] = L.Dtab[i];
		__syncwarp();
		const bool table_fast = walk_table_fast(n, s_D);
		WalkResult r;
		r.len = 0;
		r.distance = 0;
		if ((int)lane < n)
		{
			uint32_t *path = fullpath ? fullpath + (size_t)lane * cap : nullptr;
			const size_t maxlen = T * (size_t)n + 8u;
			r = walk_path(n, s_wp, s_D, table_fast, (int)lane, acc, path, cap, maxlen > 0xFFFFFFF0ull ? 0xFFFFFFF0u : (uint32_t)maxlen);
			if (fullpath && r.len > cap)
				atomicOr(status, 4); // ST_PATH_TRUNC
			fullpath_len[lane] = r.len;
			full_distance[lane] = r.distance;
			s_fd[lane] = r.distance;
		}
		__syncwarp();
		if (fullpath)
			for (int s = 0; s < n; s++)
				for (uint32_t i = __shfl_sync(FULL_MASK, r.len, s) + lane; i < cap; i += 32u)
					fullpath[(size_t)s * cap + i] = MTD_INF;
		if (lane == 0)
			walk_aggregates(n, s_fd, agg);
	}

	// buildingsPlacement.cpp:30-41 on a trial distance layer
	__global__ void k_trial_check(const uint32_t *trial_dist, const uint32_t *wp, int nwp, const uint32_t *monster_tile, const uint8_t *monster_alive, int nmonsters, int *blocked_out)
	{
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		bool bad = false;
		if (i < nwp)
			bad = trial_dist[wp[i]] == MTD_INF;
		if (i < nmonsters && (!monster_alive || monster_alive[i]))
			bad |= trial_dist[monster_tile[i]] == MTD_INF;
		if (__any_sync(FULL_MASK, bad) && lane_id() == 0)
			*blocked_out = 1;
	}

	// Grid::neighborDistance for ANY pair of tiles (sources/map/grid.cpp:141-145): the general form with
	// d = center(a) - center(b) and the dot-product sum order ((0 + d0*d0) + d1*d1) + d2*d2.
	__global__ void k_neighbor_distances(int W, int ox, int oy, const int16_t *elev, const uint32_t *a, const uint32_t *b, int n, uint32_t *out)
	{
		const int i = blockIdx.x * blockDim.x + threadIdx.x;
		if (i >= n)
			return;
		const uint32_t ta = a[i], tb = b[i];
		const float d0 = __fsub_rn((float)((int)(ta % (uint32_t)W) - ox), (float)((int)(tb % (uint32_t)W) - ox));
		const float d1 = __fsub_rn(stor(elev[ta]), stor(elev[tb]));
		const float d2 = __fsub_rn((float)((int)(ta / (uint32_t)W) - oy), (float)((int)(tb / (uint32_t)W) - oy));
		float sum = __fmul_rn(d0, d0);
		sum = __fadd_rn(sum, __fmul_rn(d1, d1));
		sum = __fadd_rn(sum, __fmul_rn(d2, d2));
		out[i] = (uint32_t)rtos32(__fsqrt_rn(sum));
	}

	__global__ void k_unpack_weights(int W, int H, const int16_t *elev, uint32_t *right, uint32_t *down)
	{
		const size_t total = (size_t)W * H;
		for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
		{
			const int x = (int)(i % (size_t)W), y = (int)(i / (size_t)W);
			right[i] = x < W - 1 ? edge_weight(elev[i], elev[i + 1]) : 0u;
			down[i] = y < H - 1 ? edge_weight(elev[i], elev[i + W]) : 0u;
		}
	}
}
