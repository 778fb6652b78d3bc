// mapgen_steps.cuh -- the stages of the procedural grid generator as device functions of a group of threads that
// owns one map in shared memory (sources/map/procedural.cpp:13-70, sources/map/grid.cpp:12-198,
// sources/map/paths.cpp:173-196). Two kernels are built from them: k_mapgen_grid (small_maps.cuh: maps only) and the
// fused mode of k_waypoints_small (waypoints_small.cuh: idle warps compute the next map's elevations while the current
// map's layers are relaxed, the rest of the generator runs in the shared memory the layers take afterwards).
//
// `tid` / `nt`: index of the calling thread inside the group and the group's size (a multiple of 32, whole warps);
// `sync()`: barrier of exactly that group.
#pragma once

#include "common.cuh"
#include "noise.cuh"

namespace mtd
{
	// i / W for i < 65536, W <= 256 with magicW = ceil(2^32 / W); W == 1 has no 32-bit magic and passes 0
	__device__ __forceinline__ uint32_t div_w(uint32_t i, uint32_t magicW)
	{
		return magicW ? __umulhi(i, magicW) : i;
	}

	__host__ __device__ inline uint32_t div_w_magic(int W)
	{
		return W > 1 ? (uint32_t)((0x100000000ull + (uint32_t)W - 1u) / (uint32_t)W) : 0u;
	}

	struct MapgenShape
	{
		int W, H, ox, oy, radius;
		uint32_t magicW;
	};

	// RNG draw order = member initialisation order of ProceduralImpl (procedural.cpp:13,24,33,43,52,61). One thread.
	// `rest` continues the stream for the spawn picking.
	__device__ __forceinline__ void mapgen_draws(uint64_t s1, uint64_t s2, int index, noise::ElevationParams &ep, Rng &rest)
	{
		Rng g;
		g.s0 = s1 + (uint64_t)index;
		g.s1 = s2;
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
		rest = g;
	}

	// elevations: grid.cpp:153-159. `out` may be shared or global memory.
	__device__ __forceinline__ void mapgen_elevations(const float2 *grad, const noise::ElevationParams &lp, const MapgenShape &S, int16_t *out, int tid, int nt)
	{
		const int T = S.W * S.H;
		for (int i = tid; i < T; i += nt)
		{
			const int y = (int)div_w((uint32_t)i, S.magicW), x = i - y * S.W;
			float e = noise::elevation(grad, lp, (float)(x - S.ox), (float)(y - S.oy));
			e = e > -100.0f ? e : -100.0f;
			e = e < 100.0f ? e : 100.0f;
			out[i] = rtos16(e);
		}
	}

	__device__ __forceinline__ uint32_t uf_find(uint32_t *parent, uint32_t i)
	{
		// parents only ever point at smaller indices, so a plain chase terminates
		uint32_t p = parent[i];
		while (p != i)
		{
			i = p;
			p = parent[i];
		}
		return i;
	}

	// find with path halving: every second node of the chain is re-pointed at its grandparent. The plain stores race with
	// the atomicMin of uf_union benignly: they only touch nodes that already have a parent (never a root, and a node never
	// becomes a root again), and they store an ancestor, so every parent stays below its node and inside its component; a
	// union that loses its own link to such a store has read old != b and goes on to merge that parent's root itself.
	__device__ __forceinline__ uint32_t uf_find_halve(uint32_t *parent, uint32_t i)
	{
		uint32_t p = parent[i];
		while (p != i)
		{
			const uint32_t g = parent[p];
			if (g == p)
				return p;
			parent[i] = g;
			i = g;
			p = parent[g];
		}
		return i;
	}

	__device__ __forceinline__ void uf_union(uint32_t *parent, uint32_t a, uint32_t b)
	{
		// lock-free union by minimum index (root = smallest tile index of the component)
		for (;;)
		{
			a = uf_find_halve(parent, a);
			b = uf_find_halve(parent, b);
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

	// biome flags, slope, radius (grid.cpp:160-183), fillUnreachable (grid.cpp:29-78), findBorder + Mana (grid.cpp:80-97,
	// 186-196). elev: [T] in, flags: [T] out, parent: [T] scratch, all in shared memory; *best_key: shared scratch word.
	// Returns whether the map has any valid tile. Ends with a sync().
	template<class Sync>
	__device__ __forceinline__ bool mapgen_flags(const MapgenShape &S, const int16_t *elev, uint16_t *flags, uint32_t *parent, unsigned long long *best_key, int tid, int nt, Sync sync)
	{
		const int W = S.W, H = S.H, T = W * H;
		const uint32_t magicW = S.magicW;
		const uint32_t lane = (uint32_t)tid & 31u;
		if (tid == 0)
			*best_key = 0ull;
		// thresholds are rtos16(-7), (1), (6), (0.35)
		for (int i = tid; i < T; i += nt)
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
			if (S.radius >= 0)
			{
				const float px = (float)(x - S.ox), py = (float)(y - S.oy);
				const float l2 = __fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py));
				if (l2 > (float)(S.radius * S.radius))
					f = F_INVALID;
			}
			flags[i] = f;
			parent[i] = (uint32_t)i;
		}
		sync();

		// ---- fillUnreachable as union-find (root = lowest tile index = discovery order of the reference's flood) ----
		// (1) horizontal runs by warp ballot: every valid tile points at the start of its run inside its 32-tile chunk
		//     (or at the last tile of the previous chunk when the run continues there): chains are <= ceil(W/32)+1 deep
		//     and no atomics are needed; (2) one union per pair of vertically overlapping runs, not per tile.
		for (int base = 0; base < T; base += nt)
		{
			const int i = base + tid;
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
				const uint32_t stops = ~lm & (0xFFFFFFFFu >> (31u - lane)); // lanes <= mine that start a run
				parent[i] = stops ? (uint32_t)(i - (int)lane + (31 - __clz((int)stops))) : (uint32_t)(i - (int)lane - 1);
			}
		}
		sync();
		for (int i = tid; i < T - W; i += nt)
		{
			if (flags[i] == F_INVALID || flags[i + W] == F_INVALID)
				continue;
			const int y = (int)div_w((uint32_t)i, magicW), x = i - y * W;
			// the pair to the left already joined the same two runs
			if (x > 0 && flags[i - 1] != F_INVALID && flags[i + W - 1] != F_INVALID)
				continue;
			uf_union(parent, (uint32_t)i, (uint32_t)(i + W));
		}
		sync();
		// flatten, then count members in the upper 16 bits of the root's word (T < 65536)
		for (int i = tid; i < T; i += nt)
			if (flags[i] != F_INVALID)
			{
				const uint32_t r = uf_find(parent, (uint32_t)i);
				if (r != (uint32_t)i)
					parent[i] = r; // benign: only shortens chains, roots are untouched
			}
		sync();
		for (int base = 0; base < T; base += nt)
		{
			const int i = base + tid;
			const bool valid = i < T && flags[i] != F_INVALID;
			const uint32_t r = valid ? (parent[i] & 0xFFFFu) : 0xFFFFFFFFu;
			// one atomic per distinct root per warp
			const uint32_t peers = __match_any_sync(FULL_MASK, r);
			if (valid && (uint32_t)(__ffs((int)peers) - 1) == lane)
				atomicAdd(&parent[r], (uint32_t)__popc(peers) << 16);
		}
		sync();
		// largest count; ties -> first discovered = smallest root (grid.cpp:68-72 uses strict '>')
		{
			unsigned long long key = 0ull;
			for (int i = tid; i < T; i += nt)
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
			if (lane == 0 && key)
				atomicMax(best_key, key);
		}
		sync();
		const unsigned long long bk = *best_key;
		const bool any_valid = bk != 0ull;
		const uint32_t best_root = 0xFFFFFFFFu - (uint32_t)(bk & 0xFFFFFFFFull);
		for (int i = tid; i < T; i += nt)
			if (flags[i] == F_INVALID || !any_valid || (parent[i] & 0xFFFFu) != best_root)
				flags[i] = F_INVALID;
		sync();

		// ---- findBorder + Mana (border bits only land on invalid tiles, so in-place is safe) ----
		for (int i = tid; i < T; i += nt)
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
		sync();
		return any_valid;
	}

	// newWaypoints spawn picking: paths.cpp:173-183, 190-196 (inherently sequential in the RNG stream). One thread.
	// out: [nwp] tile indices (MTD_INF when the map has no valid tile), shared or global memory.
	__device__ __forceinline__ void mapgen_pick_spawns(Rng g, int T, uint16_t *flags, int nwp, bool any_valid, uint32_t *out, int *status)
	{
		for (int k = 0; k < nwp; k++)
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
			out[k] = t;
		}
		if (!any_valid && nwp > 0) // without spawn picking an empty map is a legitimate newGrid result
			atomicOr(status, 2 /* ST_NO_VALID */);
	}
}
