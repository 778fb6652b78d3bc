// walk.cuh -- Waypoints::find on registers and the full-path walk of Waypoints::update
// (sources/map/paths.cpp:87-108, 141-169), written once over an accessor so the same code walks layers held in
// shared memory (one-CTA kernel, single map) and layers resident in global memory / L2 (batched maps: a dedicated
// walker warp chases the previous map while the CTA relaxes the next one; large grids: k_walk_global).
//
// Why the walk does not call find() on every hop: between two waypoint tiles the visited bits are constant, so
// every permutation cost is A_{p0}(pos) + const. One hop along the chosen layer's next-hop lowers the chosen cost
// by exactly w and every other cost by at most w (distances are shortest, the hop is an edge of weight w), so the
// first-strict-minimum in next_permutation order -- what find() returns -- cannot change until the bits do.
// neighborDistance(prev, tile) along that chain equals dist[prev] - dist[tile], so the accumulated distance of a
// segment telescopes. Both arguments need sums that do not wrap: they are applied only when every table entry and
// every A_i involved is below 2^27 (n <= 8 terms, two sums); otherwise the reference's per-hop evaluation runs.
#pragma once

#include "common.cuh"
#include "sssp_warp.cuh"

namespace mtd
{
	constexpr uint32_t WALK_FAST_LIMIT = 1u << 27;

	// Waypoints::find on registers: unvisited ids come from bit scans (no local array), k <= 3 is unrolled in
	// lexicographic permutation order, larger k falls back to find_best's next_permutation loop.
	template<class DistAtPos>
	__device__ __forceinline__ FindResult find_small(int n, uint32_t bits, const uint32_t *D, DistAtPos A)
	{
		uint32_t m = ~bits & (n >= 32 ? 0xFFFFFFFFu : ((1u << n) - 1u));
		const int k = __popc(m);
		if (k > 3)
			return find_best(n, bits, D, A);
		FindResult r;
		r.layer = MTD_INF;
		r.distance = MTD_INF;
		if (k == 0)
			return r;
		const uint32_t a = (uint32_t)__ffs((int)m) - 1u;
		m &= m - 1u;
		const uint32_t Aa = A(a);
		if (k == 1)
		{
			if (Aa < r.distance)
			{
				r.distance = Aa;
				r.layer = a;
			}
			return r;
		}
		const uint32_t b = (uint32_t)__ffs((int)m) - 1u;
		m &= m - 1u;
		const uint32_t Ab = A(b);
#define MTD_D(p, q) D[(p) * FIND_MAXN + (q)]
		if (k == 2)
		{
			const uint32_t c0 = Aa + MTD_D(a, b), c1 = Ab + MTD_D(b, a);
			if (c0 < r.distance)
			{
				r.distance = c0;
				r.layer = a;
			}
			if (c1 < r.distance)
			{
				r.distance = c1;
				r.layer = b;
			}
			return r;
		}
		const uint32_t c = (uint32_t)__ffs((int)m) - 1u;
		const uint32_t Ac = A(c);
		const uint32_t dab = MTD_D(a, b), dac = MTD_D(a, c), dba = MTD_D(b, a), dbc = MTD_D(b, c), dca = MTD_D(c, a), dcb = MTD_D(c, b);
#undef MTD_D
		const uint32_t cost[6] = { Aa + dab + dbc, Aa + dac + dcb, Ab + dba + dac, Ab + dbc + dca, Ac + dca + dab, Ac + dcb + dba };
		const uint32_t first[6] = { a, a, b, b, c, c };
#pragma unroll
		for (int p = 0; p < 6; p++)
			if (cost[p] < r.distance)
			{
				r.distance = cost[p];
				r.layer = first[p];
			}
		return r;
	}

	// true if every entry of the n x n inter-waypoint table allows the segment walk
	__device__ __forceinline__ bool walk_table_fast(int n, const uint32_t *D)
	{
		bool ok = true;
		for (int a = 0; a < n; a++)
			for (int b = 0; b < n; b++)
				ok = ok && D[a * FIND_MAXN + b] < WALK_FAST_LIMIT;
		return ok;
	}

	struct WalkResult
	{
		uint32_t len, distance;
	};

	// Acc: uint32_t dist(layer, tile); uint32_t hop(layer, tile) = tiles_layer[tile]; uint32_t edge(a, b) =
	// Grid::neighborDistance of adjacent tiles. wp[n] waypoint tiles, D the inter-waypoint table.
	template<class Acc>
	__device__ WalkResult walk_path(int n, const uint32_t *wp, const uint32_t *D, bool table_fast, int spawner, const Acc &acc, uint32_t *path, uint32_t cap, uint32_t maxlen)
	{
		uint32_t bits = 0, prev = MTD_INF, distance = 0, len = 0;
		uint32_t tile = wp[spawner];
		auto hits = [&](uint32_t t)
		{
			uint32_t h = 0;
			for (int k = 0; k < n; k++)
				if (wp[k] == t)
					h |= 1u << k;
			return h;
		};
		auto push = [&](uint32_t t)
		{
			if (path && len < cap)
				path[len] = t;
			len++;
		};
		while (tile != MTD_INF) // paths.cpp:95
		{
			bits |= hits(tile); // :97-99
			push(tile);         // :100
			if (prev != MTD_INF)
				distance += acc.edge(prev, tile); // :101-102
			prev = tile;
			const FindResult fr = find_small(n, bits, D, [&](uint32_t i) { return acc.dist(i, tile); }); // :104
			if (fr.layer == MTD_INF)
				break;
			bool fast = table_fast;
			if (fast)
			{
				uint32_t m = ~bits & ((1u << n) - 1u);
				while (m)
				{
					const uint32_t i = (uint32_t)__ffs((int)m) - 1u;
					m &= m - 1u;
					fast = fast && acc.dist(i, tile) < WALK_FAST_LIMIT;
				}
			}
			const uint32_t l = fr.layer;
			if (!fast)
			{
				tile = acc.hop(l, tile);
				if (len > maxlen)
					break; // cannot happen on consistent layers; guards against a hang on corrupted input
				continue;
			}
			// ---- segment: follow layer l until the next waypoint tile (or a dead end) ----
			const uint32_t d0 = acc.dist(l, tile);
			uint32_t t = tile;
			for (;;)
			{
				const uint32_t nxt = acc.hop(l, t);
				if (nxt == MTD_INF)
				{
					distance += d0 - acc.dist(l, t);
					tile = MTD_INF;
					break;
				}
				const uint32_t h = hits(nxt);
				if (h)
				{
					// dist_l[nxt] from the table: nxt is the tile of waypoint j
					const uint32_t j = (uint32_t)__ffs((int)h) - 1u;
					distance += d0 - D[j * FIND_MAXN + l];
					tile = nxt;
					prev = MTD_INF; // this segment's edges are already accounted
					break;
				}
				push(nxt);
				t = nxt;
				if (len > maxlen)
				{
					tile = MTD_INF;
					break;
				}
			}
		}
		WalkResult r;
		r.len = len;
		r.distance = distance;
		return r;
	}

	// layers resident in global memory (read through L2: another warp / kernel wrote them)
	struct GlobalLayersAcc
	{
		const uint32_t *gdist, *gtiles; // [n][T]
		const int16_t *gelev;           // [T]
		size_t T;
		__device__ __forceinline__ uint32_t dist(uint32_t l, uint32_t t) const { return __ldcg(gdist + (size_t)l * T + t); }
		__device__ __forceinline__ uint32_t hop(uint32_t l, uint32_t t) const { return __ldcg(gtiles + (size_t)l * T + t); }
		__device__ __forceinline__ uint32_t edge(uint32_t a, uint32_t b) const { return edge_weight(gelev[a], gelev[b]); }
	};

	// aggregates of Waypoints::update, paths.cpp:122-138
	__device__ __forceinline__ void walk_aggregates(int n, const uint32_t *fd, uint32_t *agg3)
	{
		uint32_t sum = 0;
		for (int i = 0; i < n; i++)
			sum += fd[i];
		const float avg = __fdiv_rn(stor((int32_t)sum), (float)(unsigned long long)n);
		uint32_t best = 0;
		for (int i = 0; i < n; i++)
			if (fd[i] < fd[best])
				best = (uint32_t)i;
		agg3[0] = (uint32_t)__float2ll_rz(avg);
		agg3[1] = (uint32_t)__float2ll_rz(stor((int32_t)fd[best]));
		agg3[2] = best;
	}
}
