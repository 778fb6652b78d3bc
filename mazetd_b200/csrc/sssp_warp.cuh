// sssp_warp.cuh -- warp-level shortest-path engine on a shared-memory tile, plus the pure functions that
// turn final distances into the reference's outputs (next-hop tiles, Waypoints::find).
//
// What the reference computes (sources/map/paths.cpp:24-78): single-source Dijkstra over the 4-connected grid,
// heap ordered by (distance, tile). Its result has a closed form:
//   distances[v] = exact shortest distance (unique fixpoint of relaxation, any schedule reaches it), and
//   tiles[v]     = the neighbour u with d[u] + w(u,v) == d[v] that minimises (d[u], u) lexicographically,
// because all weights are >= 255 > 0, so the heap settles nodes in ascending (d, idx) order and only the first
// settler to reach d[v] passes the strict '<' (paths.cpp:69).
//
// How it is computed here: a label-correcting wavefront run by ONE WARP on a tile held in shared memory.
//   * pending cells live in a circular queue (packed y<<8|x, uint16) + an "in queue" bitmap;
//   * a round settles every pending cell with d < dmin + 255 (255 = lower bound of every edge weight,
//     sources/map/grid.cpp:141-145, so those labels are already final -- the delta-stepping bucket with
//     delta = min weight, which needs no re-relaxation), relaxes their 4 neighbours with shared-memory
//     atomicMin, and compacts survivors / newly reached cells back into the queue with warp ballots;
//   * blocked tiles are encoded as distance 0 during the run: no relaxation (>= 255) can ever lower them, so
//     the inner loop has no flag tests. A blocked SOURCE legitimately has distance 0 and still expands, which
//     is exactly the reference's behaviour (paths.cpp:63-67 test only the neighbour, never the popped tile).
// Correctness never depends on the settle order (any improvement re-queues the cell); the order only makes
// the work near-optimal. Queue overflow is handled by leaving the cell marked in the bitmap and re-collecting
// marked cells when the queue drains.
#pragma once

#include "common.cuh"

namespace mtd
{
	constexpr uint32_t FULL_MASK = 0xFFFFFFFFu;
	constexpr uint32_t MIN_EDGE = 255u;

	template<typename WT, int QCAP>
	struct WarpTile
	{
		uint32_t *dist;  // [W*H]   shared
		const WT *wr;    // [W*H]   w(c, c+1) - 255
		const WT *wd;    // [W*H]   w(c, c+W) - 255
		uint32_t *inq;   // [ceil(W*H/32)] shared bitmap
		uint16_t *q;     // [QCAP]  shared circular queue
		int W, H;
	};

	__device__ __forceinline__ uint16_t pack_yx(int y, int x)
	{
		return (uint16_t)((y << 8) | x);
	}

	// Seeds must already be in q[0..n), marked in inq, with their distances stored; dmin = their minimum.
	// Returns true if some distance hit the uint32 guard band (reference arithmetic would wrap).
	template<typename WT, int QCAP>
	__device__ bool warp_relax(const WarpTile<WT, QCAP> t, uint32_t n, uint32_t dmin, bool dropped = false)
	{
		static_assert((QCAP & (QCAP - 1)) == 0, "QCAP must be a power of two");
		constexpr uint32_t QMASK = QCAP - 1;
		const uint32_t lane = lane_id();
		const uint32_t lt = lanemask_lt();
		const int W = t.W, H = t.H;
		const int nwords = (W * H + 31) >> 5;
		uint32_t head = 0, tail = n;
		// dropped (warp-uniform): some marked cell is not in the queue and must be re-collected from the bitmap
		bool guard = false;

		for (;;)
		{
			if (head == tail)
			{
				if (!dropped)
					break;
				// re-collect cells that are marked pending but were dropped from the queue
				dropped = false;
				uint32_t m = MTD_INF;
				for (int w0 = 0; w0 < nwords && !dropped; w0 += 32)
				{
					const int wi = w0 + (int)lane;
					uint32_t bits = wi < nwords ? t.inq[wi] : 0u;
					while (__any_sync(FULL_MASK, bits != 0))
					{
						const bool has = bits != 0;
						const uint32_t cell = (uint32_t)wi * 32u + (has ? (uint32_t)__ffs((int)bits) - 1u : 0u);
						const uint32_t bal = __ballot_sync(FULL_MASK, has);
						const uint32_t room = (uint32_t)QCAP - (tail - head);
						const uint32_t rank = __popc(bal & lt);
						if (has && rank < room)
						{
							t.q[(tail + rank) & QMASK] = pack_yx((int)(cell / (uint32_t)W), (int)(cell % (uint32_t)W));
							m = min(m, t.dist[cell]);
							bits &= bits - 1;
						}
						const uint32_t cnt = __popc(bal);
						tail += min(cnt, room);
						if (cnt > room)
						{
							dropped = true;
							break;
						}
					}
				}
				dmin = warp_min_u32(m);
				__syncwarp();
				continue;
			}

			const uint32_t thr = dmin >= MTD_GUARD ? MTD_INF : dmin + MIN_EDGE;
			uint32_t newmin = MTD_INF;
			const uint32_t round_head = head;
			const uint32_t round_n = tail - head;
			for (uint32_t base = 0; base < round_n; base += 32)
			{
				const uint32_t idx = base + lane;
				const bool active = idx < round_n;
				const uint32_t e = active ? t.q[(round_head + idx) & QMASK] : 0u;
				const int y = (int)(e >> 8), x = (int)(e & 255u);
				const uint32_t c = (uint32_t)(y * W + x);
				const uint32_t d = active ? t.dist[c] : MTD_INF;
				const bool settle = active && d < thr;
				const bool keep = active && !settle;
				if (settle)
					atomicAnd(&t.inq[c >> 5], ~(1u << (c & 31u)));
				// every lane has read its entry and cleared its mark before anyone relaxes or overwrites slots
				__syncwarp();
				const uint32_t rd = round_head + min(base + 32u, round_n); // slots before rd are free
				{
					const uint32_t kb = __ballot_sync(FULL_MASK, keep);
					if (keep)
					{
						t.q[(tail + __popc(kb & lt)) & QMASK] = (uint16_t)e;
						newmin = min(newmin, d);
					}
					tail += __popc(kb);
				}
#pragma unroll
				for (int dir = 0; dir < 4; dir++)
				{
					// 0: left, 1: right, 2: up, 3: down
					const int nx = x + (dir == 0 ? -1 : dir == 1 ? 1 : 0);
					const int ny = y + (dir == 2 ? -1 : dir == 3 ? 1 : 0);
					const bool ok = settle && nx >= 0 && nx < W && ny >= 0 && ny < H;
					bool fresh = false;
					if (ok)
					{
						const uint32_t nb = (uint32_t)(ny * W + nx);
						const uint32_t wv = dir == 0 ? t.wr[c - 1] : dir == 1 ? t.wr[c] : dir == 2 ? t.wd[c - W] : t.wd[c];
						const uint32_t nd = d + MIN_EDGE + wv;
						if (nd >= MTD_GUARD)
							guard = true;
						else
						{
							const uint32_t old = atomicMin(&t.dist[nb], nd);
							if (nd < old)
							{
								newmin = min(newmin, nd);
								const uint32_t bit = 1u << (nb & 31u);
								fresh = (atomicOr(&t.inq[nb >> 5], bit) & bit) == 0;
							}
						}
					}
					const uint32_t pb = __ballot_sync(FULL_MASK, fresh);
					if (pb)
					{
						uint32_t cnt = __popc(pb);
						const uint32_t room = (uint32_t)QCAP - (tail - rd);
						const uint32_t rank = __popc(pb & lt);
						if (fresh && rank < room)
							t.q[(tail + rank) & QMASK] = pack_yx(ny, nx);
						if (cnt > room)
						{
							dropped = true; // the cell stays marked in inq and is re-collected later
							cnt = room;
						}
						tail += cnt;
					}
				}
			}
			head = round_head + round_n;
			dmin = warp_min_u32(newmin);
			__syncwarp();
		}
		return __any_sync(FULL_MASK, guard);
	}

	// tiles[v] from final distances (blocked non-source tiles must already read MTD_INF).
	// Neighbours are visited in ascending tile index (up, left, right, down) and replace only on a strictly
	// smaller d[u], which yields the lexicographic minimum of (d[u], u). Returns a LOCAL index or MTD_M.
	template<typename WT>
	__device__ __forceinline__ uint32_t next_hop(const uint32_t *dist, const WT *wr, const WT *wd, int W, int H, int x, int y)
	{
		const uint32_t c = (uint32_t)(y * W + x);
		const uint32_t dv = dist[c];
		if (dv == MTD_INF || dv == 0u) // unreachable / blocked, or a source (paths.cpp:49: nothing is < 0)
			return MTD_INF;
		uint32_t best = MTD_INF, bestd = MTD_INF;
		if (y > 0)
		{
			const uint32_t du = dist[c - W];
			if (du != MTD_INF && du + MIN_EDGE + wd[c - W] == dv && du < bestd)
			{
				bestd = du;
				best = c - W;
			}
		}
		if (x > 0)
		{
			const uint32_t du = dist[c - 1];
			if (du != MTD_INF && du + MIN_EDGE + wr[c - 1] == dv && du < bestd)
			{
				bestd = du;
				best = c - 1;
			}
		}
		if (x < W - 1)
		{
			const uint32_t du = dist[c + 1];
			if (du != MTD_INF && du + MIN_EDGE + wr[c] == dv && du < bestd)
			{
				bestd = du;
				best = c + 1;
			}
		}
		if (y < H - 1)
		{
			const uint32_t du = dist[c + W];
			if (du != MTD_INF && du + MIN_EDGE + wd[c] == dv && du < bestd)
			{
				bestd = du;
				best = c + W;
			}
		}
		return best;
	}

	// ---- Waypoints::find, sources/map/paths.cpp:141-169 ------------------------------------------------
	// cost(perm) = dist_{p0}[pos] + dist_{p1}[wp_{p0}] + ... in uint32 (wraps like the reference); permutations
	// in std::next_permutation (lexicographic) order over the ascending list of unvisited ids; first strictly
	// smaller wins. A(i) = dist_i[pos]; D[a * MAXN + b] = dist_b[wp_a] (inter-waypoint table).
	constexpr int FIND_MAXN = 8;

	struct FindResult
	{
		uint32_t layer; // first waypoint of the best permutation (caller looks up tiles_layer[pos]); MTD_M if none
		uint32_t distance;
	};

	template<class DistAtPos>
	__device__ FindResult find_best(int n, uint32_t bits, const uint32_t *D, DistAtPos A)
	{
		FindResult r;
		r.layer = MTD_INF;
		r.distance = MTD_INF;
		uint8_t id[FIND_MAXN];
		int k = 0;
		for (int i = 0; i < n; i++)
			if ((bits & (1u << i)) == 0)
				id[k++] = (uint8_t)i;
		if (k == 0)
			return r;
		if (k <= 3)
		{
			// unrolled lexicographic enumeration, registers only
			const uint32_t a = id[0], b = k > 1 ? id[1] : 0, c = k > 2 ? id[2] : 0;
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
			const uint32_t Ab = A(b);
			if (k == 2)
			{
				const uint32_t c0 = Aa + D[a * FIND_MAXN + b];
				const uint32_t c1 = Ab + D[b * FIND_MAXN + a];
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
			const uint32_t Ac = A(c);
#define MTD_D(p, q) D[(p) * FIND_MAXN + (q)]
			const uint32_t cost[6] = {
				Aa + MTD_D(a, b) + MTD_D(b, c), // a b c
				Aa + MTD_D(a, c) + MTD_D(c, b), // a c b
				Ab + MTD_D(b, a) + MTD_D(a, c), // b a c
				Ab + MTD_D(b, c) + MTD_D(c, a), // b c a
				Ac + MTD_D(c, a) + MTD_D(a, b), // c a b
				Ac + MTD_D(c, b) + MTD_D(b, a), // c b a
			};
#undef MTD_D
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
		// general case: explicit next_permutation
		for (;;)
		{
			uint32_t dist = A(id[0]);
			for (int j = 1; j < k; j++)
				dist += D[id[j - 1] * FIND_MAXN + id[j]];
			if (dist < r.distance)
			{
				r.distance = dist;
				r.layer = id[0];
			}
			int i = k - 1;
			while (i > 0 && id[i - 1] >= id[i])
				i--;
			if (i <= 0)
				break;
			int j = k - 1;
			while (id[j] <= id[i - 1])
				j--;
			uint8_t tmp = id[i - 1];
			id[i - 1] = id[j];
			id[j] = tmp;
			for (int l = i, rr = k - 1; l < rr; l++, rr--)
			{
				tmp = id[l];
				id[l] = id[rr];
				id[rr] = tmp;
			}
		}
		return r;
	}
}
