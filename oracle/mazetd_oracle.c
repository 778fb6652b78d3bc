/*
 * mazetd_oracle.c -- TEST INFRASTRUCTURE. See mazetd_oracle.h for scope and pinning.
 *
 * Plain-C restatement of the reference algorithm; every function cites the reference lines it follows.
 * Build: oracle/Makefile (-O3, x86-64 baseline, -ffp-contract=off: float32 ops are IEEE, never fused).
 */
#include "mazetd_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#include "fnl_subset.h"

/* =============================================================================================
 * RNG -- externals/cage/sources/libcore/math/random.cpp
 * ============================================================================================= */
static inline uint64_t rotl64(uint64_t x, int k)
{
	return (x << k) | (x >> (64 - k));
}

/* random.cpp:27-38 xoroshiro128** */
uint64_t mo_rng_next(mo_rng *g)
{
	const uint64_t s0 = g->s[0];
	uint64_t s1 = g->s[1];
	const uint64_t result = rotl64(s0 * 5, 7) * 9;
	s1 ^= s0;
	g->s[0] = rotl64(s0, 24) ^ s1 ^ (s1 << 16);
	g->s[1] = rotl64(s1, 37);
	return result;
}

/* random.cpp:45-51 */
double mo_rng_chance_double(mo_rng *g)
{
	const uint64_t r = mo_rng_next(g);
	return (double)(r >> 11) * 0x1.0p-53;
}

/* random.cpp:40-43: Real(double) rounds to float32 */
float mo_rng_chance(mo_rng *g)
{
	return (float)mo_rng_chance_double(g);
}

/* random.cpp:53-66 (modulo bias kept) */
uint32_t mo_rng_range_u32(mo_rng *g, uint32_t min, uint32_t max)
{
	if (min == max)
		return min;
	const uint64_t range = (uint64_t)max - (uint64_t)min;
	const uint64_t mod = mo_rng_next(g) % range;
	return (uint32_t)(mod + min);
}

int32_t mo_rng_range_i32(mo_rng *g, int32_t min, int32_t max)
{
	if (min == max)
		return min;
	const uint64_t range = (uint64_t)(int64_t)(max - min);
	const uint64_t mod = mo_rng_next(g) % range;
	return (int32_t)(mod + (uint64_t)(int64_t)min);
}

/* random.cpp:69-79 with interpolate() of externals/cage/sources/include/cage-core/math.h:1172-1176 */
float mo_rng_range_f(mo_rng *g, float min, float max)
{
	const float c = mo_rng_chance(g);
	return min + (max - min) * c;
}

/* =============================================================================================
 * fixed point -- sources/common.h:15-29
 * ============================================================================================= */
int16_t mo_rtos16(float v)
{
	return (int16_t)(v * 255.0f);
}

int32_t mo_rtos32(float v)
{
	return (int32_t)(v * 255.0f);
}

float mo_stor(int32_t v)
{
	const float n = (float)(1.0 / 255); /* constexpr Real n = 1.0 / 255 : rounded to float once */
	return (float)v * n;
}

/* =============================================================================================
 * Grid geometry -- sources/map/grid.cpp:100-145
 * ============================================================================================= */
/* grid.cpp:100-106 */
uint32_t mo_grid_index_i(const mo_grid *g, int32_t x, int32_t y)
{
	const int32_t ax = x + g->ox, ay = y + g->oy;
	if (ax < 0 || ax >= g->W || ay < 0 || ay >= g->H)
		return MO_M;
	return (uint32_t)(ay * g->W + ax);
}

/* grid.cpp:108-111: index(Vec2i(pos + 200.5) - 200), numeric_cast truncates toward zero */
uint32_t mo_grid_index_f(const mo_grid *g, float x, float y)
{
	const int32_t ix = (int32_t)(x + 200.5f) - 200;
	const int32_t iy = (int32_t)(y + 200.5f) - 200;
	return mo_grid_index_i(g, ix, iy);
}

/* grid.cpp:118-123 */
void mo_grid_position(const mo_grid *g, uint32_t idx, int32_t *out2)
{
	const uint32_t y = idx / (uint32_t)g->W;
	const uint32_t x = idx % (uint32_t)g->W;
	out2[0] = (int32_t)x - g->ox;
	out2[1] = (int32_t)y - g->oy;
}

/* grid.cpp:125-129 */
void mo_grid_center(const mo_grid *g, uint32_t idx, float *out3)
{
	int32_t p[2];
	mo_grid_position(g, idx, p);
	out3[0] = (float)p[0];
	out3[1] = mo_stor(g->elev[idx]);
	out3[2] = (float)p[1];
}

/* grid.cpp:141-145: rtos32(distance(center(a), center(b)));
 * distance = sqrt(dot(d,d)), dot sums ((0 + d0*d0) + d1*d1) + d2*d2 (cage-core/math.h:919-926) */
uint32_t mo_neighbor_distance(const mo_grid *g, uint32_t a, uint32_t b)
{
	float ca[3], cb[3];
	mo_grid_center(g, a, ca);
	mo_grid_center(g, b, cb);
	const float d0 = ca[0] - cb[0], d1 = ca[1] - cb[1], d2 = ca[2] - cb[2];
	float sum = 0;
	sum += d0 * d0;
	sum += d1 * d1;
	sum += d2 * d2;
	return (uint32_t)mo_rtos32(sqrtf(sum));
}

void mo_edge_weights(const mo_grid *g, uint32_t *right, uint32_t *down)
{
	for (int32_t y = 0; y < g->H; y++)
		for (int32_t x = 0; x < g->W; x++)
		{
			const uint32_t i = (uint32_t)y * g->W + x;
			right[i] = x + 1 < g->W ? mo_neighbor_distance(g, i, i + 1) : 0;
			down[i] = y + 1 < g->H ? mo_neighbor_distance(g, i, i + g->W) : 0;
		}
}

/* =============================================================================================
 * Directions::update -- sources/map/paths.cpp:24-78
 * min-heap of (distance, tile) pairs ordered like std::greater<std::pair<uint32,uint32>> (paths.cpp:16-17):
 * key = distance << 32 | tile, smallest first.
 * ============================================================================================= */
typedef struct mo_heap
{
	uint64_t *a;
	size_t n, cap;
} mo_heap;

static int heap_push(mo_heap *h, uint64_t k)
{
	if (h->n == h->cap)
	{
		size_t nc = h->cap ? h->cap * 2 : 1024;
		uint64_t *na = (uint64_t *)realloc(h->a, nc * sizeof(uint64_t));
		if (!na)
			return -1;
		h->a = na;
		h->cap = nc;
	}
	size_t i = h->n++;
	while (i > 0)
	{
		const size_t p = (i - 1) >> 1;
		if (h->a[p] <= k)
			break;
		h->a[i] = h->a[p];
		i = p;
	}
	h->a[i] = k;
	return 0;
}

static uint64_t heap_pop(mo_heap *h)
{
	const uint64_t top = h->a[0];
	const uint64_t k = h->a[--h->n];
	size_t i = 0;
	for (;;)
	{
		size_t c = 2 * i + 1;
		if (c >= h->n)
			break;
		if (c + 1 < h->n && h->a[c + 1] < h->a[c])
			c++;
		if (h->a[c] >= k)
			break;
		h->a[i] = h->a[c];
		i = c;
	}
	if (h->n)
		h->a[i] = k;
	return top;
}

int mo_directions_update(const mo_grid *g, const uint32_t *sources, int nsrc, uint32_t *tiles, uint32_t *dist)
{
	const size_t total = (size_t)g->W * (size_t)g->H;
	for (size_t i = 0; i < total; i++) /* paths.cpp:40-43 */
		tiles[i] = dist[i] = MO_M;
	uint8_t *visited = (uint8_t *)calloc(total ? total : 1, 1);
	if (!visited)
		return -1;
	mo_heap q = { 0, 0, 0 };
	for (int s = 0; s < nsrc; s++) /* paths.cpp:48-49; nsrc > 1 = extension */
	{
		if (heap_push(&q, (uint64_t)sources[s]))
			goto oom;
		dist[sources[s]] = 0;
	}
	static const int32_t offs[4][2] = { { -1, 0 }, { 1, 0 }, { 0, -1 }, { 0, 1 } }; /* paths.cpp:60 */
	const uint16_t blocking = MO_FLAG_INVALID | MO_FLAG_BUILDING;                 /* paths.cpp:65 */
	while (q.n)
	{
		const uint32_t id = (uint32_t)heap_pop(&q);
		if (visited[id]) /* paths.cpp:55-57 */
			continue;
		visited[id] = 1;
		int32_t mp[2];
		mo_grid_position(g, id, mp);
		const uint32_t md = dist[id];
		for (int k = 0; k < 4; k++)
		{
			const uint32_t ni = mo_grid_index_i(g, mp[0] + offs[k][0], mp[1] + offs[k][1]);
			if (ni == MO_M)
				continue;
			if (g->flags[ni] & blocking)
				continue;
			const uint32_t nd = md + mo_neighbor_distance(g, id, ni); /* uint32 wrap kept */
			if (nd < dist[ni])                                        /* strict, paths.cpp:69 */
			{
				dist[ni] = nd;
				tiles[ni] = id;
				if (heap_push(&q, ((uint64_t)nd << 32) | ni))
					goto oom;
			}
		}
	}
	free(q.a);
	free(visited);
	return 0;
oom:
	free(q.a);
	free(visited);
	return -1;
}

/* =============================================================================================
 * Waypoints -- sources/map/paths.cpp:82-169
 * ============================================================================================= */
static int next_permutation_u32(uint32_t *a, int n)
{
	/* std::next_permutation: lexicographic successor, returns 0 after wrapping to sorted order */
	int i = n - 1;
	while (i > 0 && a[i - 1] >= a[i])
		i--;
	if (i <= 0)
	{
		for (int l = 0, r = n - 1; l < r; l++, r--)
		{
			uint32_t t = a[l];
			a[l] = a[r];
			a[r] = t;
		}
		return 0;
	}
	int j = n - 1;
	while (a[j] <= a[i - 1])
		j--;
	uint32_t t = a[i - 1];
	a[i - 1] = a[j];
	a[j] = t;
	for (int l = i, r = n - 1; l < r; l++, r--)
	{
		t = a[l];
		a[l] = a[r];
		a[r] = t;
	}
	return 1;
}

/* paths.cpp:141-169 */
void mo_waypoints_find(const mo_waypoints *w, uint32_t pos0, uint32_t bits, uint32_t *out2)
{
	const size_t T = (size_t)w->grid->W * (size_t)w->grid->H;
	uint32_t indices[32];
	int k = 0;
	for (int i = 0; i < w->n; i++)
		if ((bits & (1u << i)) == 0)
			indices[k++] = (uint32_t)i;
	out2[0] = MO_M;
	out2[1] = MO_M;
	if (k == 0)
		return;
	do
	{
		uint32_t pos = pos0;
		uint32_t dist = 0;
		for (int j = 0; j < k; j++)
		{
			const uint32_t i = indices[j];
			dist += w->dist[i * T + pos]; /* uint32 wrap kept */
			pos = w->wp_tiles[i];
		}
		if (dist < out2[1])
		{
			out2[1] = dist;
			out2[0] = w->tiles[indices[0] * T + pos0];
		}
	} while (next_permutation_u32(indices, k));
}

/* paths.cpp:87-108 */
uint32_t mo_waypoints_fullpath(const mo_waypoints *w, int spawner, uint32_t *buf, uint32_t cap, uint32_t *full_distance)
{
	uint32_t bits = 0, prev = MO_M, distance = 0, len = 0;
	uint32_t tile = w->wp_tiles[spawner];
	while (tile != MO_M)
	{
		for (int i = 0; i < w->n; i++)
			if (w->wp_tiles[i] == tile)
				bits |= 1u << i;
		if (buf && len < cap)
			buf[len] = tile;
		len++;
		if (prev != MO_M)
			distance += mo_neighbor_distance(w->grid, prev, tile);
		prev = tile;
		uint32_t r[2];
		mo_waypoints_find(w, tile, bits, r);
		tile = r[0];
	}
	*full_distance = distance;
	return len;
}

/* paths.cpp:122-138 */
void mo_waypoints_aggregates(const uint32_t *full_distance, int n, uint32_t *out3)
{
	uint32_t sum = 0;
	for (int i = 0; i < n; i++)
		sum += full_distance[i];
	/* (stor(sum) / waypoints.size()).value -> uint32 : stor takes sint32; size() converts to Real */
	out3[0] = (uint32_t)(mo_stor((int32_t)sum) / (float)(uint64_t)n);
	int best = 0;
	for (int i = 0; i < n; i++)
		if (full_distance[i] < full_distance[best])
			best = i;
	out3[1] = (uint32_t)mo_stor((int32_t)full_distance[best]);
	out3[2] = (uint32_t)best;
}

typedef struct dir_job
{
	const mo_grid *g;
	uint32_t src;
	uint32_t *tiles, *dist;
	int rc;
} dir_job;

static void *dir_job_entry(void *p)
{
	dir_job *j = (dir_job *)p;
	j->rc = mo_directions_update(j->g, &j->src, 1, j->tiles, j->dist);
	return 0;
}

/* paths.cpp:111-139 */
int mo_waypoints_update(const mo_grid *g, const uint32_t *wp_tiles, int n, int threads, uint32_t *tiles, uint32_t *dist, uint32_t *fullpath_buf, uint32_t fullpath_cap, uint32_t *fullpath_len, uint32_t *full_distance, uint32_t *agg3)
{
	if (n <= 0)
		return 0; /* paths.cpp:113-114 */
	const size_t T = (size_t)g->W * (size_t)g->H;
	dir_job jobs[32];
	pthread_t thr[32];
	for (int i = 0; i < n; i++)
	{
		jobs[i].g = g;
		jobs[i].src = wp_tiles[i];
		jobs[i].tiles = tiles + i * T;
		jobs[i].dist = dist + i * T;
		jobs[i].rc = 0;
	}
	if (threads > 1)
	{ /* tasksRunBlocking("paths directions", ..., n): one task per waypoint (paths.cpp:119) */
		for (int i = 0; i < n; i++)
			pthread_create(&thr[i], 0, dir_job_entry, &jobs[i]);
		for (int i = 0; i < n; i++)
			pthread_join(thr[i], 0);
	}
	else
		for (int i = 0; i < n; i++)
			dir_job_entry(&jobs[i]);
	for (int i = 0; i < n; i++)
		if (jobs[i].rc)
			return jobs[i].rc;
	mo_waypoints w = { g, n, wp_tiles, tiles, dist };
	for (int i = 0; i < n; i++) /* paths.cpp:120 */
		fullpath_len[i] = mo_waypoints_fullpath(&w, i, fullpath_buf ? fullpath_buf + (size_t)i * fullpath_cap : 0, fullpath_cap, &full_distance[i]);
	mo_waypoints_aggregates(full_distance, n, agg3);
	return 0;
}

/* =============================================================================================
 * procedural generator -- sources/map/procedural.cpp:13-70
 * ============================================================================================= */
/* member-initialisation order = RNG draw order (procedural.cpp:13, 24, 33, 43, 52, 61) */
void mo_procedural_init(mo_procedural *p, mo_rng *g)
{
	const float r1 = mo_rng_chance(g);
	const float r2 = mo_rng_chance(g);
	p->off_x = r1 * 1000.0f;
	p->off_y = r2 * 1000.0f;
	for (int i = 0; i < 5; i++)
		p->seeds[i] = (uint32_t)mo_rng_next(g);
	fnls_state st;
	fnls_init(&st, (int32_t)p->seeds[0]);
	st.octaves = 4;
	st.gain = 0.55f;
	fnls_calc_fractal_bounding(&st);
	p->fractal_bounding = st.fractal_bounding;
}

/* procedural.cpp:65-70; noise settings procedural.cpp:17-25 through noiseFunction.cpp:97-107 */
float mo_elevation(const mo_procedural *p, float x, float y)
{
	fnls_state st;
	fnls_init(&st, (int32_t)p->seeds[0]);
	st.frequency = 0.003f;
	st.noise_type = FNLS_NOISE_OPENSIMPLEX2;
	st.fractal_type = FNLS_FRACTAL_RIDGED;
	st.octaves = 4;
	st.lacunarity = 2.1f;
	st.gain = 0.55f;
	st.fractal_bounding = p->fractal_bounding;

	float sum = 0; /* length(pos) = sqrt(dot) */
	sum += x * x;
	sum += y * y;
	float t = sqrtf(sum) - 30.0f;
	t = t > 0.0f ? t : 0.0f; /* max(a, b) = a > b ? a : b */
	t = t * 0.2f;
	const float base = -(t * t);
	const float noise = fnls_get_noise_2d(&st, x + p->off_x, y + p->off_y) * 10.0f;
	const float v = base + noise;
	return v > -8.5f ? v : -8.5f;
}

/* =============================================================================================
 * newGrid -- sources/map/grid.cpp:12-27, 29-78, 80-97, 147-198
 * ============================================================================================= */
static inline uint32_t idx_or_m(int32_t W, int32_t H, int32_t x, int32_t y)
{
	if (x < 0 || x >= W || y < 0 || y >= H)
		return MO_M;
	return (uint32_t)(y * W + x);
}

static const int32_t OFFS[4][2] = { { -1, 0 }, { 1, 0 }, { 0, -1 }, { 0, 1 } };

/* grid.cpp:12-27 */
static int16_t max_slope(int32_t W, int32_t H, const int16_t *elev, uint32_t idx)
{
	const int16_t me = elev[idx];
	const int32_t x = (int32_t)(idx % (uint32_t)W), y = (int32_t)(idx / (uint32_t)W);
	int16_t r = 0;
	for (int k = 0; k < 4; k++)
	{
		const uint32_t i = idx_or_m(W, H, x + OFFS[k][0], y + OFFS[k][1]);
		if (i == MO_M)
			continue;
		const int16_t d = (int16_t)(me - elev[i]);
		const int16_t a = d < 0 ? (int16_t)-d : d;
		r = r > a ? r : a;
	}
	return r;
}

/* grid.cpp:160-183 */
void mo_grid_flags_from_elev(int32_t W, int32_t H, int32_t ox, int32_t oy, int32_t radius, const int16_t *elev, uint16_t *flags)
{
	const int16_t waterThreshold = mo_rtos16(-7.0f);
	const int16_t sunThreshold = mo_rtos16(1.0f);
	const int16_t windThreshold = mo_rtos16(6.0f);
	const int16_t slopeThreshold = mo_rtos16(0.35f);
	const uint32_t total = (uint32_t)W * (uint32_t)H;
	for (uint32_t i = 0; i < total; i++)
	{
		uint16_t f = 0;
		if (elev[i] < waterThreshold)
			f |= MO_FLAG_WATER;
		else if (elev[i] < sunThreshold)
			f |= MO_FLAG_SUN;
		else if (elev[i] < windThreshold)
			f |= MO_FLAG_WIND;
		else
			f |= MO_FLAG_SNOW;
		if (max_slope(W, H, elev, i) > slopeThreshold)
			f = MO_FLAG_INVALID;
		if (radius >= 0)
		{
			const float px = (float)((int32_t)(i % (uint32_t)W) - ox), py = (float)((int32_t)(i / (uint32_t)W) - oy);
			float sum = 0;
			sum += px * px;
			sum += py * py;
			if (sum > (float)(radius * radius)) /* lengthSquared(Vec2(position(i))) > sqr(45) */
				f = MO_FLAG_INVALID;
		}
		flags[i] = f;
	}
}

/* grid.cpp:29-78: DFS components over tiles with flags != Invalid (equality), ids in uint16 with 0xFFFF sentinel */
void mo_fill_unreachable(int32_t W, int32_t H, uint16_t *flags)
{
	const uint32_t total = (uint32_t)W * (uint32_t)H;
	uint16_t *comps = (uint16_t *)malloc((size_t)total * 2 + 2);
	uint32_t *stack = (uint32_t *)malloc((size_t)total * 4 * 4 + 16);
	for (uint32_t i = 0; i < total; i++)
		comps[i] = 0xFFFF;
	uint32_t largestId = MO_M, largestCnt = 0, c = 0;
	for (uint32_t s = 0; s < total; s++)
	{
		if (flags[s] == MO_FLAG_INVALID)
			continue;
		if (comps[s] != 0xFFFF)
			continue;
		comps[s] = (uint16_t)c;
		uint32_t cnt = 1;
		size_t sp = 0;
		stack[sp++] = s;
		while (sp)
		{
			const uint32_t i = stack[--sp];
			const int32_t x = (int32_t)(i % (uint32_t)W), y = (int32_t)(i / (uint32_t)W);
			for (int k = 0; k < 4; k++)
			{
				const uint32_t j = idx_or_m(W, H, x + OFFS[k][0], y + OFFS[k][1]);
				if (j == MO_M)
					continue;
				if (flags[j] == MO_FLAG_INVALID)
					continue;
				if (comps[j] != 0xFFFF)
					continue;
				comps[j] = (uint16_t)c;
				cnt++;
				stack[sp++] = j;
			}
		}
		if (cnt > largestCnt)
		{
			largestCnt = cnt;
			largestId = c;
		}
		c++;
	}
	for (uint32_t s = 0; s < total; s++)
		if (comps[s] != largestId) /* uint16 promoted vs uint32, as in the reference */
			flags[s] = MO_FLAG_INVALID;
	free(stack);
	free(comps);
}

/* grid.cpp:80-97 */
void mo_find_border(int32_t W, int32_t H, uint16_t *flags)
{
	const uint32_t total = (uint32_t)W * (uint32_t)H;
	for (uint32_t i = 0; i < total; i++)
	{
		if (!(flags[i] & MO_FLAG_INVALID))
			continue;
		const int32_t x = (int32_t)(i % (uint32_t)W), y = (int32_t)(i / (uint32_t)W);
		for (int k = 0; k < 4; k++)
		{
			const uint32_t j = idx_or_m(W, H, x + OFFS[k][0], y + OFFS[k][1]);
			if (j == MO_M)
				continue;
			if (!(flags[j] & MO_FLAG_INVALID))
				flags[i] |= MO_FLAG_BORDER;
		}
	}
}

/* grid.cpp:186-196 */
uint32_t mo_add_mana(int32_t W, int32_t H, uint16_t *flags)
{
	const uint32_t total = (uint32_t)W * (uint32_t)H;
	uint32_t valid = 0;
	for (uint32_t i = 0; i < total; i++)
		if (!(flags[i] & MO_FLAG_INVALID))
		{
			valid++;
			flags[i] |= MO_FLAG_MANA;
		}
	return valid;
}

/* grid.cpp:147-198 */
void mo_new_grid(const mo_procedural *p, int32_t W, int32_t H, int32_t ox, int32_t oy, int32_t radius, int16_t *elev, uint16_t *flags, uint32_t *valid_tiles)
{
	const uint32_t total = (uint32_t)W * (uint32_t)H;
	for (uint32_t i = 0; i < total; i++)
	{
		const float px = (float)((int32_t)(i % (uint32_t)W) - ox), py = (float)((int32_t)(i / (uint32_t)W) - oy);
		float e = mo_elevation(p, px, py);
		e = e > -100.0f ? e : -100.0f; /* clamp = min(max(v, a), b) */
		e = e < 100.0f ? e : 100.0f;
		elev[i] = mo_rtos16(e);
	}
	mo_grid_flags_from_elev(W, H, ox, oy, radius, elev, flags);
	mo_fill_unreachable(W, H, flags);
	mo_find_border(W, H, flags);
	const uint32_t v = mo_add_mana(W, H, flags);
	if (valid_tiles)
		*valid_tiles = v;
}

/* sources/map/paths.cpp:173-183, 190-196 */
void mo_pick_waypoints(mo_rng *g, int32_t W, int32_t H, uint16_t *flags, int n, uint32_t *wp_tiles)
{
	const uint32_t total = (uint32_t)W * (uint32_t)H;
	for (int k = 0; k < n; k++)
	{
		uint32_t i;
		for (;;)
		{
			i = mo_rng_range_u32(g, 0u, total);
			if (!(flags[i] & (MO_FLAG_INVALID | MO_FLAG_WAYPOINT)))
				break;
		}
		wp_tiles[k] = i;
		flags[i] |= MO_FLAG_WAYPOINT;
	}
}

/* sources/map/generate.cpp:155-165 without meshes */
int mo_mapgen(uint64_t s1, uint64_t s2, int32_t W, int32_t H, int32_t ox, int32_t oy, int32_t radius, int nwp, int16_t *elev, uint16_t *flags, uint32_t *wp_tiles, uint32_t *tiles, uint32_t *dist, uint32_t *fullpath_buf, uint32_t fullpath_cap, uint32_t *fullpath_len, uint32_t *full_distance, uint32_t *agg3)
{
	mo_rng g = { { s1, s2 } };
	mo_procedural p;
	mo_procedural_init(&p, &g);
	mo_new_grid(&p, W, H, ox, oy, radius, elev, flags, 0);
	mo_pick_waypoints(&g, W, H, flags, nwp, wp_tiles);
	mo_grid grid = { W, H, ox, oy, elev, flags };
	return mo_waypoints_update(&grid, wp_tiles, nwp, 1, tiles, dist, fullpath_buf, fullpath_cap, fullpath_len, full_distance, agg3);
}

typedef struct batch_job
{
	uint64_t s1, s2;
	int first, count, stride;
	uint64_t checksum;
} batch_job;

static void *batch_job_entry(void *p)
{
	batch_job *j = (batch_job *)p;
	enum { W = 101, T = 101 * 101, N = 4, CAP = 4096 };
	int16_t *elev = (int16_t *)malloc(T * 2);
	uint16_t *flags = (uint16_t *)malloc(T * 2);
	uint32_t *tiles = (uint32_t *)malloc((size_t)N * T * 4);
	uint32_t *dist = (uint32_t *)malloc((size_t)N * T * 4);
	uint32_t *fp = (uint32_t *)malloc((size_t)N * CAP * 4);
	uint64_t cs = 0;
	for (int k = j->first; k < j->count; k += j->stride)
	{
		uint32_t wp[N], fl[N], fd[N], agg[3];
		mo_mapgen(j->s1 + (uint64_t)k, j->s2, W, W, 50, 50, 45, N, elev, flags, wp, tiles, dist, fp, CAP, fl, fd, agg);
		cs += (uint64_t)agg[0] * 31 + agg[1] * 7 + agg[2] + fd[0];
	}
	j->checksum = cs;
	free(elev);
	free(flags);
	free(tiles);
	free(dist);
	free(fp);
	return 0;
}

uint64_t mo_mapgen_batch_checksum(uint64_t s1, uint64_t s2, int nmaps, int threads)
{
	if (threads < 1)
		threads = 1;
	if (threads > 256)
		threads = 256;
	batch_job jobs[256];
	pthread_t thr[256];
	for (int t = 0; t < threads; t++)
	{
		jobs[t].s1 = s1;
		jobs[t].s2 = s2;
		jobs[t].first = t;
		jobs[t].count = nmaps;
		jobs[t].stride = threads;
		jobs[t].checksum = 0;
		pthread_create(&thr[t], 0, batch_job_entry, &jobs[t]);
	}
	uint64_t cs = 0;
	for (int t = 0; t < threads; t++)
	{
		pthread_join(thr[t], 0);
		cs += jobs[t].checksum;
	}
	return cs;
}

/* =============================================================================================
 * monster steering -- sources/game/monsters.cpp:60-98, monsterSpawning.cpp:205-216, game.cpp:15-21
 * ============================================================================================= */
void mo_monsters_tick(const mo_waypoints *w, uint32_t gameTime, mo_monsters *ms, int64_t *damage_sum)
{
	for (int i = 0; i < ms->n; i++)
	{
		if (!ms->alive[i])
			continue;
		if (gameTime < ms->timeEnd[i]) /* monsters.cpp:64 */
			continue;
		for (int k = 0; k < w->n; k++) /* :67-69 */
			if (w->wp_tiles[k] == ms->tile[i])
				ms->bits[i] |= 1u << k;
		if (__builtin_popcount(ms->bits[i]) == w->n) /* :70-77 */
		{
			*damage_sum += ms->damage ? ms->damage[i] : 0;
			ms->alive[i] = 0;
			continue;
		}
		uint32_t go[2];
		mo_waypoints_find(w, ms->tile[i], ms->bits[i], go); /* :79 */
		ms->tileStart[i] = ms->tile[i];
		ms->tileEnd[i] = ms->tile[i] = go[0];
		ms->timeStart[i] = gameTime;
		if (ms->timeToArrive[i] > gameTime) /* :85-89, uint64 arithmetic through distNext */
		{
			const uint64_t distNext = mo_neighbor_distance(w->grid, ms->tileStart[i], ms->tileEnd[i]);
			ms->timeEnd[i] = (uint32_t)(gameTime + (ms->timeToArrive[i] - gameTime) * distNext / go[1]);
		}
		{
			const uint32_t lo = gameTime + 1; /* :90 */
			ms->timeEnd[i] = ms->timeEnd[i] > lo ? ms->timeEnd[i] : lo;
		}
		const int fire = ms->dotflags && (ms->dotflags[i] & 1), water = ms->dotflags && (ms->dotflags[i] & 2);
		if (water && !fire) /* :92-93 */
			ms->timeEnd[i] = gameTime + (ms->timeEnd[i] - gameTime) * 2;
		if (fire && !water) /* :94-95 */
			ms->timeEnd[i] = gameTime + (ms->timeEnd[i] - gameTime) / 2;
		{
			const uint32_t lo = gameTime + 1; /* :96 */
			ms->timeEnd[i] = ms->timeEnd[i] > lo ? ms->timeEnd[i] : lo;
		}
	}
}

/* monsterSpawning.cpp:205-216 */
void mo_monster_spawn(const mo_waypoints *w, uint32_t minDistanceSpawner, uint32_t gameTime, float speed, uint32_t *out3)
{
	const uint32_t position = w->wp_tiles[minDistanceSpawner];
	const uint32_t bits = 1u << minDistanceSpawner;
	uint32_t go[2];
	mo_waypoints_find(w, position, bits, go);
	out3[0] = position;
	out3[1] = bits;
	out3[2] = gameTime + (uint32_t)(30.0f * mo_stor((int32_t)go[1]) / speed);
}

/* game.cpp:15-21 */
void mo_monsters_position(const mo_grid *g, uint32_t gameTime, int n, const uint32_t *tileStart, const uint32_t *tileEnd, const uint32_t *timeStart, const uint32_t *timeEnd, float *out3n)
{
	for (int i = 0; i < n; i++)
	{
		float ca[3], cb[3];
		mo_grid_center(g, tileStart[i], ca);
		mo_grid_center(g, tileEnd[i], cb);
		const float num = (float)((int64_t)gameTime - (int64_t)timeStart[i]);
		const float den = (float)((int64_t)timeEnd[i] - (int64_t)timeStart[i]);
		float fac = num / den;
		fac = fac > 0.0f ? fac : 0.0f; /* saturate = min(max(v, 0), 1) with a>b?a:b / a<b?a:b ; NaN -> 0 */
		fac = fac < 1.0f ? fac : 1.0f;
		for (int k = 0; k < 3; k++)
			out3n[i * 3 + k] = ca[k] + (cb[k] - ca[k]) * fac;
	}
}

/* sources/player/buildingsPlacement.cpp:18-42 */
int mo_placing_blocks(const mo_grid *g, const uint32_t *wp_tiles, int n, uint32_t candidate, const uint32_t *monster_tiles, const uint8_t *monster_alive, int nmonsters, uint32_t *scratch_tiles, uint32_t *scratch_dist)
{
	const size_t T = (size_t)g->W * (size_t)g->H;
	uint16_t *fl = (uint16_t *)malloc(T * 2);
	memcpy(fl, g->flags, T * 2);
	fl[candidate] |= MO_FLAG_BUILDING;
	mo_grid trial = *g;
	trial.flags = fl;
	mo_directions_update(&trial, &wp_tiles[0], 1, scratch_tiles, scratch_dist);
	free(fl);
	for (int i = 0; i < n; i++)
		if (scratch_dist[wp_tiles[i]] == MO_M)
			return 1;
	for (int i = 0; i < nmonsters; i++)
		if ((!monster_alive || monster_alive[i]) && scratch_dist[monster_tiles[i]] == MO_M)
			return 1;
	return 0;
}

/* Maker::sdf over the marching-cubes lattice: sources/map/generate.cpp:85-96 (sdf = procedural->sdf(pos) + cliff term of
 * invalid land tiles), sources/map/procedural.cpp:134 (elevation(x, z) - y), lattice positions
 * externals/cage/sources/libcore/mesh/marchingCubes.cpp:247-257 ((b - a) * ((xyz - 2) / (res - 5)) + a), storage order
 * marchingCubes.cpp:149-165 (x fastest, then y, then z). Evaluated point by point like the reference. */
void mo_sdf_lattice(const mo_procedural *p, const mo_grid *g, const float *a, const float *b, const int32_t *res, float *out)
{
	const float dx = (float)res[0] - 5.0f, dy = (float)res[1] - 5.0f, dz = (float)res[2] - 5.0f;
	size_t o = 0;
	for (int32_t z = 0; z < res[2]; z++)
		for (int32_t y = 0; y < res[1]; y++)
			for (int32_t x = 0; x < res[0]; x++)
			{
				const float fx = ((float)x - 2.0f) / dx, fy = ((float)y - 2.0f) / dy, fz = ((float)z - 2.0f) / dz;
				const float px = (b[0] - a[0]) * fx + a[0];
				const float py = (b[1] - a[1]) * fy + a[1];
				const float pz = (b[2] - a[2]) * fz + a[2];
				float d = mo_elevation(p, px, pz) - py;
				const uint32_t i = mo_grid_index_f(g, px, pz); /* Grid::index(Vec3) = index(Vec2(x, z)), grid.cpp:113-116 */
				if (i != MO_M)
				{
					const uint16_t f = g->flags[i];
					if ((f & MO_FLAG_INVALID) && !(f & MO_FLAG_WATER))
					{
						float t = (py - -7.5f) / (-6.5f - -7.5f); /* find(value, lower, upper), generate.h:37-40 */
						t = t > 0.0f ? t : 0.0f;                  /* saturate: min(max(t, 0), 1), math.h:1021-1032 */
						t = t < 1.0f ? t : 1.0f;
						d += t;
					}
				}
				out[o++] = d;
			}
}

/* =============================================================================================
 * damage-over-time tick ("damage monsters") -- sources/game/monsters.cpp:101-183, integer part only
 * (the RNG-driven visual effects of :140-149 and the ghost entity of :176 are not state).
 * dot arrays are [5][n] in DamageTypeEnum order (game.h:18-27): Physical, Fire, Water, Poison, Magic.
 * ============================================================================================= */
static void mo_eliminate_opposing(uint32_t *dur, uint32_t *dmg, int n, int i, int a, int b) /* monsters.cpp:101-114 */
{
	if (dur[(size_t)a * n + i] > 0 && dur[(size_t)b * n + i] > 0)
	{
		dmg[(size_t)a * n + i] = 0;
		dmg[(size_t)b * n + i] = 0;
		dur[(size_t)a * n + i] = 30;
		dur[(size_t)b * n + i] = 30;
	}
}

static void mo_apply_dot(uint32_t *dur, uint32_t *dmg, int n, int i, int type, uint8_t resist, int64_t *life) /* monsters.cpp:116-150 */
{
	const int strengthen = type == 0 ? 3 : 4; /* Physical is strengthened by Poison, everything else by Magic */
	const int super = dmg[(size_t)strengthen * n + i] > 0; /* affected(): dots[].damage > 0, monsters.cpp:13-16 */
	uint32_t *du = &dur[(size_t)type * n + i], *da = &dmg[(size_t)type * n + i];
	uint32_t d = *da;
	if (*du > 1)
		d /= *du;
	if (*du > 0)
		(*du)--;
	if (d == 0)
		return;
	*da -= d;
	if (super)
		d *= 2; /* uint32, wraps like the reference */
	if (resist & (1u << type))
		d /= 4;
	*life -= (int64_t)d;
}

void mo_monsters_dots_tick(int n, uint32_t *dot_duration, uint32_t *dot_damage, const uint8_t *resist, int64_t *life, const uint32_t *money, uint8_t *alive, int64_t *money_sum, uint32_t *killed)
{
	for (int i = 0; i < n; i++)
	{
		if (!alive[i])
			continue;
		mo_eliminate_opposing(dot_duration, dot_damage, n, i, 1, 2); /* Fire, Water */
		mo_eliminate_opposing(dot_duration, dot_damage, n, i, 3, 4); /* Poison, Magic */
		for (int t = 0; t < 5; t++)
			mo_apply_dot(dot_duration, dot_damage, n, i, t, resist ? resist[i] : 0, &life[i]);
		if (life[i] <= 0) /* monsters.cpp:172-178 */
		{
			*money_sum += money ? money[i] : 0;
			alive[i] = 0;
			(*killed)++;
		}
	}
}

/* ---- "attacks": sources/game/attacks.cpp:132-160 per tower, queries by brute force (see mazetd_oracle.h) ---- */
static float mo_flat_dist2(const float *a, const float *b)
{
	/* distanceSquared(a * Vec3(1, 0, 1), b * Vec3(1, 0, 1)): dot sums ((0 + dx*dx) + 0*0) + dz*dz (cage-core/math.h:919-938) */
	const float d0 = a[0] - b[0], d2 = a[2] - b[2];
	float sum = 0;
	sum += d0 * d0;
	sum += 0.0f;
	sum += d2 * d2;
	return sum;
}

void mo_towers_attack(const mo_grid *g, uint32_t gameTime, mo_towers *tw, int nm, const uint32_t *tileStart, const uint32_t *tileEnd, const uint32_t *timeStart, const uint32_t *timeEnd, const uint8_t *alive,
					  const uint8_t *monsterClass, const uint32_t *timeToArrive, const int64_t *life, uint32_t *dot_duration, uint32_t *dot_damage)
{
	float *pos = (float *)malloc((size_t)(nm > 0 ? nm : 1) * 3 * sizeof(float));
	mo_monsters_position(g, gameTime, nm, tileStart, tileEnd, timeStart, timeEnd, pos); /* spatial.cpp:30 and attacks.cpp:80: mv.position() */
	for (int t = 0; t < tw->n; t++)
	{
		tw->target_out[t] = -1;
		tw->hits_out[t] = 0;
		if (tw->firingDelay[t] > 0) /* :136-140 */
		{
			tw->firingDelay[t]--;
			continue;
		}
		/* applyMods, :40-71 */
		uint32_t damage = tw->damage[t], overTime = tw->overTime[t], firingPeriod = tw->firingPeriod[t], manaCost = tw->manaCost[t];
		float firingRange = tw->firingRange[t], splashRadius = tw->splashRadius[t];
		const int element = tw->element[t];
		if (element != 0)
		{
			overTime = (uint32_t)(sqrtf((float)damage) * 20.0f);
			damage *= 3u;
		}
		switch (tw->bonus[t])
		{
		case 1: damage *= 2u; manaCost *= 2u; break;
		case 2: firingPeriod /= 2u; break;
		case 3: firingRange += 4.0f; break;
		case 4: splashRadius += 2.0f; manaCost *= 3u; break;
		case 5: overTime /= 5u; break;
		case 6: manaCost /= 3u; break;
		default: break;
		}
		float mp[3];
		mo_grid_center(g, tw->tile[t], mp);
		/* findMonstersForTower + pickTargetMonster: first strictly smallest value in ascending monster index */
		int best = -1, count = 0;
		float bv = 0;
		const int targeting = tw->targeting[t];
		const uint64_t rw = tw->random_words ? tw->random_words[t] : 0;
		for (int i = 0; i < nm; i++)
		{
			if (!alive[i] || (monsterClass && (monsterClass[i] & tw->invalidClasses[t])))
				continue;
			if (!(sqrtf(mo_flat_dist2(&pos[i * 3], mp)) <= firingRange))
				continue;
			count++;
			float v;
			switch (targeting)
			{
			case 1: v = (float)timeToArrive[i]; break;            /* Front */
			case 2: v = (float)(-(int32_t)timeToArrive[i]); break; /* Back: -numeric_cast<sint32>(timeToArrive) */
			case 3: v = mo_flat_dist2(&pos[i * 3], mp); break;     /* Closest */
			case 4: v = -mo_flat_dist2(&pos[i * 3], mp); break;    /* Farthest */
			case 5: v = (float)(-life[i]); break;                  /* Strongest */
			case 6: v = (float)life[i]; break;                     /* Weakest */
			default: v = 0; break;                                 /* Random: resolved below */
			}
			if (best < 0 || (targeting != 0 && v < bv))
			{
				best = i;
				bv = v;
			}
		}
		if (count == 0)
			continue; /* :152-153 */
		if (manaCost > 0) /* consumeMana, :95-105 */
		{
			if (tw->mana[t] < manaCost)
				continue;
			tw->mana[t] -= manaCost;
		}
		if (targeting == 0)
		{
			uint64_t want = rw % (uint64_t)count, seen = 0;
			for (int i = 0; i < nm; i++)
			{
				if (!alive[i] || (monsterClass && (monsterClass[i] & tw->invalidClasses[t])))
					continue;
				if (!(sqrtf(mo_flat_dist2(&pos[i * 3], mp)) <= firingRange))
					continue;
				if (seen++ == want)
				{
					best = i;
					break;
				}
			}
		}
		tw->target_out[t] = best;
		/* damageMonsters on the target, or on the splash set around it (:157-159, :120-130) */
		if (splashRadius > 0)
		{
			for (int i = 0; i < nm; i++)
			{
				if (!alive[i] || (monsterClass && (monsterClass[i] & tw->invalidClasses[t])))
					continue;
				if (!(sqrtf(mo_flat_dist2(&pos[i * 3], &pos[best * 3])) <= splashRadius))
					continue;
				dot_damage[(size_t)element * nm + i] += damage;
				dot_duration[(size_t)element * nm + i] += overTime;
				tw->hits_out[t]++;
			}
		}
		else
		{
			dot_damage[(size_t)element * nm + best] += damage;
			dot_duration[(size_t)element * nm + best] += overTime;
			tw->hits_out[t] = 1;
		}
		tw->firingDelay[t] += firingPeriod; /* :161 */
	}
	free(pos);
}
