// tilesim.c -- developer tool (not product, not oracle): counts what a tile schedule + an in-tile engine cost on a
// C4/C5-like synthetic grid, so that kernel designs can be compared offline before GPU time is spent.
//   gcc -O2 -o tilesim tilesim.c -lm && ./tilesim W H nsrc window mode
// modes: 0 = rounds with a priority window (the r01 schedule), 1 = strict priority over tiles (sequential lower bound)
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define INF 0x3FFFFFFFu
#define TS 32
static int W, H, TX, TY;
static uint8_t *blocked;
static int16_t *elev;
static uint16_t *wr, *wd; // weight to the right / down neighbour
static uint32_t *exact, *d;

static uint64_t s0 = 1346519564496ull, s1 = 42245614964156ull;
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static uint64_t rnd(void)
{
	uint64_t r = rotl(s0 * 5, 7) * 9;
	s1 ^= s0;
	s0 = rotl(s0, 24) ^ s1 ^ (s1 << 16);
	s1 = rotl(s1, 37);
	return r;
}
static uint32_t weight(int a, int b)
{
	const float n = 1.0f / 255;
	float dy = (float)elev[a] * n - (float)elev[b] * n;
	float s = sqrtf((0 + 1.0f) + dy * dy);
	return (uint32_t)(int)(s * 255.0f);
}

// ---- heap dijkstra (exact) ----
typedef struct { uint32_t k, v; } HE;
static HE *heap; static size_t hn;
static void hpush(uint32_t k, uint32_t v)
{
	size_t i = hn++;
	while (i) { size_t p = (i - 1) / 2; if (heap[p].k <= k) break; heap[i] = heap[p]; i = p; }
	heap[i].k = k; heap[i].v = v;
}
static HE hpop(void)
{
	HE top = heap[0], last = heap[--hn];
	size_t i = 0;
	for (;;) { size_t c = 2 * i + 1; if (c >= hn) break; if (c + 1 < hn && heap[c + 1].k < heap[c].k) c++; if (heap[c].k >= last.k) break; heap[i] = heap[c]; i = c; }
	heap[i] = last;
	return top;
}

static uint32_t *srcs; static int nsrc;

// ---- tile visit ----
#define RW (TS + 2)
static uint32_t F[RW][RW], FCHK[RW][RW];     // frame labels
static uint8_t FB[RW][RW];     // blocked (apron cells outside grid also blocked)
static uint16_t FR[RW][RW], FD[RW][RW]; // weight to right / down of frame cell
static uint64_t tot_sweeps, tot_steps1024, tot_steps255, tot_visits, tot_changed_sweeps, tot_cells_popped;
static uint64_t hist_sweeps[64];
static int sweep_order[4] = { 0, 3, 1, 2 }; // quadrant ids: 0=DR 1=UL 2=DL 3=UR

// one Gauss-Seidel quadrant sweep over the interior (apron fixed); returns changed
static int quad_sweep(int q)
{
	int changed = 0;
	int dy = (q == 0 || q == 2) ? 1 : -1; // D for DR, DL
	int dx = (q == 0 || q == 3) ? 1 : -1; // R for DR, UR
	for (int iy = 0; iy < TS; iy++)
	{
		int y = dy > 0 ? 1 + iy : TS - iy;
		for (int ix = 0; ix < TS; ix++)
		{
			int x = dx > 0 ? 1 + ix : TS - ix;
			if (FB[y][x]) continue;
			uint32_t v = F[y][x];
			// vertical predecessor
			uint32_t pv = F[y - dy][x];
			uint32_t wv = dy > 0 ? FD[y - 1][x] : FD[y][x];
			if (pv < INF && !(FB[y - dy][x] && 0) && pv + wv < v) v = pv + wv;
			uint32_t ph = F[y][x - dx];
			uint32_t wh = dx > 0 ? FR[y][x - 1] : FR[y][x];
			if (ph < INF && ph + wh < v) v = ph + wh;
			if (v < F[y][x]) { F[y][x] = v; changed = 1; }
		}
	}
	return changed;
}

// bucketed wavefront cost model: number of steps with bucket width delta (label-correcting within bucket counted as one step per wave)
static uint32_t LQ[RW * RW * 8];
static int wavefront_steps(uint32_t Fin[RW][RW], uint32_t delta, uint64_t *pops)
{
	// simulate: pending set; each step: take all pending with label < dmin+delta, relax their neighbours; improved ones pending
	static uint32_t G[RW][RW]; static uint8_t pend[RW][RW];
	memcpy(G, Fin, sizeof(G)); memset(pend, 0, sizeof(pend));
	// seeds: apron relax + interior sources(already finite & better than neighbours) -- treat every finite interior cell that can improve a neighbour, plus apron
	int steps = 0;
	// initial: apron cells relax inward
	for (int y = 0; y < RW; y++) for (int x = 0; x < RW; x++)
	{
		int ap = (y == 0 || y == RW - 1 || x == 0 || x == RW - 1);
		if (!ap) continue;
		if ((y == 0 || y == RW - 1) && (x == 0 || x == RW - 1)) continue;
		if (G[y][x] >= INF) continue;
		int ny = y == 0 ? 1 : (y == RW - 1 ? RW - 2 : y), nx = x == 0 ? 1 : (x == RW - 1 ? RW - 2 : x);
		if (FB[ny][nx]) continue;
		uint32_t w = y == 0 ? FD[0][x] : y == RW - 1 ? FD[RW - 2][x] : x == 0 ? FR[y][0] : FR[y][RW - 2];
		if (G[y][x] + w < G[ny][nx]) { G[ny][nx] = G[y][x] + w; pend[ny][nx] = 1; }
	}
	// interior cells that are finite and could improve a neighbour (e.g. sources or left over from an earlier visit)
	for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++) if (!FB[y][x] && G[y][x] < INF && !pend[y][x])
	{
		int imp = 0;
		if (y > 1 && !FB[y - 1][x] && G[y][x] + FD[y - 1][x] < G[y - 1][x]) imp = 1;
		if (y < TS && !FB[y + 1][x] && G[y][x] + FD[y][x] < G[y + 1][x]) imp = 1;
		if (x > 1 && !FB[y][x - 1] && G[y][x] + FR[y][x - 1] < G[y][x - 1]) imp = 1;
		if (x < TS && !FB[y][x + 1] && G[y][x] + FR[y][x] < G[y][x + 1]) imp = 1;
		if (imp) pend[y][x] = 1;
	}
	for (;;)
	{
		uint32_t dmin = INF;
		for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++) if (pend[y][x] && G[y][x] < dmin) dmin = G[y][x];
		if (dmin == INF) break;
		steps++;
		int n = 0;
		for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++) if (pend[y][x] && G[y][x] < dmin + delta) { LQ[n++] = y * RW + x; pend[y][x] = 0; }
		*pops += n;
		for (int i = 0; i < n; i++)
		{
			int y = LQ[i] / RW, x = LQ[i] % RW; uint32_t v = G[y][x];
			if (y > 1 && !FB[y - 1][x] && v + FD[y - 1][x] < G[y - 1][x]) { G[y - 1][x] = v + FD[y - 1][x]; pend[y - 1][x] = 1; }
			if (y < TS && !FB[y + 1][x] && v + FD[y][x] < G[y + 1][x]) { G[y + 1][x] = v + FD[y][x]; pend[y + 1][x] = 1; }
			if (x > 1 && !FB[y][x - 1] && v + FR[y][x - 1] < G[y][x - 1]) { G[y][x - 1] = v + FR[y][x - 1]; pend[y][x - 1] = 1; }
			if (x < TS && !FB[y][x + 1] && v + FR[y][x] < G[y][x + 1]) { G[y][x + 1] = v + FR[y][x]; pend[y][x + 1] = 1; }
		}
	}
	memcpy(Fin, G, sizeof(G));
	return steps;
}


// ---- toroidal 32-step skewed passes (register engine model). type 0=DR 1=UL (layout A), 2=DL 3=UR (layout B) ----
static uint64_t tot_passes, tot_conv, hist_passes[128];
static int tor_pass(uint32_t G[RW][RW], int type)
{
	int changed = 0;
	static uint32_t N[TS];
	for (int st = 0; st < TS; st++)
	{
		int s = (type == 0 || type == 2) ? st : TS - 1 - st;
		for (int c = 0; c < TS; c++)
		{
			int r = (type < 2) ? ((s - c) & 31) : ((s + c) & 31);
			int y = r + 1, x = c + 1;
			uint32_t v = G[y][x];
			N[c] = v;
			if (FB[y][x]) continue;
			int dy = (type == 0 || type == 2) ? -1 : 1; // source row offset
			int dx = (type == 0 || type == 3) ? -1 : 1; // source col offset
			int ry = r + dy, cx = c + dx;
			if (ry >= 0 && ry < TS && !FB[y + dy][x]) { uint32_t w = dy < 0 ? FD[y - 1][x] : FD[y][x]; uint32_t p = G[y + dy][x]; if (p < INF && p + w < v) v = p + w; }
			if (cx >= 0 && cx < TS && !FB[y][x + dx]) { uint32_t w = dx < 0 ? FR[y][x - 1] : FR[y][x]; uint32_t p = G[y][x + dx]; if (p < INF && p + w < v) v = p + w; }
			N[c] = v;
		}
		for (int c = 0; c < TS; c++)
		{
			int r = (type < 2) ? ((s - c) & 31) : ((s + c) & 31);
			if (N[c] < G[r + 1][c + 1]) { G[r + 1][c + 1] = N[c]; changed = 1; }
		}
	}
	return changed;
}
static int tor_engine(uint32_t G[RW][RW], int *conv)
{
	// seed border cells from the apron once
	for (int i = 1; i <= TS; i++)
	{
		if (!FB[1][i] && G[0][i] < INF && G[0][i] + FD[0][i] < G[1][i]) G[1][i] = G[0][i] + FD[0][i];
		if (!FB[TS][i] && G[RW - 1][i] < INF && G[RW - 1][i] + FD[TS][i] < G[TS][i]) G[TS][i] = G[RW - 1][i] + FD[TS][i];
		if (!FB[i][1] && G[i][0] < INF && G[i][0] + FR[i][0] < G[i][1]) G[i][1] = G[i][0] + FR[i][0];
		if (!FB[i][TS] && G[i][RW - 1] < INF && G[i][RW - 1] + FR[i][TS] < G[i][TS]) G[i][TS] = G[i][RW - 1] + FR[i][TS];
	}
	int clean[4] = { 0, 0, 0, 0 }, n = 0, layout = 0, t = 0; *conv = 0;
	extern int tor_order[4];
	int pos = 0;
	while (!(clean[0] && clean[1] && clean[2] && clean[3]) && n < 500)
	{
		// next dirty type in order
		while (clean[tor_order[pos & 3]]) pos++;
		t = tor_order[pos & 3];
		int lay = t >> 1;
		if (lay != layout) { (*conv)++; layout = lay; }
		int c = tor_pass(G, t); n++;
		if (c) { clean[0] = clean[1] = clean[2] = clean[3] = 0; /* stay on the same type: a toroidal pass is not idempotent */ }
		else { clean[t] = 1; pos++; }
	}
	return n;
}
int tor_order[4] = { 0, 1, 2, 3 };

static uint32_t *key; // per tile offered key
static int first_quadrant_hint = -1;

// visits tile (tx,ty); returns offers to 4 neighbours in off[4] (INF = none): up, down, left, right
static void visit(int tx, int ty, uint32_t off[4], int engine_stats)
{
	int gx0 = tx * TS - 1, gy0 = ty * TS - 1;
	for (int y = 0; y < RW; y++) for (int x = 0; x < RW; x++)
	{
		int gx = gx0 + x, gy = gy0 + y;
		if (gx < 0 || gx >= W || gy < 0 || gy >= H) { F[y][x] = INF; FB[y][x] = 1; FR[y][x] = FD[y][x] = 0; continue; }
		size_t i = (size_t)gy * W + gx;
		F[y][x] = d[i]; FB[y][x] = blocked[i]; FR[y][x] = wr[i]; FD[y][x] = wd[i];
		if (blocked[i]) F[y][x] = INF;
	}
	// sources blocked or not keep label 0 (they're in d already)
	uint32_t ap_old[4][TS];
	for (int i = 0; i < TS; i++) { ap_old[0][i] = F[0][1 + i]; ap_old[1][i] = F[RW - 1][1 + i]; ap_old[2][i] = F[1 + i][0]; ap_old[3][i] = F[1 + i][RW - 1]; }
	tot_visits++;
	if (engine_stats)
	{
		static uint32_t G2[RW][RW];
		memcpy(G2, F, sizeof(F));
		uint64_t pops = 0;
		tot_steps1024 += wavefront_steps(G2, 1024, &pops);
		tot_cells_popped += pops;
		memcpy(G2, F, sizeof(F));
		uint64_t p2 = 0;
		tot_steps255 += wavefront_steps(G2, 255, &p2);
	}
	if (engine_stats)
	{
		static uint32_t G3[RW][RW];
		memcpy(G3, F, sizeof(F));
		int conv; int np = tor_engine(G3, &conv);
		tot_passes += np; tot_conv += conv; hist_passes[np < 127 ? np : 127]++;
		for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++) G3[0][0] = 0;
		memcpy(FCHK, G3, sizeof(G3));
	}
	// sweep engine (this one produces the result)
	int nosw = 0, ns = 0, q = 0, nch = 0;
	while (nosw < 3 && ns < 400)
	{
		int c = quad_sweep(sweep_order[q & 3]); q++; ns++;
		if (c) { nosw = 0; nch++; } else nosw++;
	}
	tot_sweeps += ns; tot_changed_sweeps += nch; hist_sweeps[ns < 63 ? ns : 63]++;
	if (engine_stats) for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++) if (!FB[y][x] && F[y][x] != FCHK[y][x]) { printf("toroidal engine mismatch\n"); exit(1); }
	// write back
	for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++)
	{
		int gx = gx0 + x, gy = gy0 + y;
		if (gx >= W || gy >= H) continue;
		size_t i = (size_t)gy * W + gx;
		if (!blocked[i] && F[y][x] < d[i]) d[i] = F[y][x];
	}
	for (int s = 0; s < 4; s++) off[s] = INF;
	for (int i = 0; i < TS; i++)
	{
		// up: border cell F[1][1+i] offers to apron F[0][1+i]
		if (!FB[0][1 + i] && F[1][1 + i] < INF && F[1][1 + i] + FD[0][1 + i] < ap_old[0][i]) { uint32_t k = F[1][1 + i] + FD[0][1 + i]; if (k < off[0]) off[0] = k; }
		if (!FB[RW - 1][1 + i] && F[TS][1 + i] < INF && F[TS][1 + i] + FD[TS][1 + i] < ap_old[1][i]) { uint32_t k = F[TS][1 + i] + FD[TS][1 + i]; if (k < off[1]) off[1] = k; }
		if (!FB[1 + i][0] && F[1 + i][1] < INF && F[1 + i][1] + FR[1 + i][0] < ap_old[2][i]) { uint32_t k = F[1 + i][1] + FR[1 + i][0]; if (k < off[2]) off[2] = k; }
		if (!FB[1 + i][RW - 1] && F[1 + i][TS] < INF && F[1 + i][TS] + FR[1 + i][TS] < ap_old[3][i]) { uint32_t k = F[1 + i][TS] + FR[1 + i][TS]; if (k < off[3]) off[3] = k; }
	}
}

int main(int argc, char **argv)
{
	W = argc > 1 ? atoi(argv[1]) : 1024; H = argc > 2 ? atoi(argv[2]) : 1024; nsrc = argc > 3 ? atoi(argv[3]) : 4;
	uint32_t window = argc > 4 ? (uint32_t)atoi(argv[4]) : 98304u;
	int mode = argc > 5 ? atoi(argv[5]) : 0;
	int wall = argc > 6 ? atoi(argv[6]) : 30;
	int stats = argc > 7 ? atoi(argv[7]) : 1;
	size_t T = (size_t)W * H;
	TX = (W + TS - 1) / TS; TY = (H + TS - 1) / TS;
	blocked = malloc(T); elev = malloc(2 * T); wr = calloc(T, 2); wd = calloc(T, 2); exact = malloc(4 * T); d = malloc(4 * T);
	for (size_t i = 0; i < T; i++) blocked[i] = (rnd() % 100) < (uint64_t)wall;
	for (size_t i = 0; i < T; i++) elev[i] = (int16_t)((int)(rnd() % 512) - 256);
	for (int y = 0; y < H; y++) for (int x = 0; x < W; x++)
	{
		size_t i = (size_t)y * W + x;
		if (x + 1 < W) wr[i] = (uint16_t)weight(i, i + 1);
		if (y + 1 < H) wd[i] = (uint16_t)weight(i, i + W);
	}
	srcs = malloc(4 * nsrc);
	for (int s = 0; s < nsrc;) { uint32_t i = (uint32_t)(rnd() % T); if (!blocked[i]) srcs[s++] = i; }
	// exact
	heap = malloc(sizeof(HE) * T * 4); hn = 0;
	for (size_t i = 0; i < T; i++) exact[i] = INF;
	for (int s = 0; s < nsrc; s++) { exact[srcs[s]] = 0; hpush(0, srcs[s]); }
	while (hn)
	{
		HE e = hpop();
		if (e.k != exact[e.v]) continue;
		int x = e.v % W, y = e.v / W;
		if (x > 0 && !blocked[e.v - 1] && e.k + wr[e.v - 1] < exact[e.v - 1]) { exact[e.v - 1] = e.k + wr[e.v - 1]; hpush(exact[e.v - 1], e.v - 1); }
		if (x + 1 < W && !blocked[e.v + 1] && e.k + wr[e.v] < exact[e.v + 1]) { exact[e.v + 1] = e.k + wr[e.v]; hpush(exact[e.v + 1], e.v + 1); }
		if (y > 0 && !blocked[e.v - W] && e.k + wd[e.v - W] < exact[e.v - W]) { exact[e.v - W] = e.k + wd[e.v - W]; hpush(exact[e.v - W], e.v - W); }
		if (y + 1 < H && !blocked[e.v + W] && e.k + wd[e.v] < exact[e.v + W]) { exact[e.v + W] = e.k + wd[e.v]; hpush(exact[e.v + W], e.v + W); }
	}
	uint32_t dmax = 0; size_t reach = 0;
	for (size_t i = 0; i < T; i++) if (exact[i] < INF) { reach++; if (exact[i] > dmax) dmax = exact[i]; }
	printf("grid %dx%d src %d reach %zu dmax %u (%.0f min-hops)\n", W, H, nsrc, reach, dmax, dmax / 255.0);

	// tile schedule
	for (size_t i = 0; i < T; i++) d[i] = INF;
	for (int s = 0; s < nsrc; s++) d[srcs[s]] = 0;
	int NT = TX * TY;
	key = malloc(4 * NT);
	for (int t = 0; t < NT; t++) key[t] = INF;
	for (int s = 0; s < nsrc; s++) { int x = srcs[s] % W, y = srcs[s] / W; key[(y / TS) * TX + x / TS] = 0; }
	int rounds = 0; uint64_t maxper = 0, sum_active = 0;
	if (mode == 0)
	{
		uint32_t *cur = malloc(4 * NT);
		for (;;)
		{
			uint32_t kmin = INF; for (int t = 0; t < NT; t++) if (key[t] < kmin) kmin = key[t];
			if (kmin == INF) break;
			int n = 0; for (int t = 0; t < NT; t++) if (key[t] < kmin + window) cur[n++] = t;
			rounds++; sum_active += n; if ((uint64_t)n > maxper) maxper = n;
			// all visits of a round see the state at round start for offers?? (GPU: concurrent; here sequential, slightly optimistic)
			for (int i = 0; i < n; i++) key[cur[i]] = INF;
			for (int i = 0; i < n; i++)
			{
				int t = cur[i], tx = t % TX, ty = t / TX; uint32_t off[4];
				visit(tx, ty, off, stats);
				if (off[0] < INF && ty > 0 && off[0] < key[t - TX]) key[t - TX] = off[0];
				if (off[1] < INF && ty + 1 < TY && off[1] < key[t + TX]) key[t + TX] = off[1];
				if (off[2] < INF && tx > 0 && off[2] < key[t - 1]) key[t - 1] = off[2];
				if (off[3] < INF && tx + 1 < TX && off[3] < key[t + 1]) key[t + 1] = off[3];
			}
		}
	}
	else
	{
		for (;;)
		{
			uint32_t kmin = INF; int bt = -1; for (int t = 0; t < NT; t++) if (key[t] < kmin) { kmin = key[t]; bt = t; }
			if (bt < 0) break;
			rounds++;
			int t = bt, tx = t % TX, ty = t / TX; uint32_t off[4];
			key[t] = INF;
			visit(tx, ty, off, stats);
			if (off[0] < INF && ty > 0 && off[0] < key[t - TX]) key[t - TX] = off[0];
			if (off[1] < INF && ty + 1 < TY && off[1] < key[t + TX]) key[t + TX] = off[1];
			if (off[2] < INF && tx > 0 && off[2] < key[t - 1]) key[t - 1] = off[2];
			if (off[3] < INF && tx + 1 < TX && off[3] < key[t + 1]) key[t + 1] = off[3];
		}
	}
	size_t bad = 0; for (size_t i = 0; i < T; i++) if (!blocked[i] && d[i] != exact[i]) bad++;
	printf("mode %d window %u: rounds %d visits %llu (%.2f per tile) avg active %.0f max %llu mismatches %zu\n", mode, window, rounds, (unsigned long long)tot_visits, (double)tot_visits / NT, rounds ? (double)sum_active / rounds : 0, (unsigned long long)maxper, bad);
	printf("sweep engine: %.2f sweeps/visit (%.2f changing); wavefront d1024: %.1f steps/visit, %.0f pops/visit; d255: %.1f steps/visit\n", (double)tot_sweeps / tot_visits, (double)tot_changed_sweeps / tot_visits, (double)tot_steps1024 / tot_visits, (double)tot_cells_popped / tot_visits, (double)tot_steps255 / tot_visits);
	printf("toroidal engine: %.2f passes/visit, %.2f layout conversions/visit\n passes hist:", (double)tot_passes / tot_visits, (double)tot_conv / tot_visits);
	for (int i = 0; i < 128; i++) if (hist_passes[i]) printf(" %d:%llu", i, (unsigned long long)hist_passes[i]); printf("\n");
	printf("sweeps hist:"); for (int i = 0; i < 64; i++) if (hist_sweeps[i]) printf(" %d:%llu", i, (unsigned long long)hist_sweeps[i]); printf("\n");
	return 0;
}
