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

// ---- column-owner engine: lane c owns column c of the tile; per step every lane pops ONE pending cell of its column,
// relaxes up / down itself and hands the left / right candidates to the neighbour lanes (no atomics, no shared queue).
// variant 0: lowest pending row; 1: elevator (continue in the current direction); 2: pending cell with the smallest label;
// 3: as 2 but only cells below (warp-min pending label + delta) are popped
static uint64_t col_steps[8], col_pops[8];
static void col_engine(uint32_t Fin[RW][RW], int variant, uint32_t delta)
{
	static uint32_t G[RW][RW];
	uint32_t pend[TS + 2];
	int dir[TS + 2], last[TS + 2];
	memcpy(G, Fin, sizeof(G));
	memset(pend, 0, sizeof(pend));
	for (int c = 0; c < TS + 2; c++) { dir[c] = 1; last[c] = 0; }
	// seeds as in wavefront_steps
	for (int i = 1; i <= TS; i++)
	{
		if (!FB[1][i] && G[0][i] < INF && G[0][i] + FD[0][i] < G[1][i]) { G[1][i] = G[0][i] + FD[0][i]; pend[i] |= 1u << 0; }
		if (!FB[TS][i] && G[RW - 1][i] < INF && G[RW - 1][i] + FD[TS][i] < G[TS][i]) { G[TS][i] = G[RW - 1][i] + FD[TS][i]; pend[i] |= 1u << (TS - 1); }
		if (!FB[i][1] && G[i][0] < INF && G[i][0] + FR[i][0] < G[i][1]) { G[i][1] = G[i][0] + FR[i][0]; pend[1] |= 1u << (i - 1); }
		if (!FB[i][TS] && G[i][RW - 1] < INF && G[i][RW - 1] + FR[i][TS] < G[i][TS]) { G[i][TS] = G[i][RW - 1] + FR[i][TS]; pend[TS] |= 1u << (i - 1); }
	}
	for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++) if (!FB[y][x] && G[y][x] < INF)
	{
		int imp = 0;
		if (y > 1 && !FB[y - 1][x] && G[y][x] + FD[y - 1][x] < G[y - 1][x]) imp = 1;
		if (y < TS && !FB[y + 1][x] && G[y][x] + FD[y][x] < G[y + 1][x]) imp = 1;
		if (x > 1 && !FB[y][x - 1] && G[y][x] + FR[y][x - 1] < G[y][x - 1]) imp = 1;
		if (x < TS && !FB[y][x + 1] && G[y][x] + FR[y][x] < G[y][x + 1]) imp = 1;
		if (imp) pend[x] |= 1u << (y - 1);
	}
	for (;;)
	{
		int any = 0;
		for (int c = 1; c <= TS; c++) any |= pend[c] != 0;
		if (!any) break;
		col_steps[variant]++;
		uint32_t thr = 0xFFFFFFFFu;
		if (variant == 3)
		{
			uint32_t m = INF;
			for (int c = 1; c <= TS; c++) for (int r = 0; r < TS; r++) if ((pend[c] >> r) & 1u) if (G[r + 1][c] < m) m = G[r + 1][c];
			thr = m + delta;
		}
		int pr[TS + 2]; uint32_t pd[TS + 2];
		for (int c = 1; c <= TS; c++)
		{
			pr[c] = -1;
			if (!pend[c]) continue;
			int r = -1;
			if (variant == 0) r = __builtin_ctz(pend[c]);
			else if (variant == 1 || variant >= 4)
			{
				// next pending row at or beyond `last` in direction dir, else turn round
				for (int pass = 0; pass < 2 && r < 0; pass++)
				{
					if (dir[c] > 0) { uint32_t m = pend[c] & ~((last[c] >= 32 ? 0xFFFFFFFFu : (1u << last[c])) - 1u); if (last[c] < 32 && m) r = __builtin_ctz(m); else dir[c] = -1; }
					else { uint32_t m = pend[c] & ((last[c] >= 31 ? 0xFFFFFFFFu : ((2u << last[c]) - 1u))); if (m) r = 31 - __builtin_clz(m); else dir[c] = 1; }
				}
				if (r < 0) r = __builtin_ctz(pend[c]);
			}
			else
			{
				uint32_t best = 0xFFFFFFFFu;
				for (int q = 0; q < TS; q++) if ((pend[c] >> q) & 1u) if (G[q + 1][c] < best) { best = G[q + 1][c]; r = q; }
				if (variant == 3 && best >= thr) r = -1;
			}
			if (r < 0) continue;
			pr[c] = r; pd[c] = G[r + 1][c];
			if (variant < 4) { last[c] = r; pend[c] &= ~(1u << r); col_pops[variant]++; }
		}
		if (variant >= 4)
		{
			uint32_t m = 0xFFFFFFFFu;
			for (int c = 1; c <= TS; c++) if (pr[c] >= 0 && pd[c] < m) m = pd[c];
			for (int c = 1; c <= TS; c++)
			{
				if (pr[c] < 0) continue;
				if (pd[c] - m >= delta) { pr[c] = -1; /* candidate stays pending; look at another one next step */ last[c] = (last[c] + dir[c]) & 31; continue; }
				last[c] = pr[c]; pend[c] &= ~(1u << pr[c]); col_pops[variant]++;
			}
		}
		// relax: vertical by the owner, lateral handed to the neighbours (applied after all lanes popped: same step)
		for (int c = 1; c <= TS; c++)
		{
			if (pr[c] < 0) continue;
			int y = pr[c] + 1; uint32_t v = pd[c];
			if (y > 1 && !FB[y - 1][c] && v + FD[y - 1][c] < G[y - 1][c]) { G[y - 1][c] = v + FD[y - 1][c]; pend[c] |= 1u << (y - 2); }
			if (y < TS && !FB[y + 1][c] && v + FD[y][c] < G[y + 1][c]) { G[y + 1][c] = v + FD[y][c]; pend[c] |= 1u << y; }
			if (c > 1 && !FB[y][c - 1] && v + FR[y][c - 1] < G[y][c - 1]) { G[y][c - 1] = v + FR[y][c - 1]; pend[c - 1] |= 1u << (y - 1); }
			if (c < TS && !FB[y][c + 1] && v + FR[y][c] < G[y][c + 1]) { G[y][c + 1] = v + FR[y][c]; pend[c + 1] |= 1u << (y - 1); }
		}
	}
	for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++) if (!FB[y][x] && G[y][x] != FCHK[y][x]) { printf("column engine mismatch v%d\n", variant); exit(1); }
}

static uint32_t *key; // per tile offered key
static int first_quadrant_hint = -1;

// visits tile (tx,ty); returns offers to 4 neighbours in off[4] (INF = none): up, down, left, right
static uint32_t RES[TS][TS]; static int defer_wb = 0; static int last_steps = 0;
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
		static uint32_t G4[RW][RW];
		for (int v = 0; v < 4; v++) { memcpy(G4, F, sizeof(F)); col_engine(G4, v, 1024); }
		memcpy(G4, F, sizeof(F)); col_engine(G4, 4, 1024); memcpy(G4, F, sizeof(F)); col_engine(G4, 5, 2048); memcpy(G4, F, sizeof(F)); col_engine(G4, 6, 4096); memcpy(G4, F, sizeof(F)); col_engine(G4, 7, 8192);
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
	if (defer_wb) { uint64_t pp = 0; static uint32_t G5[RW][RW]; memcpy(G5, F, sizeof(F)); (void)pp; for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++) RES[y - 1][x - 1] = F[y][x]; last_steps = ns; }
	// write back
	if (!defer_wb) for (int y = 1; y <= TS; y++) for (int x = 1; x <= TS; x++)
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

	else if (mode == 2)
	{
		// event-driven: P workers, a visit stages at its start and writes back / offers at its end (duration in ticks grows
		// with the work it did); policy bit 0: window; bit 1: a tile does not start while a neighbour tile is in flight or
		// pending with a smaller key (margin)
		int P = argc > 8 ? atoi(argv[8]) : 37;
		int policy = argc > 9 ? atoi(argv[9]) : 1;
		uint32_t margin = argc > 10 ? (uint32_t)atoi(argv[10]) : 0;
		typedef struct { int tile, end; uint32_t key; uint32_t off[4]; uint32_t res[TS][TS]; } Job;
		Job *jobs = calloc(P, sizeof(Job));
		for (int w = 0; w < P; w++) jobs[w].tile = -1;
		uint8_t *busy = calloc(NT, 1);
		int now = 0; uint64_t busy_ticks = 0; int inflight = 0;
		defer_wb = 1;
		for (;;)
		{
			// finish
			for (int w = 0; w < P; w++) if (jobs[w].tile >= 0 && jobs[w].end <= now)
			{
				int t = jobs[w].tile, tx = t % TX, ty = t / TX;
				for (int y = 0; y < TS; y++) for (int x = 0; x < TS; x++)
				{
					int gx = tx * TS + x, gy = ty * TS + y; if (gx >= W || gy >= H) continue;
					size_t i = (size_t)gy * W + gx; if (!blocked[i] && jobs[w].res[y][x] < d[i]) d[i] = jobs[w].res[y][x];
				}
				uint32_t *off = jobs[w].off;
				if (off[0] < INF && ty > 0 && off[0] < key[t - TX]) key[t - TX] = off[0];
				if (off[1] < INF && ty + 1 < TY && off[1] < key[t + TX]) key[t + TX] = off[1];
				if (off[2] < INF && tx > 0 && off[2] < key[t - 1]) key[t - 1] = off[2];
				if (off[3] < INF && tx + 1 < TX && off[3] < key[t + 1]) key[t + 1] = off[3];
				busy[t] = 0; jobs[w].tile = -1; inflight--;
			}
			// watermark
			uint32_t wm = INF; int npend = 0;
			for (int t = 0; t < NT; t++) if (key[t] < INF) { npend++; if (key[t] < wm) wm = key[t]; }
			for (int w = 0; w < P; w++) if (jobs[w].tile >= 0 && jobs[w].key < wm) wm = jobs[w].key;
			if (npend == 0 && inflight == 0) break;
			// assign
			for (int w = 0; w < P; w++) if (jobs[w].tile < 0)
			{
				uint32_t bk = INF; int bt = -1;
				for (int t = 0; t < NT; t++)
				{
					if (key[t] >= INF || busy[t] || key[t] >= bk) continue;
					if ((policy & 1) && key[t] - wm >= window) continue;
					if (policy & 2)
					{
						int tx = t % TX, ty = t / TX, ok = 1;
						int nb[4] = { ty > 0 ? t - TX : -1, ty + 1 < TY ? t + TX : -1, tx > 0 ? t - 1 : -1, tx + 1 < TX ? t + 1 : -1 };
						for (int q = 0; q < 4 && ok; q++) if (nb[q] >= 0)
						{
							if (busy[nb[q]]) ok = 0;
							else if (key[nb[q]] < INF && key[nb[q]] + margin < key[t]) ok = 0;
						}
						if (!ok) continue;
					}
					bk = key[t]; bt = t;
				}
				if (bt < 0) break;
				jobs[w].tile = bt; jobs[w].key = key[bt]; key[bt] = INF; busy[bt] = 1; inflight++;
				visit(bt % TX, bt / TX, jobs[w].off, 0);
				memcpy(jobs[w].res, RES, sizeof(RES));
				jobs[w].end = now + 6 + last_steps; // ticks: fixed part + engine
				rounds++;
			}
			busy_ticks += inflight;
			now++;
		}
		printf("async P=%d policy=%d margin=%u: makespan %d ticks, avg busy workers %.1f\n", P, policy, margin, now, (double)busy_ticks / (now ? now : 1));
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
	for (int v = 0; v < 8; v++) printf("column engine v%d: %.1f steps/visit, %.0f pops/visit\n", v, (double)col_steps[v] / tot_visits, (double)col_pops[v] / tot_visits);
	printf("sweeps hist:"); for (int i = 0; i < 64; i++) if (hist_sweeps[i]) printf(" %d:%llu", i, (unsigned long long)hist_sweeps[i]); printf("\n");
	return 0;
}
