/*
 * fnl_subset.h -- TEST INFRASTRUCTURE (oracle). Not part of the shipped product path.
 *
 * CPU restatement of the subset of FastNoiseLite that mazetd's elevation function reaches.
 *
 * Third-party dependency, ABSENT from /root/reference:
 *   Auburn/FastNoiseLite (https://github.com/Auburns/FastNoise.git),
 *   pinned by the reference at commit 14ff80df8c817b4bb1976d84f3b25ed972ef8be0
 *   (externals/cage/externals/fastnoise/CMakeLists.txt:7-8, .SUBMODULES.json "missing").
 * Reference call sites that fix WHICH subset is needed:
 *   externals/cage/sources/libcore/noiseFunction.cpp:97-107  (ctor(seed), SetFrequency, SetNoiseType,
 *       SetFractalType, SetFractalOctaves, SetFractalLacunarity, SetFractalGain, SetCellular*)
 *   externals/cage/sources/libcore/noiseFunction.cpp:117-121 (GetNoise(x, y) with float arguments)
 *   sources/map/procedural.cpp:15-26,68 (OpenSimplex2, Ridged, 4 octaves, gain .55, lacunarity 2.1, freq .003)
 *
 * PARITY UNPINNED: no test in the reference pins a single noise value
 * (externals/cage/sources/test-core/noiseFunction.cpp:18-50 only writes images), and the upstream header is
 * not available offline. This file restates the *published* algorithm (FastNoiseLite.h "GetNoise(x,y)" ->
 * TransformNoiseCoordinate -> GenFractalRidged -> SingleSimplex -> GradCoord) from its documentation and
 * well-known structure. The GPU kernel is held bit-exact to THIS restatement; equivalence with upstream bits
 * cannot be verified here.
 *
 * All arithmetic is IEEE float32 with no FMA contraction (build with -ffp-contract=off), matching the
 * reference build flags (externals/cage/cmake/cage_build_configuration.cmake:1-93: no -march, no fast-math).
 */
#ifndef MTD_ORACLE_FNL_SUBSET_H
#define MTD_ORACLE_FNL_SUBSET_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { FNLS_NOISE_OPENSIMPLEX2 = 0, FNLS_NOISE_OTHER = 1 };
enum { FNLS_FRACTAL_NONE = 0, FNLS_FRACTAL_FBM = 1, FNLS_FRACTAL_RIDGED = 2, FNLS_FRACTAL_PINGPONG = 3 };

typedef struct fnls_state
{
	int32_t seed;
	float frequency;
	int32_t noise_type;
	int32_t fractal_type;
	int32_t octaves;
	float lacunarity;
	float gain;
	float weighted_strength;
	float fractal_bounding;
} fnls_state;

#define FNLS_PRIME_X 501125321
#define FNLS_PRIME_Y 1136930381

/* 24 unit gradients at 7.5deg + k*15deg, walking clockwise from +y, repeated 5x (=120), then 8 at 22.5+45j. */
static inline void fnls_gradient(int idx /*0..127*/, float *gx, float *gy)
{
	static const float mag[6] = { 0.130526192220052f, 0.38268343236509f, 0.608761429008721f, 0.793353340291235f, 0.923879532511287f, 0.99144486137381f };
	int k;
	if (idx < 120)
		k = idx % 24;
	else
		k = (idx - 120) * 3 + 1;
	/* quadrant q: 0: (+s,+c) 1: (+c,-s) 2: (-s,-c) 3: (-c,+s) with position p in 0..5 */
	const int q = k / 6, p = k % 6;
	const float s = mag[p], c = mag[5 - p];
	switch (q)
	{
		case 0: *gx = s; *gy = c; break;
		case 1: *gx = c; *gy = -s; break;
		case 2: *gx = -s; *gy = -c; break;
		default: *gx = -c; *gy = s; break;
	}
}

static inline void fnls_calc_fractal_bounding(fnls_state *st)
{
	float gain = st->gain < 0 ? -st->gain : st->gain;
	float amp = gain;
	float amp_fractal = 1.0f;
	for (int i = 1; i < st->octaves; i++)
	{
		amp_fractal += amp;
		amp *= gain;
	}
	st->fractal_bounding = 1 / amp_fractal;
}

static inline void fnls_init(fnls_state *st, int32_t seed)
{
	st->seed = seed;
	st->frequency = 0.01f;
	st->noise_type = FNLS_NOISE_OPENSIMPLEX2;
	st->fractal_type = FNLS_FRACTAL_NONE;
	st->octaves = 3;
	st->lacunarity = 2.0f;
	st->gain = 0.5f;
	st->weighted_strength = 0.0f;
	st->fractal_bounding = 1 / 1.75f;
}

static inline int32_t fnls_fast_floor(float f)
{
	return f >= 0 ? (int32_t)f : (int32_t)f - 1;
}

static inline int32_t fnls_hash(int32_t seed, int32_t xp, int32_t yp)
{
	uint32_t h = (uint32_t)(seed ^ xp ^ yp);
	h *= 0x27d4eb2du;
	return (int32_t)h;
}

static inline float fnls_grad_coord(int32_t seed, int32_t xp, int32_t yp, float xd, float yd)
{
	int32_t hash = fnls_hash(seed, xp, yp);
	hash ^= hash >> 15; /* arithmetic shift */
	hash &= 127 << 1;
	float xg, yg;
	fnls_gradient(hash >> 1, &xg, &yg);
	return xd * xg + yd * yg;
}

static inline float fnls_single_simplex(int32_t seed, float x, float y)
{
	const float SQRT3 = 1.7320508075688772935274463415059f;
	const float G2 = (3 - SQRT3) / 6;

	int32_t i = fnls_fast_floor(x);
	int32_t j = fnls_fast_floor(y);
	const float xi = (float)(x - i);
	const float yi = (float)(y - j);

	const float t = (xi + yi) * G2;
	const float x0 = (float)(xi - t);
	const float y0 = (float)(yi - t);

	i = (int32_t)((uint32_t)i * (uint32_t)FNLS_PRIME_X);
	j = (int32_t)((uint32_t)j * (uint32_t)FNLS_PRIME_Y);

	float n0, n1, n2;

	const float a = 0.5f - x0 * x0 - y0 * y0;
	if (a <= 0)
		n0 = 0;
	else
		n0 = (a * a) * (a * a) * fnls_grad_coord(seed, i, j, x0, y0);

	const float c = (float)(2 * (1 - 2 * G2) * (1 / G2 - 2)) * t + ((float)(-2 * (1 - 2 * G2) * (1 - 2 * G2)) + a);
	if (c <= 0)
		n2 = 0;
	else
	{
		const float x2 = x0 + (2 * (float)G2 - 1);
		const float y2 = y0 + (2 * (float)G2 - 1);
		n2 = (c * c) * (c * c) * fnls_grad_coord(seed, (int32_t)((uint32_t)i + (uint32_t)FNLS_PRIME_X), (int32_t)((uint32_t)j + (uint32_t)FNLS_PRIME_Y), x2, y2);
	}

	if (y0 > x0)
	{
		const float x1 = x0 + (float)G2;
		const float y1 = y0 + ((float)G2 - 1);
		const float b = 0.5f - x1 * x1 - y1 * y1;
		if (b <= 0)
			n1 = 0;
		else
			n1 = (b * b) * (b * b) * fnls_grad_coord(seed, i, (int32_t)((uint32_t)j + (uint32_t)FNLS_PRIME_Y), x1, y1);
	}
	else
	{
		const float x1 = x0 + ((float)G2 - 1);
		const float y1 = y0 + (float)G2;
		const float b = 0.5f - x1 * x1 - y1 * y1;
		if (b <= 0)
			n1 = 0;
		else
			n1 = (b * b) * (b * b) * fnls_grad_coord(seed, (int32_t)((uint32_t)i + (uint32_t)FNLS_PRIME_X), j, x1, y1);
	}

	return (n0 + n1 + n2) * 99.83685446303647f;
}

static inline float fnls_lerp(float a, float b, float t)
{
	return a + t * (b - a);
}

static inline float fnls_gen_fractal_ridged(const fnls_state *st, float x, float y)
{
	int32_t seed = st->seed;
	float sum = 0;
	float amp = st->fractal_bounding;
	for (int i = 0; i < st->octaves; i++)
	{
		float noise = fnls_single_simplex(seed++, x, y);
		noise = noise < 0 ? -noise : noise;
		sum += (noise * -2 + 1) * amp;
		amp *= fnls_lerp(1.0f, 1 - noise, st->weighted_strength);
		x *= st->lacunarity;
		y *= st->lacunarity;
		amp *= st->gain;
	}
	return sum;
}

/* GetNoise(x, y) for noise_type OpenSimplex2 and fractal None / Ridged (the only combination the path uses). */
static inline float fnls_get_noise_2d(const fnls_state *st, float x, float y)
{
	x *= st->frequency;
	y *= st->frequency;
	{
		const float SQRT3 = (float)1.7320508075688772935274463415059;
		const float F2 = 0.5f * (SQRT3 - 1);
		const float t = (x + y) * F2;
		x += t;
		y += t;
	}
	if (st->fractal_type == FNLS_FRACTAL_RIDGED)
		return fnls_gen_fractal_ridged(st, x, y);
	return fnls_single_simplex(st->seed, x, y);
}

#ifdef __cplusplus
}
#endif

#endif
