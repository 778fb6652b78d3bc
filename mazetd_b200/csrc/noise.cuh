// noise.cuh -- ProceduralImpl::elevation on the device (sources/map/procedural.cpp:65-70).
//
// The noise itself is FastNoiseLite (OpenSimplex2 2-D, Ridged fractal, 4 octaves, gain .55, lacunarity 2.1,
// frequency .003 -- sources/map/procedural.cpp:17-25 via externals/cage/sources/libcore/noiseFunction.cpp:97-121),
// a third-party header absent from the reference snapshot (Auburn/FastNoiseLite @ 14ff80df8c81). This is a
// from-scratch device formulation of its published algorithm; it is held bit-exact to the CPU restatement
// used by the oracle. Parity against upstream bits is UNPINNED (no reference test pins a noise value).
//
// All float32 arithmetic is unfused, round-to-nearest (explicit intrinsics + -fmad=false).
#pragma once

#include "common.cuh"

namespace mtd
{
	namespace noise
	{
		constexpr int PRIME_X = 501125321;
		constexpr int PRIME_Y = 1136930381;

		// 128 unit gradients: 24 directions at 7.5deg + 15deg*k (clockwise from +y) five times, then 8 at 22.5deg + 45deg*j
		__device__ __forceinline__ float2 gradient(int idx)
		{
			const float mag[6] = { 0.130526192220052f, 0.38268343236509f, 0.608761429008721f, 0.793353340291235f, 0.923879532511287f, 0.99144486137381f };
			const int k = idx < 120 ? idx % 24 : (idx - 120) * 3 + 1;
			const int q = k / 6, p = k % 6;
			const float s = mag[p], c = mag[5 - p];
			switch (q)
			{
				case 0: return make_float2(s, c);
				case 1: return make_float2(c, -s);
				case 2: return make_float2(-s, -c);
				default: return make_float2(-c, s);
			}
		}

		// fills a 128-entry table (shared memory) cooperatively
		__device__ __forceinline__ void fill_gradients(float2 *table)
		{
			for (int i = threadIdx.x; i < 128; i += blockDim.x)
				table[i] = gradient(i);
		}

		__device__ __forceinline__ int fast_floor(float f)
		{
			return f >= 0 ? __float2int_rz(f) : __float2int_rz(f) - 1;
		}

		__device__ __forceinline__ float grad_coord(const float2 *table, int seed, int xp, int yp, float xd, float yd)
		{
			int hash = seed ^ xp ^ yp;
			hash = (int)((uint32_t)hash * 0x27d4eb2du);
			hash ^= hash >> 15;
			hash &= 127 << 1;
			const float2 g = table[hash >> 1];
			return __fadd_rn(__fmul_rn(xd, g.x), __fmul_rn(yd, g.y));
		}

		__device__ __forceinline__ float pow4_times(float a, float g)
		{
			const float a2 = __fmul_rn(a, a);
			return __fmul_rn(__fmul_rn(a2, a2), g);
		}

		__device__ __forceinline__ float single_simplex(const float2 *table, int seed, float x, float y)
		{
			const float SQRT3 = 1.7320508075688772935274463415059f;
			const float G2 = (3 - SQRT3) / 6;
			const float C1 = (float)(2 * (1 - 2 * G2) * (1 / G2 - 2));
			const float C2 = (float)(-2 * (1 - 2 * G2) * (1 - 2 * G2));
			const float G2M1 = (float)G2 - 1;
			const float G22M1 = 2 * (float)G2 - 1;

			int i = fast_floor(x);
			int j = fast_floor(y);
			const float xi = __fsub_rn(x, (float)i);
			const float yi = __fsub_rn(y, (float)j);

			const float t = __fmul_rn(__fadd_rn(xi, yi), G2);
			const float x0 = __fsub_rn(xi, t);
			const float y0 = __fsub_rn(yi, t);

			i = (int)((uint32_t)i * (uint32_t)PRIME_X);
			j = (int)((uint32_t)j * (uint32_t)PRIME_Y);

			// The three corner terms are evaluated without branches: a corner whose falloff is not positive contributes a zero
			// (the reference skips it and adds its initial 0). The zero may come out negative here (0 * negative gradient);
			// that cannot reach a result bit: a zero term changes no non-zero sum, and an all-zero sum is only used through
			// |n| * -2 + 1, which is 1 for either sign.
			const float a = __fsub_rn(__fsub_rn(0.5f, __fmul_rn(x0, x0)), __fmul_rn(y0, y0));
			const float n0 = pow4_times(a > 0 ? a : 0.0f, grad_coord(table, seed, i, j, x0, y0));

			const float c = __fadd_rn(__fmul_rn(C1, t), __fadd_rn(C2, a));
			const float x2 = __fadd_rn(x0, G22M1);
			const float y2 = __fadd_rn(y0, G22M1);
			const float n2 = pow4_times(c > 0 ? c : 0.0f, grad_coord(table, seed, (int)((uint32_t)i + (uint32_t)PRIME_X), (int)((uint32_t)j + (uint32_t)PRIME_Y), x2, y2));

			const bool up = y0 > x0;
			const float x1 = __fadd_rn(x0, up ? G2 : G2M1);
			const float y1 = __fadd_rn(y0, up ? G2M1 : G2);
			const float b = __fsub_rn(__fsub_rn(0.5f, __fmul_rn(x1, x1)), __fmul_rn(y1, y1));
			const float n1 = pow4_times(b > 0 ? b : 0.0f, grad_coord(table, seed, up ? i : (int)((uint32_t)i + (uint32_t)PRIME_X), up ? (int)((uint32_t)j + (uint32_t)PRIME_Y) : j, x1, y1));

			return __fmul_rn(__fadd_rn(__fadd_rn(n0, n1), n2), 99.83685446303647f);
		}

		// CalculateFractalBounding for (octaves, gain)
		__device__ __forceinline__ float fractal_bounding(int octaves, float gain_in)
		{
			const float gain = gain_in < 0 ? -gain_in : gain_in;
			float amp = gain;
			float amp_fractal = 1.0f;
			for (int i = 1; i < octaves; i++)
			{
				amp_fractal = __fadd_rn(amp_fractal, amp);
				amp = __fmul_rn(amp, gain);
			}
			return __fdiv_rn(1.0f, amp_fractal);
		}

		struct ElevationParams
		{
			float off_x, off_y; // elevationOffset, procedural.cpp:13
			int seed;           // elevation noise seed, procedural.cpp:24
			float bounding;     // fractal bounding for 4 octaves, gain .55
		};

		// GetNoise(x, y): frequency scale, OpenSimplex2 skew, ridged fractal
		__device__ __forceinline__ float ridged_noise(const float2 *table, const ElevationParams &p, float x, float y)
		{
			const float frequency = 0.003f, lacunarity = 2.1f, gain = 0.55f;
			x = __fmul_rn(x, frequency);
			y = __fmul_rn(y, frequency);
			{
				const float SQRT3 = (float)1.7320508075688772935274463415059;
				const float F2 = 0.5f * (SQRT3 - 1);
				const float t = __fmul_rn(__fadd_rn(x, y), F2);
				x = __fadd_rn(x, t);
				y = __fadd_rn(y, t);
			}
			int seed = p.seed;
			float sum = 0;
			float amp = p.bounding;
#pragma unroll
			for (int o = 0; o < 4; o++)
			{
				float n = single_simplex(table, seed++, x, y);
				n = n < 0 ? -n : n;
				sum = __fadd_rn(sum, __fmul_rn(__fadd_rn(__fmul_rn(n, -2.0f), 1.0f), amp));
				// amp *= Lerp(1, 1 - n, weightedStrength = 0)  ==  amp * (1 + 0 * ((1 - n) - 1))
				amp = __fmul_rn(amp, __fadd_rn(1.0f, __fmul_rn(0.0f, __fsub_rn(__fsub_rn(1.0f, n), 1.0f))));
				x = __fmul_rn(x, lacunarity);
				y = __fmul_rn(y, lacunarity);
				amp = __fmul_rn(amp, gain);
			}
			return sum;
		}

		// ProceduralImpl::elevation, procedural.cpp:65-70
		__device__ __forceinline__ float elevation(const float2 *table, const ElevationParams &p, float px, float py)
		{
			const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(px, px), __fmul_rn(py, py)));
			float t = __fsub_rn(len, 30.0f);
			t = t > 0.0f ? t : 0.0f;
			t = __fmul_rn(t, 0.2f);
			const float base = -__fmul_rn(t, t);
			const float n = __fmul_rn(ridged_noise(table, p, __fadd_rn(px, p.off_x), __fadd_rn(py, p.off_y)), 10.0f);
			const float v = __fadd_rn(base, n);
			return v > -8.5f ? v : -8.5f;
		}
	}
}
