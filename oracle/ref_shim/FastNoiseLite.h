/*
 * FastNoiseLite.h -- TEST INFRASTRUCTURE (oracle/_ref build only).
 *
 * Stand-in for the third-party header the reference includes at
 * externals/cage/sources/libcore/noiseFunction.cpp:3 and which is ABSENT from /root/reference
 * (Auburn/FastNoiseLite @ 14ff80df8c817b4bb1976d84f3b25ed972ef8be0, see oracle/fnl_subset.h).
 * It exposes the class surface noiseFunction.cpp:14-127 touches and forwards the one
 * configuration the hot path uses (OpenSimplex2 2-D, fractal None/Ridged) to the restatement in
 * oracle/fnl_subset.h. Every other configuration aborts: it is not on the path and must not be
 * silently approximated.
 */
#ifndef MTD_ORACLE_FASTNOISELITE_SHIM_H
#define MTD_ORACLE_FASTNOISELITE_SHIM_H

#include <cstdio>
#include <cstdlib>

#include "../fnl_subset.h"

class FastNoiseLite
{
public:
	enum NoiseType
	{
		NoiseType_OpenSimplex2,
		NoiseType_OpenSimplex2S,
		NoiseType_Cellular,
		NoiseType_Perlin,
		NoiseType_ValueCubic,
		NoiseType_Value
	};
	enum FractalType
	{
		FractalType_None,
		FractalType_FBm,
		FractalType_Ridged,
		FractalType_PingPong
	};
	enum CellularDistanceFunction
	{
		CellularDistanceFunction_Euclidean,
		CellularDistanceFunction_EuclideanSq,
		CellularDistanceFunction_Manhattan,
		CellularDistanceFunction_Hybrid
	};
	enum CellularReturnType
	{
		CellularReturnType_CellValue,
		CellularReturnType_Distance,
		CellularReturnType_Distance2,
		CellularReturnType_Distance2Add,
		CellularReturnType_Distance2Sub,
		CellularReturnType_Distance2Mul,
		CellularReturnType_Distance2Div
	};

	explicit FastNoiseLite(int seed = 1337) { fnls_init(&st, seed); }

	void SetFrequency(float f) { st.frequency = f; }
	void SetNoiseType(NoiseType t) { st.noise_type = t == NoiseType_OpenSimplex2 ? FNLS_NOISE_OPENSIMPLEX2 : FNLS_NOISE_OTHER; }
	void SetFractalType(FractalType t) { st.fractal_type = (int)t; }
	void SetFractalOctaves(int o)
	{
		st.octaves = o;
		fnls_calc_fractal_bounding(&st);
	}
	void SetFractalLacunarity(float l) { st.lacunarity = l; }
	void SetFractalGain(float g)
	{
		st.gain = g;
		fnls_calc_fractal_bounding(&st);
	}
	void SetCellularDistanceFunction(CellularDistanceFunction) {}
	void SetCellularReturnType(CellularReturnType) {}

	float GetNoise(float x, float y) const
	{
		if (st.noise_type != FNLS_NOISE_OPENSIMPLEX2 || (st.fractal_type != FNLS_FRACTAL_NONE && st.fractal_type != FNLS_FRACTAL_RIDGED))
			off_path("GetNoise(x,y) with a noise/fractal type outside the restated subset");
		return fnls_get_noise_2d(&st, x, y);
	}

	float GetNoise(float, float, float) const
	{
		off_path("GetNoise(x,y,z): 3-D noise feeds texturing only (sources/map/procedural.cpp:79-132), out of scope");
		return 0;
	}

private:
	fnls_state st;

	[[noreturn]] static void off_path(const char *what)
	{
		std::fprintf(stderr, "FastNoiseLite shim: %s\n", what);
		std::abort();
	}
};

#endif
