// common.cuh -- shared device helpers for the mazetd flow-field kernels (sm_100a).
//
// Bit-exactness contract: every float32 operation that feeds an integer result is written with an explicit
// round-to-nearest intrinsic (and the library is additionally compiled with -fmad=false), so no FMA
// contraction can change a rounding. This mirrors the reference build, which has no -march / fast-math / FMA
// (externals/cage/cmake/cage_build_configuration.cmake:1-93).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#define MTD_INF 0xFFFFFFFFu
// distances at or above this are reported as overflow: the warp scans add up to 2^22 of slack to keys
#define MTD_GUARD 0xFF000000u

namespace mtd
{
	// TileFlags, sources/grid.h:15-28
	enum : uint16_t
	{
		F_INVALID = 1u << 0,
		F_BORDER = 1u << 1,
		F_WAYPOINT = 1u << 2,
		F_WATER = 1u << 5,
		F_SUN = 1u << 6,
		F_WIND = 1u << 7,
		F_SNOW = 1u << 8,
		F_MANA = 1u << 13,
		F_BUILDING = 1u << 14,
		F_TRAP = 1u << 15,
		F_BLOCKING = F_INVALID | F_BUILDING, // sources/map/paths.cpp:65
	};

	// ---- sources/common.h:15-29 ----
	__device__ __forceinline__ float stor(int32_t v)
	{
		const float n = (float)(1.0 / 255); // constexpr Real n = 1.0 / 255
		return __fmul_rn((float)v, n);
	}

	__device__ __forceinline__ int32_t rtos32(float v)
	{
		return __float2int_rz(__fmul_rn(v, 255.0f));
	}

	__device__ __forceinline__ int16_t rtos16(float v)
	{
		return (int16_t)__float2int_rz(__fmul_rn(v, 255.0f));
	}

	// Grid::neighborDistance for 4-neighbours, sources/map/grid.cpp:141-145:
	// rtos32(sqrt(((0 + dx*dx) + dy*dy) + dz*dz)) with dy = stor(ea) - stor(eb) and {dx,dz} = {+-1, 0}.
	// fl(1 + dy^2) == fl(dy^2 + 1), so one expression serves horizontal and vertical edges; always >= 255.
	__device__ __forceinline__ uint32_t edge_weight(int16_t ea, int16_t eb)
	{
		const float dy = __fsub_rn(stor(ea), stor(eb));
		const float s = __fadd_rn(1.0f, __fmul_rn(dy, dy));
		return (uint32_t)rtos32(__fsqrt_rn(s));
	}

	// ---- xoroshiro128**, externals/cage/sources/libcore/math/random.cpp:27-66 ----
	struct Rng
	{
		uint64_t s0, s1;

		__device__ __forceinline__ uint64_t next()
		{
			const uint64_t a = s0;
			uint64_t b = s1;
			const uint64_t r5 = a * 5;
			const uint64_t result = ((r5 << 7) | (r5 >> 57)) * 9;
			b ^= a;
			s0 = ((a << 24) | (a >> 40)) ^ b ^ (b << 16);
			s1 = (b << 37) | (b >> 27);
			return result;
		}

		__device__ __forceinline__ float chance() // randomChance(): Real(double)
		{
			const uint64_t r = next();
			return __double2float_rn(__dmul_rn((double)(r >> 11), 0x1.0p-53));
		}

		__device__ __forceinline__ uint32_t range_u32(uint32_t lo, uint32_t hi)
		{
			if (lo == hi)
				return lo;
			const uint64_t range = (uint64_t)hi - (uint64_t)lo;
			return (uint32_t)(next() % range + lo);
		}
	};

	__device__ __forceinline__ uint32_t lane_id()
	{
		return threadIdx.x & 31u;
	}

	__device__ __forceinline__ uint32_t lanemask_lt()
	{
		uint32_t m;
		asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
		return m;
	}

	__device__ __forceinline__ uint32_t warp_min_u32(uint32_t v)
	{
		return __reduce_min_sync(0xFFFFFFFFu, v);
	}

	// streaming (evict-first) stores for write-once outputs
	__device__ __forceinline__ void st_stream_u32(uint32_t *p, uint32_t v)
	{
		__stcs(p, v);
	}
}
