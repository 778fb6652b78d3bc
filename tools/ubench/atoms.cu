// micro-benchmark: latency of dependent shared-memory operations on sm_100a (one warp, or W warps contending)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template<int mode>
__global__ void k(int iters, long long *out)
{
	__shared__ uint32_t s[4096];
	for (int i = threadIdx.x; i < 4096; i += blockDim.x)
		s[i] = 0xFFFFFFFFu - i;
	__syncthreads();
	uint32_t idx = (threadIdx.x * 33u) & 4095u, v = threadIdx.x;
	long long t0 = clock64();
	for (int i = 0; i < iters; i++)
	{
		if (mode == 0) // dependent LDS chain
			idx = (s[idx] + idx) & 4095u;
		else if (mode == 1) // dependent ATOMS.MIN chain (result feeds the next address)
			idx = (atomicMin(&s[idx], v + i) + idx + 1u) & 4095u;
		else if (mode == 2) // 4 independent ATOMS then one dependent use
		{
			uint32_t a = atomicMin(&s[idx], v + i), b = atomicMin(&s[(idx + 1) & 4095u], v + i), c = atomicMin(&s[(idx + 64) & 4095u], v + i), d = atomicMin(&s[(idx + 101) & 4095u], v + i);
			idx = (a + b + c + d + idx + 1u) & 4095u;
		}
		else if (mode == 3) // ballot chain
		{
			uint32_t b = __ballot_sync(0xFFFFFFFFu, (idx & 1u) != 0);
			idx = (idx + __popc(b) + 1u) & 4095u;
		}
		else if (mode == 4) // redux chain
			idx = (__reduce_min_sync(0xFFFFFFFFu, idx) + threadIdx.x + 1u) & 4095u;
		else if (mode == 5) // named barrier, 64 threads
		{
			asm volatile("bar.sync 1, 64;" ::: "memory");
			idx++;
		}
		else if (mode == 6) // STS then dependent LDS of another lane's slot
		{
			s[threadIdx.x] = idx;
			__syncwarp();
			idx = (s[(threadIdx.x + 1) & 31] + 1u) & 4095u;
			__syncwarp();
		}
		else if (mode == 7 || mode == 8) // the relaxation step's chain: queue -> label + 4 weights -> 4 atomics (7) / 4 plain load+store (8) -> 5 ballots -> queue
		{
			const uint16_t *q16 = (const uint16_t *)s;
			const uint32_t c = (q16[(idx * 2u) & 8191u] + idx) & 2047u;
			const uint32_t dd = s[1024 + c];
			const uint32_t w0 = q16[4096 + ((c + 7) & 2047u)], w1 = q16[4096 + ((c + 1) & 2047u)], w2 = q16[4096 + c], w3 = q16[4096 + ((c + 9) & 2047u)];
			uint32_t o0, o1, o2, o3;
			if (mode == 7)
			{
				o0 = atomicMin(&s[1024 + ((c + 101) & 2047u)], dd + w0);
				o1 = atomicMin(&s[1024 + ((c + 2047) & 2047u)], dd + w1);
				o2 = atomicMin(&s[1024 + ((c + 1) & 2047u)], dd + w2);
				o3 = atomicMin(&s[1024 + ((c + 1946) & 2047u)], dd + w3);
			}
			else
			{
				o0 = s[1024 + ((c + 101) & 2047u)];
				o1 = s[1024 + ((c + 2047) & 2047u)];
				o2 = s[1024 + ((c + 1) & 2047u)];
				o3 = s[1024 + ((c + 1946) & 2047u)];
				if (dd + w0 < o0) s[1024 + ((c + 101) & 2047u)] = dd + w0;
				if (dd + w1 < o1) s[1024 + ((c + 2047) & 2047u)] = dd + w1;
				if (dd + w2 < o2) s[1024 + ((c + 1) & 2047u)] = dd + w2;
				if (dd + w3 < o3) s[1024 + ((c + 1946) & 2047u)] = dd + w3;
			}
			const uint32_t b0 = __ballot_sync(0xFFFFFFFFu, o0 & 1u), b1 = __ballot_sync(0xFFFFFFFFu, o1 & 1u), b2 = __ballot_sync(0xFFFFFFFFu, o2 & 1u), b3 = __ballot_sync(0xFFFFFFFFu, o3 & 1u), b4 = __ballot_sync(0xFFFFFFFFu, dd & 1u);
			const uint32_t lt = (1u << (threadIdx.x & 31u)) - 1u;
			uint32_t pos = __popc(b0 & lt);
			((uint16_t *)s)[(pos + idx) & 1023u] = (uint16_t)c;
			pos += __popc(b0) + __popc(b1 & lt);
			((uint16_t *)s)[(pos + idx) & 1023u] = (uint16_t)(c + 1);
			pos += __popc(b1) + __popc(b2 & lt);
			((uint16_t *)s)[(pos + idx) & 1023u] = (uint16_t)(c + 2);
			pos += __popc(b2) + __popc(b3 & lt) + __popc(b4);
			idx = (idx + pos + 1u) & 4095u;
		}
	}
	long long t1 = clock64();
	if (threadIdx.x == 0)
		out[blockIdx.x] = t1 - t0;
	if (idx == 0x12345678u)
		out[1] = s[5];
}
int main()
{
	long long *d, h;
	cudaMalloc(&d, 64);
	const char *names[] = { "LDS chain", "ATOMS.MIN chain (32 lanes, spread)", "4x ATOMS + use", "ballot+popc chain", "redux.min chain", "bar.sync 64 thr", "STS+syncwarp+LDS", "relax-step chain, atomics", "relax-step chain, plain ld/st" };
	for (int threads : { 32, 64, 256, 512 })
		for (int mode = 0; mode < 9; mode++)
		{
			if (mode == 5 && threads != 64)
				continue;
			const int iters = 2000;
			switch (mode)
			{
			case 0: k<0><<<1, threads>>>(iters, d); break;
			case 1: k<1><<<1, threads>>>(iters, d); break;
			case 2: k<2><<<1, threads>>>(iters, d); break;
			case 3: k<3><<<1, threads>>>(iters, d); break;
			case 4: k<4><<<1, threads>>>(iters, d); break;
			case 5: k<5><<<1, threads>>>(iters, d); break;
			case 6: k<6><<<1, threads>>>(iters, d); break;
			case 7: k<7><<<1, threads>>>(iters, d); break;
			default: k<8><<<1, threads>>>(iters, d); break;
			}
			cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
			printf("threads=%3d %-38s %.1f cycles/iter\n", threads, names[mode], (double)h / iters);
		}
	return 0;
}
