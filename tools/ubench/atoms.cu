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
	const char *names[] = { "LDS chain", "ATOMS.MIN chain (32 lanes, spread)", "4x ATOMS + use", "ballot+popc chain", "redux.min chain", "bar.sync 64 thr", "STS+syncwarp+LDS" };
	for (int threads : { 32, 64, 256, 512 })
		for (int mode = 0; mode < 7; mode++)
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
			default: k<6><<<1, threads>>>(iters, d); break;
			}
			cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
			printf("threads=%3d %-38s %.1f cycles/iter\n", threads, names[mode], (double)h / iters);
		}
	return 0;
}
