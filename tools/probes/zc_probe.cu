// Zero-copy probe: how fast can SMs read ray planes straight from pinned host memory, and write 16-B hit records back?
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o zc_probe zc_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void readKernel(const float4* __restrict__ O, const float4* __restrict__ D, int n, int raysPerWarpFetch, unsigned* counter, float4* sink)
{
	const int lane = threadIdx.x & 31;
	float4 acc = make_float4(0, 0, 0, 0);
	while (true)
	{
		unsigned base = 0;
		if (lane == 0) base = atomicAdd(counter, raysPerWarpFetch);
		base = __shfl_sync(0xffffffffu, base, 0);
		if (base >= (unsigned)n) break;
		if (lane < raysPerWarpFetch && base + lane < (unsigned)n)
		{
			float4 o = __ldcs(O + base + lane), d = __ldcs(D + base + lane);
			acc.x += o.x + d.x, acc.y += o.y + d.y, acc.z += o.z + d.z, acc.w += o.w + d.w;
		}
	}
	if (acc.x == 12345.678f) sink[0] = acc;
}

// each lane writes records in a pseudo-random order within a window of `window` records (models rays retiring out of order)
__global__ void writeKernel(float4* __restrict__ H, int n, int window, int streaming)
{
	const int tid = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
	for (int i = tid; i < n; i += stride)
	{
		int w0 = (i / window) * window, k = i - w0;
		int len = n - w0 < window ? n - w0 : window;
		int j = w0 + (int)(((unsigned long long)k * 2654435761ull) % (unsigned)len); // not a bijection in general; good enough for traffic
		float4 v = make_float4((float)i, 1, 2, 3);
		if (streaming) __stcs(H + j, v); else H[j] = v;
	}
}

int main()
{
	const int n = 2 << 20;
	float4 *hO, *hD, *hH, *dSink; unsigned* dCounter;
	CK(cudaHostAlloc(&hO, (size_t)n * 16, cudaHostAllocDefault)); CK(cudaHostAlloc(&hD, (size_t)n * 16, cudaHostAllocDefault));
	CK(cudaHostAlloc(&hH, (size_t)n * 16, cudaHostAllocDefault));
	for (int i = 0; i < n; i++) hO[i] = make_float4(i, 1, 2, 3), hD[i] = make_float4(3, 2, 1, i);
	CK(cudaMalloc(&dSink, 64)); CK(cudaMalloc(&dCounter, 4));
	cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
	float4* dBuf; CK(cudaMalloc(&dBuf, (size_t)n * 16));
	// DMA reference
	for (int r = 0; r < 3; r++)
	{
		cudaEventRecord(a); CK(cudaMemcpyAsync(dBuf, hO, (size_t)n * 16, cudaMemcpyHostToDevice)); cudaEventRecord(b); CK(cudaEventSynchronize(b));
		float ms; cudaEventElapsedTime(&ms, a, b); if (r == 2) printf("DMA H2D %.1f GB/s\n", n * 16.0 / ms / 1e6);
		cudaEventRecord(a); CK(cudaMemcpyAsync(hH, dBuf, (size_t)n * 16, cudaMemcpyDeviceToHost)); cudaEventRecord(b); CK(cudaEventSynchronize(b));
		cudaEventElapsedTime(&ms, a, b); if (r == 2) printf("DMA D2H %.1f GB/s\n", n * 16.0 / ms / 1e6);
	}
	for (int ctas = 148; ctas <= 888; ctas *= 2)
		for (int rpf = 32; rpf >= 4; rpf /= 2)
		{
			float best = 1e9;
			for (int r = 0; r < 4; r++)
			{
				CK(cudaMemset(dCounter, 0, 4));
				cudaEventRecord(a); readKernel<<<ctas, 128>>>(hO, hD, n, rpf, dCounter, dSink); cudaEventRecord(b); CK(cudaEventSynchronize(b));
				float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
			}
			printf("read  ctas %4d rays/fetch %2d: %.3f ms  %.1f GB/s\n", ctas, rpf, best, n * 32.0 / best / 1e6);
		}
	for (int streaming = 0; streaming < 2; streaming++)
		for (int window = 1; window <= 65536; window *= 16)
		{
			float best = 1e9;
			for (int r = 0; r < 4; r++)
			{
				cudaEventRecord(a); writeKernel<<<888, 128>>>(hH, n, window, streaming); cudaEventRecord(b); CK(cudaEventSynchronize(b));
				float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
			}
			printf("write window %6d %s: %.3f ms  %.1f GB/s\n", window, streaming ? "stcs" : "st  ", best, n * 16.0 / best / 1e6);
		}
	return 0;
}
