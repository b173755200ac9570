// Dev probe: issue rate of the min/max instructions the filter's epilogue is made of (per SM sub-partition, 4 warps each, 8 independent
// chains per thread), next to FFMA for scale.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I ../../boofcv_b200/csrc -o alu_probe alu_probe.cu
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include "bs_sm100.cuh"
using namespace bs::sm100;

template <int OP>
__global__ void __launch_bounds__(512, 1) k(int iters, float seed, long long* cycles, float* sink) {
	float a[8];
#pragma unroll
	for (int i = 0; i < 8; i++) a[i] = seed * (float)(threadIdx.x + i);
	float b = seed + 1.0f, c = seed + 2.0f;
	uint32_t u[8];
#pragma unroll
	for (int i = 0; i < 8; i++) u[i] = __float_as_uint(a[i]);
	const long long t0 = clock64();
	for (int it = 0; it < iters; it++) {
#pragma unroll
		for (int i = 0; i < 8; i++) {
			if (OP == 0) a[i] = fmax3(a[i], b, c);
			if (OP == 1) a[i] = fmaxf(a[i], b);
			if (OP == 2) a[i] = fmaf(a[i], b, c);
			if (OP == 3) { asm volatile("max.bf16x2 %0, %0, %1;" : "+r"(u[i]) : "r"(__float_as_uint(b))); }
			if (OP == 4) { asm volatile("max.f16x2 %0, %0, %1;" : "+r"(u[i]) : "r"(__float_as_uint(b))); }
			if (OP == 5) { int v = (int)u[i]; asm volatile("max.s32 %0, %0, %1;" : "+r"(v) : "r"((int)__float_as_uint(b))); u[i] = (uint32_t)v; }
			if (OP == 6) { a[i] = fmax3(a[i], b, c); a[i] = fmaf(a[i], b, c); }          // one of each pipe
			if (OP == 7) { a[i] = fmax3(a[i], b, c); asm volatile("max.bf16x2 %0, %0, %1;" : "+r"(u[i]) : "r"(__float_as_uint(b))); }   // do they share a pipe?
		}
		b += 1e-9f;
	}
	const long long dt = clock64() - t0;
	if (threadIdx.x == 0) cycles[blockIdx.x] = dt;
	float s = 0;
#pragma unroll
	for (int i = 0; i < 8; i++) s += a[i] + __uint_as_float(u[i]);
	if (s == 1.2345f) sink[0] = s;
}

int main() {
	long long* d; cudaMalloc(&d, 148 * 8);
	float* sink; cudaMalloc(&sink, 4);
	const int iters = 20000;
	const char* names[] = {"fmax3 (FMNMX3)", "fmaxf (FMNMX)", "fmaf (FFMA)", "max.bf16x2", "max.f16x2", "max.s32", "fmax3 + fmaf pairs (per pair)", "fmax3 + max.bf16x2 pairs (per pair)"};
	for (int op = 0; op < 8; op++)
		for (int threads : {128, 256, 512}) {
			cudaMemset(d, 0, 148 * 8);
			switch (op) {
				case 0: k<0><<<148, threads>>>(iters, 0.5f, d, sink); break;
				case 1: k<1><<<148, threads>>>(iters, 0.5f, d, sink); break;
				case 2: k<2><<<148, threads>>>(iters, 0.5f, d, sink); break;
				case 3: k<3><<<148, threads>>>(iters, 0.5f, d, sink); break;
				case 4: k<4><<<148, threads>>>(iters, 0.5f, d, sink); break;
				case 5: k<5><<<148, threads>>>(iters, 0.5f, d, sink); break;
				case 6: k<6><<<148, threads>>>(iters, 0.5f, d, sink); break;
				case 7: k<7><<<148, threads>>>(iters, 0.5f, d, sink); break;
			}
			cudaError_t e = cudaDeviceSynchronize();
			if (e != cudaSuccess) { printf("%s: %s\n", names[op], cudaGetErrorString(e)); return 1; }
			long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
			const double per = (double)h / ((double)iters * 8 * (threads / 128));    // cycles per warp instruction per sub-partition
			printf("%-32s %d warps per sub-partition: %.2f cycles per warp instruction\n", names[op], threads / 128, per);
		}
	return 0;
}
