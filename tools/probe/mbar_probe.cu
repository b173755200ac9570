// Dev probe: what a wait on an ALREADY COMPLETED mbarrier phase costs one thread: mbarrier.try_wait, mbarrier.test_wait, and a plain
// shared-memory load for scale.  One thread per CTA arrives on its own barrier (count 1: the phase completes at once) and then waits for it.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I ../../boofcv_b200/csrc -o mbar_probe mbar_probe.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "bs_sm100.cuh"
using namespace bs::sm100;

__global__ void k(int mode, int iters, long long* cycles, uint32_t* sink) {
	__shared__ uint64_t bar;
	__shared__ uint32_t word;
	if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); word = 0; }
	__syncthreads();
	if (threadIdx.x != 0) return;
	const uint32_t addr = smem_u32(&bar), waddr = smem_u32(&word);
	uint32_t acc = 0;
	const long long t0 = clock64();
	for (int it = 0; it < iters; it++) {
		const uint32_t parity = it & 1;
		if (mode != 3) mbar_arrive(&bar);                 // completes phase `it`
		uint32_t done = 0;
		if (mode == 0) {                                  // try_wait until true
			while (!done) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(addr), "r"(parity) : "memory");
		} else if (mode == 1) {                           // test_wait until true
			while (!done) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(addr), "r"(parity) : "memory");
		} else if (mode == 2) {                           // arrive only
		} else {                                          // store + volatile load of a word
			asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(waddr), "r"((uint32_t)it) : "memory");
			asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(done) : "r"(waddr) : "memory");
		}
		acc += done;
	}
	cycles[blockIdx.x] = clock64() - t0;
	if (acc == 0x12345u) sink[0] = acc;
}
int main() {
	long long* d; cudaMalloc(&d, 148 * 8); uint32_t* s; cudaMalloc(&s, 4);
	const char* names[] = {"arrive + try_wait on the completed phase", "arrive + test_wait on the completed phase", "arrive only", "st.volatile + ld.volatile of a shared word"};
	for (int mode = 0; mode < 4; mode++) {
		k<<<148, 64>>>(mode, 20000, d, s);
		cudaError_t e = cudaDeviceSynchronize();
		if (e != cudaSuccess) { printf("%s: %s\n", names[mode], cudaGetErrorString(e)); return 1; }
		long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
		printf("%-50s %.1f cycles per iteration\n", names[mode], (double)h / 20000);
	}
	return 0;
}
