// Dev probe (not part of the library): the accumulator hand-off of the association filter in isolation.  One MMA issuer
// (128 x N x 64 per accumulator = four tcgen05.mma, A in TMEM, B static in shared memory) and 16 epilogue warps in four groups
// (A half x column half, as in filter_kernel) pass S accumulator slots back and forth through mbarriers.  Measures cycles per
// accumulator for S = 2..7 slots, fp32 (N columns per slot) or fp16 accumulators (N/2 columns per slot), with and without the
// TMEM read and a stand-in for the reduction — i.e. how much hand-off latency the slots in flight have to cover.
// Also checks functionally that kind::f16 with an F16 accumulator packs columns (2j, 2j+1) into TMEM column j.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I ../../boofcv_b200/csrc -o hs_probe hs_probe.cu
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "bs_sm100.cuh"

using namespace bs::sm100;

struct Bars { uint64_t full[8], empty[8], done; uint32_t tmem; uint32_t pad; };

// D format: 1 = F32, 0 = F16; A/B format: 1 = BF16, 0 = F16
__host__ __device__ constexpr uint32_t idesc_of(int M, int N, int dfmt, int abfmt) {
	return ((uint32_t)dfmt << 4) | ((uint32_t)abfmt << 7) | ((uint32_t)abfmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// slots: accumulators in flight; cols: TMEM columns per slot (128 for fp32 N=128, 64 for fp16 N=128); ld: read TMEM; work: dummy
// reduction steps per accumulator and thread
__global__ void __launch_bounds__(576, 1) hs_kernel(int slots, int cols, int dfmt, int ld, int work, int iters, long long* cycles, int flags) {
	extern __shared__ unsigned char raw[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
	unsigned char* sB = smem;         // 128 rows x 128 B
	Bars* bar = reinterpret_cast<Bars*>(smem + 16384);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	for (int i = threadIdx.x; i < 16384 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
	if (threadIdx.x == 0) {
		for (int s = 0; s < 8; s++) { mbar_init(&bar->full[s], 1); mbar_init(&bar->empty[s], 8); }
		mbar_init(&bar->done, 1);
		mbar_fence_init();
	}
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	__syncwarp();
	if (warp == 0) tmem_alloc<512>(&bar->tmem);
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = bar->tmem;
	if (warp == 1) {
		if (elect_one()) {
			const uint32_t idesc = idesc_of(128, cols, dfmt, 1);
			const uint64_t db = umma_desc_sw128(smem_u32(sB));
			const uint32_t a_t = tmem + 448;
			const long long t0 = clock64();
			for (int it = 0; it < iters; it++) {
				const int slot = it % slots;
				if (!(flags & 2)) mbar_wait(&bar->empty[slot], ((it / slots) & 1) ^ 1);
				if (!(flags & 1)) tc_fence_after();
				const uint32_t d = tmem + (uint32_t)(slot * cols);
#pragma unroll
				for (int kk = 0; kk < 4; kk++) umma_f16_ts(d, a_t + (uint32_t)((it & 1) * 32) + 8 * kk, db + 2 * kk, idesc, kk != 0);
				if (!(flags & 8)) umma_commit(&bar->full[slot]);
			}
			umma_commit(&bar->done);
			mbar_wait(&bar->done, 0);
			cycles[blockIdx.x] = clock64() - t0;
		}
	} else if (warp >= 2) {
		const int q = warp & 3, g = (warp - 2) >> 2, h = g >> 1, ch = g & 1;
		const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
		float run = 0.0f;
		for (int it = h; it < iters; it += 2) {
			const int slot = it % slots;
			if (!(flags & 4)) mbar_wait(&bar->full[slot], (it / slots) & 1);
			tc_fence_after();
			uint32_t r0[32], r1[32];
#pragma unroll
			for (int k = 0; k < 32; k++) { r0[k] = k + it; r1[k] = k * 3 + it; }
			if (ld) {
				const uint32_t tb = lane_base + (uint32_t)(slot * cols + ch * (cols / 2));
				tmem_ld32_issue(tb, r0);
				if (cols >= 112) tmem_ld32_issue(tb + (cols / 2 - 32), r1);
				tmem_ld_wait(r0);
				if (cols >= 112) tmem_ld_wait(r1);
			}
			tc_fence_before();
			__syncwarp();
			if (lane == 0) mbar_arrive(&bar->empty[slot]);
			if (work) {
				float a[11];
#pragma unroll
				for (int i = 0; i < 10; i++) a[i] = fmax3(__uint_as_float(r0[3 * i]), __uint_as_float(r0[3 * i + 1]), __uint_as_float(r0[3 * i + 2]));
				a[10] = fmaxf(__uint_as_float(r0[30]), __uint_as_float(r0[31]));
				float m = fmaxf(fmax3(fmax3(a[0], a[1], a[2]), fmax3(a[3], a[4], a[5]), fmax3(a[6], a[7], a[8])), fmaxf(a[9], a[10]));
				if (work > 1) {
#pragma unroll
					for (int i = 0; i < 10; i++) a[i] = fmax3(__uint_as_float(r1[3 * i]), __uint_as_float(r1[3 * i + 1]), __uint_as_float(r1[3 * i + 2]));
					a[10] = fmaxf(__uint_as_float(r1[30]), __uint_as_float(r1[31]));
					m = fmaxf(m, fmaxf(fmax3(fmax3(a[0], a[1], a[2]), fmax3(a[3], a[4], a[5]), fmax3(a[6], a[7], a[8])), fmaxf(a[9], a[10])));
				}
				run = fmaxf(run, m);
			}
		}
		if (run == 123.456f) cycles[0] = 0;
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

// functional: D[128 x 128] (F16 accumulator) = A[128 x 64] . B[128 x 64]^T, A/B F16 (abfmt 0) or BF16 (abfmt 1)
__global__ void __launch_bounds__(128, 1) f16_check_kernel(const uint16_t* A, const uint16_t* B, uint32_t* D, int abfmt) {
	extern __shared__ unsigned char raw[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
	unsigned char* sB = smem;
	Bars* bar = reinterpret_cast<Bars*>(smem + 16384);
	const int warp = threadIdx.x >> 5;
	for (int i = threadIdx.x; i < 128 * 8; i += blockDim.x) {
		const int r = i >> 3, c = i & 7;
		const uint4 v = *reinterpret_cast<const uint4*>(B + r * 64 + c * 8);
		*reinterpret_cast<uint4*>(sB + r * 128 + ((c ^ (r & 7)) << 4)) = v;
	}
	if (threadIdx.x == 0) { mbar_init(&bar->done, 1); mbar_fence_init(); }
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	__syncwarp();
	if (warp == 0) tmem_alloc<512>(&bar->tmem);
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = bar->tmem;
	{
		uint32_t r[32];
		const uint32_t* a = reinterpret_cast<const uint32_t*>(A + threadIdx.x * 64);
		for (int j = 0; j < 32; j++) r[j] = a[j];
		tmem_st32(tmem + ((uint32_t)(warp * 32) << 16) + 480, r);
	}
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	if (threadIdx.x == 0) {
		const uint32_t idesc = idesc_of(128, 128, 0, abfmt);
		const uint64_t db = umma_desc_sw128(smem_u32(sB));
		for (int kk = 0; kk < 4; kk++) umma_f16_ts(tmem, tmem + 480 + 8 * kk, db + 2 * kk, idesc, kk != 0);
		umma_commit(&bar->done);
	}
	mbar_wait(&bar->done, 0);
	tc_fence_after();
	for (int c = 0; c < 4; c++) {      // dump 128 TMEM columns raw
		uint32_t r[32];
		tmem_ld32_issue(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, r);
		tmem_ld_wait(r);
		for (int k = 0; k < 32; k++) D[threadIdx.x * 128 + c * 32 + k] = r[k];
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

int main() {
	const int smem_bytes = 1024 + 16384 + 256;
	CK(cudaFuncSetAttribute(hs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
	CK(cudaFuncSetAttribute(f16_check_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
	long long* d_cycles; CK(cudaMalloc(&d_cycles, 148 * 8));
	for (int abfmt = 0; abfmt < 1; abfmt++) {   // BF16 inputs with an F16 accumulator: illegal instruction (measured)
		std::vector<uint16_t> hA(128 * 64), hB(128 * 64);
		std::vector<float> fA(128 * 64), fB(128 * 64);
		srand(5);
		for (int i = 0; i < 128 * 64; i++) {
			fA[i] = (float)((rand() % 17) - 8) / 8.0f; fB[i] = (float)((rand() % 13) - 6) / 4.0f;
			if (abfmt) { __nv_bfloat16 a = __float2bfloat16(fA[i]), b = __float2bfloat16(fB[i]); hA[i] = *reinterpret_cast<uint16_t*>(&a); hB[i] = *reinterpret_cast<uint16_t*>(&b); }
			else { __half a = __float2half(fA[i]), b = __float2half(fB[i]); hA[i] = *reinterpret_cast<uint16_t*>(&a); hB[i] = *reinterpret_cast<uint16_t*>(&b); }
		}
		uint16_t *dA, *dB; uint32_t* dD;
		CK(cudaMalloc(&dA, 128 * 64 * 2)); CK(cudaMalloc(&dB, 128 * 64 * 2)); CK(cudaMalloc(&dD, 128 * 128 * 4));
		CK(cudaMemset(dD, 0, 128 * 128 * 4));
		CK(cudaMemcpy(dA, hA.data(), 128 * 64 * 2, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB.data(), 128 * 64 * 2, cudaMemcpyHostToDevice));
		f16_check_kernel<<<1, 128, smem_bytes>>>(dA, dB, dD, abfmt);
		cudaError_t e = cudaDeviceSynchronize();
		if (e != cudaSuccess) { printf("F16 accumulator with %s inputs: kernel failed: %s\n", abfmt ? "BF16" : "F16", cudaGetErrorString(e)); return 1; }
		std::vector<uint32_t> hD(128 * 128);
		CK(cudaMemcpy(hD.data(), dD, 128 * 128 * 4, cudaMemcpyDeviceToHost));
		// hypothesis P: packed, TMEM column j = {element 2j (low half), element 2j+1 (high half)}, 64 columns used
		// hypothesis U: one element per column in the low half, 128 columns used
		int badP = 0, badU = 0;
		for (int i = 0; i < 128; i++) for (int j = 0; j < 128; j++) {
			float s = 0; for (int k = 0; k < 64; k++) s += fA[i * 64 + k] * fB[j * 64 + k];
			const uint32_t wp = hD[i * 128 + j / 2];
			const uint16_t hp = (j & 1) ? (uint16_t)(wp >> 16) : (uint16_t)(wp & 0xffff);
			const uint16_t hu = (uint16_t)(hD[i * 128 + j] & 0xffff);
			if (__half2float(*reinterpret_cast<const __half*>(&hp)) != s) badP++;
			if (__half2float(*reinterpret_cast<const __half*>(&hu)) != s) badU++;
		}
		printf("F16 accumulator, %s inputs: packed-pairs layout %d mismatches, one-per-column layout %d mismatches (D[0][0..3] raw: %08x %08x %08x %08x)\n",
		       abfmt ? "BF16" : "F16", badP, badU, hD[0], hD[1], hD[2], hD[3]);
	}
	struct Case { int slots, cols, dfmt, ld, work, flags; const char* name; } cases[] = {
	    {3, 128, 1, 1, 0, 0, "3 slots, ld"}, {3, 128, 1, 1, 0, 1, "3 slots, ld, no issuer fence"},
	    {3, 128, 1, 0, 0, 2 + 4, "no empty wait, no full wait"}, {3, 128, 1, 0, 0, 2 + 4 + 8, "no waits, no commit"},
	    {3, 128, 1, 0, 0, 4, "empty wait only (epilogue free-runs)"}, {3, 128, 1, 0, 0, 2, "full wait only (issuer free-runs)"}};
	const int iters = 20000;
	for (int grid : {148})
		for (auto& c : cases) {
			CK(cudaMemset(d_cycles, 0, 148 * 8));
			hs_kernel<<<grid, 576, smem_bytes>>>(c.slots, c.cols, c.dfmt, c.ld, c.work, iters, d_cycles, c.flags);
			CK(cudaDeviceSynchronize());
			long long h; CK(cudaMemcpy(&h, d_cycles, 8, cudaMemcpyDeviceToHost));
			printf("grid %3d  %-30s %.1f cycles per accumulator (128x128x64; MMA floor 256, measured alone 292)\n", grid, c.name, (double)h / iters);
		}
	return 0;
}
