// Dev probe (not part of the library): tcgen05 rates on this GPU — TMEM read bandwidth (tcgen05.ld), UMMA issue rate with
// the A operand in shared memory (SS) or in TMEM (TS), N = 128 / 256, alone and with concurrent TMEM reads — and a functional
// check of the TMEM layout of a bf16 A operand (lane = row, 32-bit column j = elements k = 2j (low half), 2j+1 (high half)).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I ../../boofcv_b200/csrc -o tc_probe tc_probe.cu
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "bs_sm100.cuh"

using namespace bs::sm100;

struct Bars { uint64_t done; uint32_t tmem; uint32_t pad; };

// mma: 0 none, 1 SS, 2 TS; N: columns per MMA; chains: accumulators the consecutive MMAs rotate over (1 = the four K steps of one
// tile back to back on one accumulator, 2 = K steps of two tiles interleaved); ld_warps: warps reading TMEM concurrently
__global__ void __launch_bounds__(320, 1) rate_kernel(int mma, int N, int chains, int ld_warps, int iters, long long* cycles) {
	extern __shared__ unsigned char raw[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
	unsigned char* sA = smem;                 // 128 rows x 128 B
	unsigned char* sB = smem + 16384;         // 256 rows x 128 B
	Bars* bar = reinterpret_cast<Bars*>(smem + 16384 + 32768);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
	if (threadIdx.x == 0) { mbar_init(&bar->done, 1); mbar_fence_init(); }
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	__syncwarp();
	if (warp == 0) tmem_alloc<512>(&bar->tmem);
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = bar->tmem;
	long long t0 = 0, t1 = 0;
	if (warp == 1 && mma) {
		if (elect_one()) {
			const uint32_t idesc = umma_idesc_bf16(128, N);
			const uint64_t da = umma_desc_sw128(smem_u32(sA)), db = umma_desc_sw128(smem_u32(sB));
			const uint32_t a_t = tmem + 480;
			t0 = clock64();
			for (int it = 0; it < iters; it++) {
				if (chains == 1) {
					const uint32_t d = tmem + (uint32_t)((it & 1) * (2 * N <= 480 ? N : 0));
#pragma unroll
					for (int kk = 0; kk < 4; kk++) {
						if (mma == 1) umma_bf16(d, da + 2 * kk, db + 2 * kk, idesc, kk != 0);
						else umma_f16_ts(d, a_t + 8 * kk, db + 2 * kk, idesc, kk != 0);
					}
				} else {
#pragma unroll
					for (int kk = 0; kk < 4; kk++)
						for (int c = 0; c < 2; c++) {       // two iterations' worth of MMAs per loop trip
							const uint32_t d = tmem + (uint32_t)(c * N);
							if (mma == 1) umma_bf16(d, da + 2 * kk, db + 2 * kk, idesc, kk != 0);
							else umma_f16_ts(d, a_t + 8 * kk, db + 2 * kk, idesc, kk != 0);
						}
					it++;
				}
			}
			umma_commit(&bar->done);
			mbar_wait(&bar->done, 0);
			t1 = clock64();
			cycles[blockIdx.x * 2 + 0] = t1 - t0;
		}
	} else if (warp >= 2 && warp < 2 + ld_warps) {
		const int q = warp & 3;
		const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + 256;
		uint32_t ra[32], rb[32];
		uint32_t acc = 0;
		t0 = clock64();
		for (int it = 0; it < iters; it++) {
			tmem_ld32_issue(tbase, ra);
			tmem_ld32_issue(tbase + 32, rb);
			tmem_ld_wait(ra); tmem_ld_wait(rb);
			acc += ra[3] ^ rb[17];
			tmem_ld32_issue(tbase + 64, ra);
			tmem_ld32_issue(tbase + 96, rb);
			tmem_ld_wait(ra); tmem_ld_wait(rb);
			acc += ra[5] ^ rb[30];
		}
		t1 = clock64();
		if (lane == 0 && warp == 2) cycles[blockIdx.x * 2 + 1] = t1 - t0;
		if (acc == 0x12345678u) cycles[0] = 0;
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

// functional: D[128 x 128] = A[128 x 64] . B[128 x 64]^T with A written to TMEM by tcgen05.st, B in smem (128-byte swizzle)
__global__ void __launch_bounds__(128, 1) ts_check_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D) {
	extern __shared__ unsigned char raw[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
	unsigned char* sB = smem;
	Bars* bar = reinterpret_cast<Bars*>(smem + 16384);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	// B: row r at r*128 bytes, 16-byte chunk c stored at chunk (c ^ (r & 7))
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
	{   // A row = TMEM lane = threadIdx.x; column j = (k = 2j, 2j+1)
		uint32_t r[32];
		const uint32_t* a = reinterpret_cast<const uint32_t*>(A + threadIdx.x * 64);
		for (int j = 0; j < 32; j++) r[j] = a[j];
		tmem_st32(tmem + ((uint32_t)(warp * 32) << 16) + 480, r);
	}
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	if (threadIdx.x == 0) {
		const uint32_t idesc = umma_idesc_bf16(128, 128);
		const uint64_t db = umma_desc_sw128(smem_u32(sB));
		for (int kk = 0; kk < 4; kk++) umma_f16_ts(tmem, tmem + 480 + 8 * kk, db + 2 * kk, idesc, kk != 0);
		umma_commit(&bar->done);
	}
	mbar_wait(&bar->done, 0);
	tc_fence_after();
	for (int c = 0; c < 4; c++) {
		uint32_t r[32];
		tmem_ld32_issue(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, r);
		tmem_ld_wait(r);
		for (int k = 0; k < 32; k++) D[threadIdx.x * 128 + c * 32 + k] = __uint_as_float(r[k]);
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

int main() {
	const int smem_bytes = 1024 + 16384 + 32768 + 64;
	CK(cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
	CK(cudaFuncSetAttribute(ts_check_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
	long long* d_cycles; CK(cudaMalloc(&d_cycles, 148 * 2 * 8));
	// ---- functional check
	{
		std::vector<__nv_bfloat16> hA(128 * 64), hB(128 * 64);
		std::vector<float> fA(128 * 64), fB(128 * 64);
		srand(5);
		for (int i = 0; i < 128 * 64; i++) {
			fA[i] = (float)((rand() % 17) - 8) / 8.0f; fB[i] = (float)((rand() % 13) - 6) / 4.0f;
			hA[i] = __float2bfloat16(fA[i]); hB[i] = __float2bfloat16(fB[i]);
		}
		__nv_bfloat16 *dA, *dB; float* dD;
		CK(cudaMalloc(&dA, 128 * 64 * 2)); CK(cudaMalloc(&dB, 128 * 64 * 2)); CK(cudaMalloc(&dD, 128 * 128 * 4));
		CK(cudaMemcpy(dA, hA.data(), 128 * 64 * 2, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB.data(), 128 * 64 * 2, cudaMemcpyHostToDevice));
		ts_check_kernel<<<1, 128, smem_bytes>>>(dA, dB, dD);
		CK(cudaDeviceSynchronize());
		std::vector<float> hD(128 * 128);
		CK(cudaMemcpy(hD.data(), dD, 128 * 128 * 4, cudaMemcpyDeviceToHost));
		int bad = 0;
		for (int i = 0; i < 128; i++) for (int j = 0; j < 128; j++) {
			float s = 0; for (int k = 0; k < 64; k++) s += fA[i * 64 + k] * fB[j * 64 + k];
			if (s != hD[i * 128 + j]) { if (bad < 5) printf("mismatch D[%d][%d] = %g expected %g\n", i, j, hD[i * 128 + j], s); bad++; }
		}
		printf("TS functional check (A in TMEM, lane=row, col j = k 2j|2j+1): %s (%d mismatches)\n", bad ? "FAILED" : "ok", bad);
	}
	// ---- rates
	struct Case { int mma, N, chains, ld; const char* name; } cases[] = {
	    {0, 128, 1, 1, "LDTM only, 1 warp"}, {0, 128, 1, 4, "LDTM only, 4 warps"}, {0, 128, 1, 8, "LDTM only, 8 warps"},
	    {1, 128, 1, 0, "SS N=128, 1 chain"}, {1, 128, 2, 0, "SS N=128, 2 chains"}, {1, 256, 1, 0, "SS N=256, 1 chain"}, {1, 256, 2, 0, "SS N=256, 2 chains"},
	    {2, 128, 1, 0, "TS N=128, 1 chain"}, {2, 128, 2, 0, "TS N=128, 2 chains"}, {2, 192, 2, 0, "TS N=192, 2 chains"},
	    {2, 224, 1, 0, "TS N=224, 1 chain"}, {2, 224, 2, 0, "TS N=224, 2 chains"}, {2, 240, 2, 0, "TS N=240, 2 chains"}, {2, 256, 1, 0, "TS N=256, 1 chain"},
	    {1, 64, 2, 0, "SS N=64, 2 chains"}, {2, 64, 2, 0, "TS N=64, 2 chains"},
	    {1, 128, 2, 8, "SS N=128, 2 chains + LDTM 8 warps"}, {2, 128, 2, 8, "TS N=128, 2 chains + LDTM 8 warps"}, {2, 240, 2, 8, "TS N=240, 2 chains + LDTM 8 warps"}};
	const int iters = 20000;
	for (int grid : {1, 148})
		for (auto& c : cases) {
			CK(cudaMemset(d_cycles, 0, 148 * 2 * 8));
			rate_kernel<<<grid, 320, smem_bytes>>>(c.mma, c.N, c.chains, c.ld, iters, d_cycles);
			CK(cudaDeviceSynchronize());
			long long h[2]; CK(cudaMemcpy(h, d_cycles, 16, cudaMemcpyDeviceToHost));
			printf("grid %3d  %-36s", grid, c.name);
			if (c.mma) printf("  %.1f cycles per 128x%dx16 MMA (floor %d)", (double)h[0] / (iters * 4.0), c.N, 128 * c.N / 256);
			if (c.ld) printf("  LDTM %.1f B/cycle per warp, %.1f B/cycle per SM", iters * 16384.0 / h[1], iters * 16384.0 / h[1] * c.ld);
			printf("\n");
		}
	return 0;
}
