// Dev probe: would F16 accumulators help the association filter?  kind::f16 with F16 inputs and an F16 accumulator keeps one 16-bit
// element per 32-bit TMEM column (hs_probe.cu), so TMEM holds no more of them — but tcgen05.ld .pack::16b would bring two columns per
// register (half the registers per row, packed max at twice the element rate).  Checks: (1) the packed load's layout, (2) the error of
// the fp16-accumulated 64-term dot product of unit vectors against fp64, next to the bf16-input / fp32-accumulator error of today's filter.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I ../../boofcv_b200/csrc -o f16acc_probe f16acc_probe.cu
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "bs_sm100.cuh"
using namespace bs::sm100;

struct Bars { uint64_t done; uint32_t tmem; uint32_t pad; };
__host__ __device__ constexpr uint32_t idesc_of(int M, int N, int dfmt, int abfmt) {
	return ((uint32_t)dfmt << 4) | ((uint32_t)abfmt << 7) | ((uint32_t)abfmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[128 x 128] = A[128 x 64] . B[128 x 64]^T; dfmt 0 = F16 accumulator (abfmt 0 = F16 inputs), dfmt 1 = F32 (abfmt 1 = BF16 inputs)
// out_raw: 128 x 128 words read with plain 32x32b.x32; out_packed: 128 x 64 words read with .pack::16b (only meaningful for dfmt 0)
__global__ void __launch_bounds__(128, 1) k(const uint16_t* A, const uint16_t* B, uint32_t* out_raw, uint32_t* out_packed, int dfmt, int abfmt) {
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
		const uint32_t idesc = idesc_of(128, 128, dfmt, abfmt);
		const uint64_t db = umma_desc_sw128(smem_u32(sB));
		for (int kk = 0; kk < 4; kk++) umma_f16_ts(tmem, tmem + 480 + 8 * kk, db + 2 * kk, idesc, kk != 0);
		umma_commit(&bar->done);
	}
	mbar_wait(&bar->done, 0);
	tc_fence_after();
	const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
	for (int c = 0; c < 4; c++) {
		uint32_t r[32];
		tmem_ld32_issue(lane_base + c * 32, r);
		tmem_ld_wait(r);
		for (int j = 0; j < 32; j++) out_raw[threadIdx.x * 128 + c * 32 + j] = r[j];
	}
	for (int c = 0; c < 2; c++) {      // 64 columns -> 32 registers
		uint32_t r[32];
#define O(j) "=r"(r[j])
		asm volatile(
		    "tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "
		    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
		    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
		    : O(0), O(1), O(2), O(3), O(4), O(5), O(6), O(7), O(8), O(9), O(10), O(11), O(12), O(13), O(14), O(15), O(16), O(17), O(18), O(19), O(20),
		      O(21), O(22), O(23), O(24), O(25), O(26), O(27), O(28), O(29), O(30), O(31)
		    : "r"(lane_base + c * 64)
		    : "memory");
#undef O
		tmem_ld_wait(r);
		for (int j = 0; j < 32; j++) out_packed[threadIdx.x * 64 + c * 32 + j] = r[j];
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
static float h2f(uint16_t h) { __half x; *reinterpret_cast<uint16_t*>(&x) = h; return __half2float(x); }

int main() {
	const int smem_bytes = 1024 + 16384 + 64;
	CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
	std::vector<double> A(128 * 64), B(128 * 64);
	srand(11);
	auto unit = [&](std::vector<double>& v) {
		for (int r = 0; r < 128; r++) {
			double n = 0;
			for (int c = 0; c < 64; c++) { double x = (rand() / (double)RAND_MAX) * 2 - 1; x = x * x * x; v[r * 64 + c] = x; n += x * x; }   // peaky like SURF sums
			n = std::sqrt(n);
			for (int c = 0; c < 64; c++) v[r * 64 + c] /= n;
		}
	};
	unit(A); unit(B);
	// rounding of the F16 accumulator: rows 120.. of A against row 120 of B = (1, 1, ..., 1): exact sums 1 + 3 * 2^-12 (round to nearest -> 1 + 2^-10,
	// truncation -> 1) inside one MMA (k < 16) and across two MMAs (second term at k = 16)
	for (int c = 0; c < 64; c++) { B[120 * 64 + c] = 1.0; A[120 * 64 + c] = 0.0; A[121 * 64 + c] = 0.0; A[122 * 64 + c] = 0.0; A[123 * 64 + c] = 0.0; }
	A[120 * 64 + 0] = 1.0; A[120 * 64 + 1] = 3.0 / 4096.0;
	A[121 * 64 + 0] = 1.0; A[121 * 64 + 16] = 3.0 / 4096.0;
	A[122 * 64 + 0] = 1.0; A[122 * 64 + 1] = 1.0 / 4096.0;      // 1 + 2^-12: nearest -> 1, truncation -> 1, round up -> 1 + 2^-10
	A[123 * 64 + 0] = -1.0; A[123 * 64 + 16] = -3.0 / 4096.0;   // negative: truncation (toward zero) -> -1, toward -inf -> -(1 + 2^-10)
	for (int r = 0; r < 32; r++) for (int c = 0; c < 64; c++) B[r * 64 + c] = A[r * 64 + c] + 0.02 * B[r * 64 + c];   // near matches (dot ~ 1)
	uint16_t *dA, *dB; uint32_t *dR, *dP;
	CK(cudaMalloc(&dA, 128 * 64 * 2)); CK(cudaMalloc(&dB, 128 * 64 * 2)); CK(cudaMalloc(&dR, 128 * 128 * 4)); CK(cudaMalloc(&dP, 128 * 64 * 4));
	for (int mode = 0; mode < 2; mode++) {      // 0: F16 in / F16 acc, 1: BF16 in / F32 acc
		std::vector<uint16_t> hA(128 * 64), hB(128 * 64);
		for (int i = 0; i < 128 * 64; i++) {
			if (mode == 0) { __half a = __double2half(A[i]), b = __double2half(B[i]); hA[i] = *reinterpret_cast<uint16_t*>(&a); hB[i] = *reinterpret_cast<uint16_t*>(&b); }
			else { __nv_bfloat16 a = __double2bfloat16(A[i]), b = __double2bfloat16(B[i]); hA[i] = *reinterpret_cast<uint16_t*>(&a); hB[i] = *reinterpret_cast<uint16_t*>(&b); }
		}
		CK(cudaMemcpy(dA, hA.data(), 128 * 64 * 2, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB.data(), 128 * 64 * 2, cudaMemcpyHostToDevice));
		CK(cudaMemset(dR, 0, 128 * 128 * 4)); CK(cudaMemset(dP, 0, 128 * 64 * 4));
		k<<<1, 128, smem_bytes>>>(dA, dB, dR, dP, mode == 0 ? 0 : 1, mode == 0 ? 0 : 1);
		cudaError_t e = cudaDeviceSynchronize();
		if (e != cudaSuccess) { printf("mode %d: kernel failed: %s\n", mode, cudaGetErrorString(e)); return 1; }
		std::vector<uint32_t> R(128 * 128), P(128 * 64);
		CK(cudaMemcpy(R.data(), dR, 128 * 128 * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(P.data(), dP, 128 * 64 * 4, cudaMemcpyDeviceToHost));
		double worst = 0, sum = 0; int bad_pack = 0;
		for (int i = 0; i < 128; i++) for (int j = 0; j < 128; j++) {
			double s = 0, na = 0, nb = 0;
			for (int c = 0; c < 64; c++) { s += A[i * 64 + c] * B[j * 64 + c]; na += A[i * 64 + c] * A[i * 64 + c]; nb += B[j * 64 + c] * B[j * 64 + c]; }
			double got;
			if (mode == 0) {
				got = h2f((uint16_t)(R[i * 128 + j] & 0xffff));
				const uint32_t w = P[i * 64 + j / 2];
				const uint16_t hp = (j & 1) ? (uint16_t)(w >> 16) : (uint16_t)(w & 0xffff);
				if (hp != (uint16_t)(R[i * 128 + j] & 0xffff)) bad_pack++;
			} else {
				got = *reinterpret_cast<float*>(&R[i * 128 + j]);
			}
			const double err = std::fabs(got - s) / std::sqrt(na * nb);
			worst = err > worst ? err : worst; sum += err;
		}
		printf("%s: |s - a.b| / (|a||b|): worst %.3e, mean %.3e (budget of the filter's tolerance: 4e-3)%s\n", mode == 0 ? "F16 inputs, F16 accumulator " : "BF16 inputs, F32 accumulator",
		       worst, sum / (128 * 128), "");
		if (mode == 0) printf("  accumulator rounding: 1 + 3*2^-12 in one MMA -> %.6f, across two MMAs -> %.6f (nearest: 1.000977, truncation: 1.000000); 1 + 2^-12 -> %.6f; -(1 + 3*2^-12) across two -> %.6f\n",
		                      h2f((uint16_t)(R[120 * 128 + 120] & 0xffff)), h2f((uint16_t)(R[121 * 128 + 120] & 0xffff)), h2f((uint16_t)(R[122 * 128 + 120] & 0xffff)), h2f((uint16_t)(R[123 * 128 + 120] & 0xffff)));
		if (mode == 0) printf("  .pack::16b load: register j of the first load = {column 2j (low half), column 2j+1 (high half)}: %d mismatches of 16384\n", bad_pack);
	}
	return 0;
}
