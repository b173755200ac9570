// Dev probe (not part of the library): what the accumulator hand-off of filter_kernel costs, piece by piece.
//  micro 1   one thread: MMAs + tcgen05.commit + wait for that commit              (completion signal: ~245 cycles; ~95 per dependent MMA)
//  micro 3/4 a pipeline of D groups in flight, the issuing thread waits (mbarrier try_wait) for the group D back, directly or relayed by
//            another warp: 625 cycles per group for EVERY D — a try_wait of the thread that issued a commit does not return before that
//            commit has arrived, so the pipe drains between two groups of one issuer; 380 without the wait
//  micro 5   a 200-cycle pause of the issuing thread per group costs 134 cycles: the pipe takes MMAs from the thread almost synchronously
//  micro 6   the issuer polls a shared-memory word a relay warp writes after ITS mbarrier wait: 366 cycles per group at D >= 2 (pipe-bound).
//            In filter_kernel the extra relay hop lengthened the slot round trip by more than it saved (3.27 against 2.69 ms): not kept.
//  micro 2   mbarrier arrive -> try_wait hop between two warps: 85-100 cycles
//  ldtm      tcgen05.ld throughput with 1..16 warps
// The first section (a skeleton of the whole hand-off: two or four issuers, four slots, sixteen epilogue warps) is kept for reference only:
// its result moved between 1200 and 2060 cycles per tile from one build of this file to the next (register arrays in local memory, code
// layout), so nothing is concluded from it.  Results: profiles/r2c_hs2_probe.txt.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I ../../boofcv_b200/csrc -o hs2_probe hs2_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include "bs_sm100.cuh"

using namespace bs::sm100;

// wait variants: 0 = try_wait spin (the library's), 1 = try_wait with a long suspend-time hint, 2 = test_wait + nanosleep(20),
// 3 = lane 0 polls, the warp waits at __syncwarp
__device__ __forceinline__ void wait_v(uint64_t* bar, uint32_t parity, int mode) {
	const uint32_t addr = smem_u32(bar);
	if (mode == 0) { mbar_wait(bar, parity); return; }
	uint32_t spin = 0;
	if (mode == 1) {
		uint32_t done = 0;
		while (!done) {
			asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(addr), "r"(parity), "r"(0x989680u) : "memory");
			if (++spin == (1u << 22)) __trap();
		}
	} else if (mode == 2) {
		uint32_t done = 0;
		while (true) {
			asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(addr), "r"(parity) : "memory");
			if (done) break;
			__nanosleep(20);
			if (++spin == (1u << 22)) __trap();
		}
	} else {
		if ((threadIdx.x & 31) == 0) mbar_wait(bar, parity);
		__syncwarp();
	}
}

struct Bars { uint64_t full[4], empty[4], done[2]; uint32_t tmem; uint32_t pad; };

__host__ __device__ constexpr uint32_t idesc_of(int M, int N) {
	return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// flags: 1 = no TMEM reads, 2 = no reduction, 4 = release the slot after the FIRST round (wrong, timing only), 8 = one round of 64 columns only,
// 16 = issuers do not wait for empty (epilogue still waits for full), 32 = spin with nanosleep-free plain test_wait
__global__ void __launch_bounds__(736, 1) hs2_kernel(int cols, int iters, long long* cycles, int flags, int wmode, int nwarps) {
	extern __shared__ unsigned char raw[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
	unsigned char* sB = smem;         // 128 rows x 128 B, zeros
	Bars* bar = reinterpret_cast<Bars*>(smem + 16384);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	for (int i = threadIdx.x; i < 16384 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
	if (threadIdx.x == 0) {
		for (int s = 0; s < 4; s++) { mbar_init(&bar->full[s], 1); mbar_init(&bar->empty[s], nwarps); }
		mbar_init(&bar->done[0], (flags & 128) ? 4 : 1); mbar_init(&bar->done[1], 1);
		mbar_fence_init();
	}
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	__syncwarp();
	if (warp == 0) tmem_alloc<512>(&bar->tmem);
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = bar->tmem;
	if ((flags & 128) && (warp == 1 || (warp >= 18 && warp <= 20))) {
		// four issuers: one per slot (half, tile parity)
		if (elect_one()) {
			const uint32_t id = warp == 1 ? 0u : (uint32_t)(warp - 17);
			const uint32_t h = id >> 1, par = id & 1u, slot = id;
			const uint32_t idesc = idesc_of(128, cols);
			const uint64_t db = umma_desc_sw128(smem_u32(sB));
			const uint32_t a_t = tmem + 448 + h * 32;
			const uint32_t d = tmem + slot * (uint32_t)cols;
			const long long t0 = clock64();
			uint32_t k = 0;
			for (int n = (int)par; n < iters; n += 2, k++) {
				if (k > 0) wait_v(&bar->empty[slot], (k - 1) & 1, wmode);
				tc_fence_after();
#pragma unroll
				for (int kk = 0; kk < 4; kk++) umma_f16_ts(d, a_t + 8 * kk, db + 2 * kk, idesc, kk != 0);
				umma_commit(&bar->full[slot]);
			}
			umma_commit(&bar->done[0]);
			if (id == 0) { mbar_wait(&bar->done[0], 0); cycles[blockIdx.x * 2] = clock64() - t0; }
		}
	} else if ((warp == 1 || warp == 18) && !((flags & 32) && warp == 18)) {
		if (elect_one()) {
			const uint32_t h = warp == 1 ? 0u : 1u;
			const uint32_t idesc = idesc_of(128, cols);
			const uint64_t db = umma_desc_sw128(smem_u32(sB));
			const uint32_t a_t = tmem + 448 + h * 32;
			const long long t0 = clock64();
			for (int n = 0; n < iters; n++) {
				const uint32_t slot = 2u * h + ((uint32_t)n & 1u);
				const uint32_t u = (uint32_t)n >> 1;
				if (u > 0) wait_v(&bar->empty[slot], (u - 1) & 1, wmode);
				tc_fence_after();
				const uint32_t d = tmem + slot * (uint32_t)cols;
#pragma unroll
				for (int kk = 0; kk < 4; kk++) if (kk == 0 || !(flags & 64)) umma_f16_ts(d, a_t + 8 * kk, db + 2 * kk, idesc, kk != 0);
				umma_commit(&bar->full[slot]);
			}
			umma_commit(&bar->done[h]);
			mbar_wait(&bar->done[h], 0);
			cycles[blockIdx.x * 2 + h] = clock64() - t0;
			if (flags & 32) cycles[blockIdx.x * 2 + 1] = 0;
		}
	} else if (warp >= 2 && warp < 18 && ((warp - 2) & 3) < nwarps && !((flags & 32) && warp >= 10)) {
		const int q = warp & 3, g = (warp - 2) >> 2, h = g >> 1, par = g & 1;
		const uint32_t slot = (uint32_t)(2 * h + par);
		const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16) + slot * (uint32_t)cols;
		float run = 0.0f;
		uint32_t k = 0;
		for (int n = par; n < iters; n += 2, k++) {
			wait_v(&bar->full[slot], k & 1, wmode);
			tc_fence_after();
#pragma unroll 1
			for (int c2 = 0; c2 < 2; c2++) {
				uint32_t r0[32], r1[32];
#pragma unroll
				for (int i = 0; i < 32; i++) { r0[i] = i + n; r1[i] = i * 3 + n; }
				if (!(flags & 1) && !((flags & 8) && c2 == 1)) {
					const uint32_t tb = lane_base + (uint32_t)(c2 * 64);
					tmem_ld32_issue(tb, r0);
					if (c2 == 0) tmem_ld32_issue(tb + 32, r1);
					else tmem_ld16_issue(tb + 32, r1);
					tmem_ld_wait(r0); tmem_ld_wait(r1);
				}
				if (c2 == ((flags & 4) ? 0 : 1)) {
					tc_fence_before();
					__syncwarp();
					if (lane == 0) mbar_arrive(&bar->empty[slot]);
				}
				if (!(flags & 2)) {
					float a[11];
#pragma unroll
					for (int i = 0; i < 10; i++) a[i] = fmax3(__uint_as_float(r0[3 * i]), __uint_as_float(r0[3 * i + 1]), __uint_as_float(r0[3 * i + 2]));
					a[10] = fmaxf(__uint_as_float(r0[30]), __uint_as_float(r0[31]));
					float m = fmaxf(fmax3(fmax3(a[0], a[1], a[2]), fmax3(a[3], a[4], a[5]), fmax3(a[6], a[7], a[8])), fmaxf(a[9], a[10]));
#pragma unroll
					for (int i = 0; i < 10; i++) a[i] = fmax3(__uint_as_float(r1[3 * i]), __uint_as_float(r1[3 * i + 1]), __uint_as_float(r1[3 * i + 2]));
					a[10] = fmaxf(__uint_as_float(r1[30]), __uint_as_float(r1[31]));
					m = fmaxf(m, fmaxf(fmax3(fmax3(a[0], a[1], a[2]), fmax3(a[3], a[4], a[5]), fmax3(a[6], a[7], a[8])), fmaxf(a[9], a[10])));
					run = fmaxf(run, m);
				}
			}
		}
		if (run == 123.456f) cycles[0] = 0;
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}


// micro 1: one thread: (MMAs; commit; wait for the commit) in a loop -> MMA + completion signal + wake-up per iteration
// micro 2: two warps ping-pong through two mbarriers (arrive / try_wait) -> two signal hops per iteration
// `pollers` other warps spin on a barrier that never completes (until the end), as the waiting epilogue warps of the filter do
__global__ void __launch_bounds__(608, 1) micro_kernel(int which, int cols, int nmma, int iters, int pollers, long long* cycles, int mflags = 0) {
	// mflags & 16: F16 inputs and an F16 accumulator (all format fields of the instruction descriptor 0) instead of BF16 inputs / F32 accumulator
	extern __shared__ unsigned char raw[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
	unsigned char* sB = smem;
	Bars* bar = reinterpret_cast<Bars*>(smem + 16384);
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	for (int i = threadIdx.x; i < 16384 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
	if (threadIdx.x == 0) {
		for (int s = 0; s < 4; s++) { mbar_init(&bar->full[s], 1); mbar_init(&bar->empty[s], 1); }
		mbar_init(&bar->done[0], 1); mbar_init(&bar->done[1], 1);
		bar->pad = 0;
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
		if (which == 1) {
			if (elect_one()) {
				const uint32_t idesc = idesc_of(128, cols);
				const uint64_t db = umma_desc_sw128(smem_u32(sB));
				const long long t0 = clock64();
				for (int n = 0; n < iters; n++) {
					for (int kk = 0; kk < nmma; kk++) umma_f16_ts(tmem, tmem + 448 + 8 * (kk & 3), db + 2 * (kk & 3), idesc, kk != 0);
					umma_commit(&bar->full[0]);
					mbar_wait(&bar->full[0], n & 1);
					tc_fence_after();
				}
				cycles[blockIdx.x] = clock64() - t0;
				mbar_arrive(&bar->done[0]);
			}
		} else if (which >= 60) {
			// micro 6: the issuer never executes an mbarrier wait: a relay warp (warp 2) waits for the commits and publishes how many groups
			// have completed in a shared-memory word the issuer polls with plain volatile loads; depth D = which - 60 groups in flight
			const int D = which - 60;
			volatile uint32_t* done_ctr = reinterpret_cast<volatile uint32_t*>(&bar->pad);
			if (elect_one()) {
				const uint32_t idesc = idesc_of(128, cols);
				const uint64_t db = umma_desc_sw128(smem_u32(sB));
				const long long t0 = clock64();
				for (int n = 0; n < iters; n++) {
					const int sl = n & 3;
					if (n >= D) { uint32_t spin = 0; while ((int)*done_ctr < n - D + 1) { if (++spin == (1u << 24)) __trap(); } tc_fence_after(); }
					for (int kk = 0; kk < nmma; kk++) umma_f16_ts(tmem + sl * cols, tmem + 448 + 8 * (kk & 3), db + 2 * (kk & 3), idesc, kk != 0);
					umma_commit(&bar->full[sl]);
				}
				uint32_t spin = 0; while ((int)*done_ctr < iters) { if (++spin == (1u << 26)) __trap(); }
				cycles[blockIdx.x] = clock64() - t0;
				mbar_arrive(&bar->done[0]);
			}
		} else if (which >= 50) {
			// micro 5: what exactly costs ~250 cycles per group when the issuer waits?
			//  50: wait on a barrier a free-running partner keeps completing (never blocks for long), commit to a barrier nobody reads
			//  51: no wait, a ~200-cycle clock spin instead
			//  52: MMAs + commit first, THEN wait for the commit of 3 groups ago (depth 4)
			//  53: depth 4 self-paced, waiting with a plain ld.acquire poll of a flag the partner sets after ITS wait (software relay through smem)
			if (elect_one()) {
				const uint32_t idesc = idesc_of(128, cols);
				const uint64_t db = umma_desc_sw128(smem_u32(sB));
				const long long t0 = clock64();
				for (int n = 0; n < iters; n++) {
					const int sl = n & 3;
					if (which == 50) { mbar_wait(&bar->empty[0], n & 1); }
					if (which == 51) { const long long c0 = clock64(); while (clock64() - c0 < 200) {} }
					for (int kk = 0; kk < nmma; kk++) umma_f16_ts(tmem + sl * cols, tmem + 448 + 8 * (kk & 3), db + 2 * (kk & 3), idesc, kk != 0);
					umma_commit(&bar->full[sl]);
					if (which == 52 && n >= 3) { const int m = n - 3; mbar_wait(&bar->full[m & 3], (m >> 2) & 1); }
				}
				umma_commit(&bar->done[1]);
				mbar_wait(&bar->done[1], 0);
				cycles[blockIdx.x] = clock64() - t0;
				mbar_arrive(&bar->done[0]);
			}
		} else if (which >= 30) {
			// micro 3: self-paced pipeline of depth D = which - 30: wait for the commit of D groups ago, then 4 MMAs + commit
			// micro 4 (which >= 40): same, but a partner warp (warp 2) relays the completion: full[s] -> (partner) -> empty[s] -> issuer
			const int relay = which >= 40;
			const int D = which - (relay ? 40 : 30);
			if (elect_one()) {
				const uint32_t idesc = (mflags & 16) ? (((uint32_t)(cols >> 3) << 17) | (8u << 24)) : idesc_of(128, cols);
				const uint64_t db = umma_desc_sw128(smem_u32(sB));
				const long long t0 = clock64();
				for (int n = 0; n < iters; n++) {
					const int sl = n % D, u = n / D;
					// mflags: 1 = no fence after the wait, 2 = no wait, 4 = no commit
					if (u > 0 && !(mflags & 2)) { mbar_wait(relay ? &bar->empty[sl] : &bar->full[sl], (u - 1) & 1); if (!(mflags & 1)) tc_fence_after(); }
					for (int kk = 0; kk < nmma; kk++) umma_f16_ts(tmem + sl * cols, tmem + 448 + 8 * (kk & 3), db + 2 * (kk & 3), idesc, kk != 0);
					if (!(mflags & 4)) umma_commit(&bar->full[sl]);
				}
				umma_commit(&bar->done[1]);
				mbar_wait(&bar->done[1], 0);
				cycles[blockIdx.x] = clock64() - t0;
				mbar_arrive(&bar->done[0]);
			}
		} else {
			const long long t0 = clock64();
			for (int n = 0; n < iters; n++) {
				if (lane == 0) mbar_arrive(&bar->full[0]);
				mbar_wait(&bar->empty[0], n & 1);
			}
			if (lane == 0) { cycles[blockIdx.x] = clock64() - t0; mbar_arrive(&bar->done[0]); }
		}
	} else if (warp == 2 && which >= 60) {
		volatile uint32_t* done_ctr = reinterpret_cast<volatile uint32_t*>(&bar->pad);
		for (int n = 0; n < iters; n++) {
			mbar_wait(&bar->full[n & 3], (n >> 2) & 1);
			if (lane == 0) *done_ctr = (uint32_t)(n + 1);
		}
	} else if (warp == 2 && which == 50) {
		// free-running partner: completes a phase of empty[0] as soon as the issuer has consumed the previous one (ping-pong through full[1])
		for (int n = 0; n < iters; n++) {
			if (lane == 0) mbar_arrive(&bar->empty[0]);
			__nanosleep(0);
			// do not run more than one phase ahead: wait until the issuer's wait can have seen it (no back-channel: sleep a little)
			const long long c0 = clock64(); while (clock64() - c0 < 300) {}
		}
	} else if (warp == 2 && which >= 40 && which < 50) {
		const int D = which - 40;
		for (int n = 0; n < iters; n++) {
			const int sl = n % D, u = n / D;
			mbar_wait(&bar->full[sl], u & 1);
			tc_fence_after();
			tc_fence_before();
			__syncwarp();
			if (lane == 0) mbar_arrive(&bar->empty[sl]);
		}
	} else if (warp == 2 && which == 2) {
		for (int n = 0; n < iters; n++) {
			mbar_wait(&bar->full[0], n & 1);
			if (lane == 0) mbar_arrive(&bar->empty[0]);
		}
	} else if (warp >= 3 && warp < 3 + pollers) {
		mbar_wait(&bar->done[0], 0);
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}


// micro 6: TMEM read throughput.  `nw` warps (lane quarter = warp % 4) read `cols` columns per round with tcgen05.ld 32x32b.x32 (four
// loads in flight when cols = 128, two when 64, one when 32) and consume the registers with one xor each round.
__global__ void __launch_bounds__(512, 1) ldtm_kernel(int nw, int cols, int iters, long long* cycles, uint32_t* sink) {
	__shared__ uint32_t tmem_s;
	const int warp = threadIdx.x >> 5;
	if (warp == 0) tmem_alloc<512>(&tmem_s);
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = tmem_s;
	if (warp < nw) {
		const uint32_t base = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 128 % 512);
		uint32_t acc = 0;
		uint32_t r0[32], r1[32], r2[32], r3[32];
		const long long t0 = clock64();
		for (int it = 0; it < iters; it++) {
			tmem_ld32_issue(base, r0);
			if (cols >= 64) tmem_ld32_issue(base + 32, r1);
			if (cols >= 128) { tmem_ld32_issue(base + 64, r2); tmem_ld32_issue(base + 96, r3); }
			tmem_ld_wait(r0);
			acc ^= r0[it & 31];
			if (cols >= 64) { tmem_ld_wait(r1); acc ^= r1[5]; }
			if (cols >= 128) { tmem_ld_wait(r2); tmem_ld_wait(r3); acc ^= r2[7] ^ r3[9]; }
		}
		const long long dt = clock64() - t0;
		if (threadIdx.x == 0) cycles[blockIdx.x] = dt;
		if (acc == 0x12345u) sink[0] = acc;
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

int main() {
	const int smem_bytes = 1024 + 16384 + 256;
	CK(cudaFuncSetAttribute(hs2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
	long long* d_cycles; CK(cudaMalloc(&d_cycles, 148 * 16));
	struct Case { int cols, flags, wmode, nwarps; const char* name; } cases[] = {
	    {112, 128, 0, 4, "FOUR issuers, full skeleton"}, {112, 128 + 3, 0, 4, "FOUR issuers, hand-off only"}, {112, 128, 1, 4, "FOUR issuers, full skeleton, suspend hint"},
	    {112, 128, 2, 4, "FOUR issuers, full skeleton, test_wait + nanosleep"},
	    {112, 0, 0, 4, "full skeleton"}, {112, 3, 0, 4, "hand-off only (no TMEM reads, no reduction)"},
	    {112, 3, 0, 1, "hand-off only, ONE warp per group"}, {112, 3, 0, 2, "hand-off only, two warps per group"},
	    {32, 3, 0, 4, "hand-off only, 32-column accumulators"}, {64, 3, 0, 4, "hand-off only, 64-column accumulators"}, {128, 3, 0, 4, "hand-off only, 128-column accumulators (slot 3 overlaps A: timing only)"},
	    {112, 3 + 32, 0, 4, "hand-off only, half 0 alone"}, {32, 3 + 32, 0, 4, "hand-off only, half 0 alone, 32 columns"},
	    {112, 3 + 64, 0, 4, "hand-off only, one MMA per accumulator instead of four"}};
	const int iters = 20000;
	for (auto& c : cases) {
		CK(cudaMemset(d_cycles, 0, 148 * 16));
		for (int threads : {608, 640, 672, 736}) {
			if ((c.flags & 128) && threads < 672) continue;
			CK(cudaMemset(d_cycles, 0, 148 * 16));
			hs2_kernel<<<148, threads, smem_bytes>>>(c.cols, iters, d_cycles, c.flags, c.wmode, c.nwarps);
			CK(cudaDeviceSynchronize());
			long long h[2]; CK(cudaMemcpy(h, d_cycles, 16, cudaMemcpyDeviceToHost));
			printf("%-60s %d threads: %.1f cycles per tile\n", c.name, threads, (double)(h[0] > h[1] ? h[0] : h[1]) / iters);
		}
	}
	CK(cudaFuncSetAttribute(micro_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
	struct M { int which, cols, nmma, pollers; const char* name; int mflags; } ms[] = {
	    {1, 112, 4, 0, "4 MMAs 128x112x16 + commit + own wait"}, {1, 112, 1, 0, "1 MMA 128x112x16 + commit + own wait"}, {1, 32, 1, 0, "1 MMA 128x32x16 + commit + own wait"},
	    {1, 112, 0, 0, "commit alone + own wait"}, {1, 112, 4, 15, "4 MMAs + commit + own wait, 15 warps polling"}, {1, 112, 0, 15, "commit alone + own wait, 15 warps polling"},
	    {34, 112, 4, 0, "self-paced depth 4, no fence", 1}, {34, 112, 4, 0, "depth 4, no wait (commit every group)", 2}, {34, 112, 4, 0, "depth 4, no wait, no commit", 6},
	    {34, 112, 4, 0, "depth 4, wait+fence but no commit (waits on phase never... skipped)", 2 + 4}, {34, 128, 4, 0, "depth 4 (cols 128: timing only), no fence", 1},
	    {34, 112, 1, 0, "depth 4, ONE MMA, no fence", 1}, {34, 112, 8, 0, "depth 4, EIGHT MMAs, no fence", 1}, {34, 112, 8, 0, "depth 4, EIGHT MMAs, fence", 0},
	    {61, 112, 4, 0, "issuer polls a shared-memory counter (relay warp waits), depth 1"}, {62, 112, 4, 0, "shared-memory counter, depth 2"}, {63, 112, 4, 0, "shared-memory counter, depth 3"},
	    {62, 112, 1, 0, "shared-memory counter, depth 2, ONE MMA"}, {62, 112, 4, 12, "shared-memory counter, depth 2, 12 warps polling"},
	    {51, 112, 4, 0, "no wait, 200-cycle clock spin per group"}, {52, 112, 4, 0, "MMAs + commit, then wait for the commit of 3 groups ago"},
	    {52, 112, 1, 0, "ONE MMA + commit, then wait for the commit of 3 groups ago"},
	    {34, 112, 4, 0, "free-running groups of four 128x112x16 MMAs, BF16 in / F32 accumulator", 2}, {34, 112, 4, 0, "free-running groups of four 128x112x16 MMAs, F16 in / F16 accumulator", 2 + 16},
	    {34, 128, 4, 0, "free-running groups of four 128x128x16 MMAs (timing only), BF16 / F32", 2}, {34, 64, 4, 0, "free-running groups of four 128x64x16 MMAs, BF16 / F32", 2},
	    {34, 112, 1, 0, "free-running single 128x112x16 MMAs into four accumulators in turn, BF16 / F32", 2}, {34, 112, 1, 0, "the same, F16 / F16", 2 + 16},
	    {31, 112, 4, 0, "self-paced, depth 1"}, {32, 112, 4, 0, "self-paced, depth 2"}, {33, 112, 4, 0, "self-paced, depth 3"}, {34, 112, 4, 0, "self-paced, depth 4"},
	    {41, 112, 4, 0, "relayed by one warp, depth 1"}, {42, 112, 4, 0, "relayed by one warp, depth 2"}, {43, 112, 4, 0, "relayed, depth 3"}, {44, 112, 4, 0, "relayed, depth 4"},
	    {42, 112, 4, 12, "relayed, depth 2, 12 warps polling"}, {42, 32, 1, 0, "relayed, depth 2, one 128x32x16 MMA"},
	    {2, 0, 0, 0, "arrive/wait ping-pong between two warps (two hops)"}, {2, 0, 0, 8, "ping-pong, 8 warps polling"}, {2, 0, 0, 15, "ping-pong, 15 warps polling"}};
	for (auto& c : ms) {
		CK(cudaMemset(d_cycles, 0, 148 * 16));
		micro_kernel<<<148, 608, smem_bytes>>>(c.which, c.cols, c.nmma, 20000, c.pollers, d_cycles, c.mflags);
		CK(cudaDeviceSynchronize());
		long long h; CK(cudaMemcpy(&h, d_cycles, 8, cudaMemcpyDeviceToHost));
		printf("%-70s %.1f cycles per iteration\n", c.name, (double)h / 20000);
	}
	uint32_t* d_sink; CK(cudaMalloc(&d_sink, 4));
	for (int nw : {1, 4, 8, 16})
		for (int cols : {32, 64, 128}) {
			CK(cudaMemset(d_cycles, 0, 148 * 16));
			ldtm_kernel<<<148, 512>>>(nw, cols, 20000, d_cycles, d_sink);
			CK(cudaDeviceSynchronize());
			long long h; CK(cudaMemcpy(&h, d_cycles, 8, cudaMemcpyDeviceToHost));
			printf("LDTM: %2d warps x %3d columns per round: %.1f cycles per round, %.1f bytes per cycle per SM\n", nw, cols, (double)h / 20000,
			       (double)nw * 32 * cols * 4 * 20000 / (double)h);
		}
	return 0;
}
