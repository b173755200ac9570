// bs_assoc_tc.cu — tensor-core candidate filter for greedy association (SURVEY.md §8a rows 14-15 at 100k x 100k scale).
//
// Reference semantics (unchanged): AssociateGreedy.associate (F/alg/feature/associate/AssociateGreedy.java:65-118) with
// DescriptorDistance.euclideanSq (F/alg/descriptor/DescriptorDistance.java:55-64): per source row the destination with the
// smallest fp64 squared distance (later index wins ties), optional backward validation by the strict column minimum.
//
// The reference evaluates all N*M distances in fp64.  Here the N*M work runs on the 5th-generation tensor cores and only
// a handful of pairs per row is re-scored in fp64, in the reference's summation order, so the result is bit-identical:
//   1. split_kernel: every descriptor x is split into bf16 hi + bf16 lo (x = hi + lo + r, |r| <= 2^-18 |x|) and stored as a
//      K = 192 row: P = [hi | hi | lo] or Q = [hi | lo | hi], so that P(a) . Q(b) = hi.hi + hi.lo + lo.hi.
//   2. filter_kernel (tcgen05): s = P(a).Q(b) for 256 rows x 128 columns per step, fp32 accumulators in TMEM;
//      d' = |b|^2 - 2 s.  Each row keeps the running minimum m of d' and emits column j as a CANDIDATE whenever
//      d'_j <= m + tol.  tol = 2 * (bound on |d' - exact|), so the exact minimum — and every exact tie — is always among
//      the candidates (the bound and its margin are stated at tolerance_of()).
//   3. rescore_kernel: exact fp64 distance of every candidate, sequential over k without FMA like the reference; minimum,
//      first and last index of the minimum -> pairs/fit (forward) or the column statistics used by backward validation.
// Roles inside the filter CTA (320 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one thread), warps 2-9 =
// epilogue (one TMEM lane = one row per thread, one warp set per resident A tile).  A (2 x 128 rows) stays resident per work item, B tiles stream through a
// 2-stage ring; two accumulator buffers in TMEM let the epilogue of tile t overlap the MMAs of tile t+1.
#include "bs_common.cuh"
#include <cstdlib>
#include "bs_sm100.cuh"

#include <cuda_bf16.h>
#include <math_constants.h>
#include <cfloat>
#include <cmath>

namespace bs {

using namespace sm100;

constexpr int TC_DOF = 64;                 // descriptor length this path is built for (SURF-64)
constexpr int TC_K = 3 * TC_DOF;           // 192 bf16 per operand row
constexpr int TC_ATOMS = TC_K / 64;        // 128-byte swizzle atoms along K
constexpr int TC_BM = 128, TC_BN = 128;    // one UMMA: 128 x 128 x 16
constexpr int TC_MROWS = 2 * TC_BM;        // rows per work item (two resident A tiles)
constexpr int TC_STAGES = 2;
constexpr int TC_ATOM_BYTES = TC_BM * 128; // 16 KB: 128 rows x 128 B
constexpr int TC_TILE_BYTES = TC_ATOMS * TC_ATOM_BYTES;   // 48 KB
constexpr int TC_EPI_WARPS = 8;           // two per TMEM lane quarter: one per resident A tile
constexpr int TC_THREADS = 64 + 32 * TC_EPI_WARPS;
constexpr int TC_CAND_CAP = 64;            // candidates kept per row; more -> the caller falls back to the exact kernel
constexpr int TC_TMEM_COLS = 512;

// ---------------------------------------------------------------- 1. split
// arrangement 0: P = [hi | hi | lo], 1: Q = [hi | lo | hi].  Rows >= n (padding) are zero with norm = +inf.
__global__ void split_kernel(const double* __restrict__ x, int n, int n_pad, int arrangement, __nv_bfloat16* __restrict__ out,
                             float* __restrict__ norm, unsigned* __restrict__ max_norm_bits) {
	const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (row >= n_pad) return;
	__nv_bfloat16* o = out + (size_t)row * TC_K;
	double sq = 0.0;
	for (int k = lane; k < TC_DOF; k += 32) {
		const double v = row < n ? x[(size_t)row * TC_DOF + k] : 0.0;
		const __nv_bfloat16 hi = __double2bfloat16(v);
		const __nv_bfloat16 lo = __double2bfloat16(v - (double)__bfloat162float(hi));
		o[k] = hi;
		o[TC_DOF + k] = arrangement == 0 ? hi : lo;
		o[2 * TC_DOF + k] = arrangement == 0 ? lo : hi;
		sq += v * v;
	}
	for (int d = 16; d > 0; d >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, d);
	if (lane == 0) {
		const float f = row < n ? (float)sq : INFINITY;
		norm[row] = f;
		if (row < n) atomicMax(max_norm_bits, __float_as_uint(f));      // non-negative floats order like their bit patterns
	}
}

// Bound on |d'_computed - d'_exact| for one pair, d' = |b|^2 - 2 a.b:
//   split: dropped terms lo.lo, r_a.b, a.r_b               <= 3 * 2^-18 |a||b|      = 1.2e-5 |a||b|
//   fp32 accumulation of 192 exact bf16 x bf16 products    <= 192 * 2^-23 * 3|a||b| = 6.9e-5 |a||b|   (truncating adder assumed)
//   => |s - a.b| <= 8.1e-5 |a||b|; budgeted 3e-4 (3.7x margin) ; d' doubles it: 6e-4 |a||b|
//   fp32 rounding of |b|^2 and of the final fma            <= 2.4e-7 (|a|^2 + |b|^2); budgeted 1e-6
// tol = 2 * bound (tests/test_gpu_assoc_tc.py measures the real error and asserts it stays below a quarter of the budget)
__device__ __forceinline__ float tolerance_of(float na, float nb_max) {
	return 1.2e-3f * sqrtf(na * nb_max) + 2e-6f * (na + nb_max);
}

// ---------------------------------------------------------------- 2. filter
struct FilterArgs {
	int nA, nB;                 // real rows / columns
	int a_tiles256;             // ceil(nA / 256)
	int b_tiles;                // ceil(nB / 128)
	int chunks, tiles_per_chunk;
	const float* norm_a;        // [nA_pad]
	const float* norm_b;        // [nB_pad] (+inf in the padding)
	const unsigned* max_norm_b_bits;
	int* cand_count;            // [nA]
	int* cand;                  // [nA][TC_CAND_CAP]
	float* dbg;                 // optional [nA][nB] dump of d' + |a|^2 (tests)
	const double* row_bound;    // optional [nA]: an exact |a-b_j|^2 of some column j per row (an upper bound of the row's minimum), else huge
};

struct __align__(8) FilterBarriers {
	uint64_t a_full, a_empty;
	uint64_t b_full[TC_STAGES], b_empty[TC_STAGES];
	uint64_t acc_full[2], acc_empty[2];
	uint32_t tmem_base;
	uint32_t pad;
};

constexpr int TC_PEND = 8;                // candidates an epilogue thread collects before it reserves slots in its row's list
constexpr int TC_SMEM_BYTES = 1024 /*alignment slack*/ + 2 * TC_TILE_BYTES + TC_STAGES * TC_TILE_BYTES + 2 * TC_BN * 4 + (int)sizeof(FilterBarriers) +
                              TC_PEND * 32 * TC_EPI_WARPS * 4;

// Row items that run side by side would all stream the same B tiles at the same moment (every CTA starts at the first tile of its
// chunk): each row item walks its chunk from a different starting tile instead.  The running minimum and the candidate lists do not
// depend on the order of the columns.
__device__ __forceinline__ int rotated_tile(int t, int t0, int t1, int rp) {
#ifdef BS_TC_NO_ROTATE
	return t;
#else
	const int len = t1 - t0;
	int tt = t + (int)(((unsigned)rp * 11u) % (unsigned)len);
	return tt >= t1 ? tt - len : tt;
#endif
}

__global__ void __launch_bounds__(TC_THREADS, 1)
filter_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, FilterArgs P) {
	extern __shared__ unsigned char smem_raw[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
	unsigned char* sA = smem;                                        // [2][3 atoms][128 rows][128 B]
	unsigned char* sB = sA + 2 * TC_TILE_BYTES;                      // [stages][3 atoms][128 rows][128 B]
	float* s_nb = reinterpret_cast<float*>(sB + TC_STAGES * TC_TILE_BYTES);   // [2][128]
	FilterBarriers* bar = reinterpret_cast<FilterBarriers*>(s_nb + 2 * TC_BN);
	int* s_pend = reinterpret_cast<int*>(bar + 1);                   // [TC_PEND][256]: pending candidates, one column per epilogue thread

	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0) {
		mbar_init(&bar->a_full, 1); mbar_init(&bar->a_empty, 1);
		for (int s = 0; s < TC_STAGES; s++) { mbar_init(&bar->b_full[s], 1); mbar_init(&bar->b_empty[s], 1); }
		for (int s = 0; s < 2; s++) { mbar_init(&bar->acc_full[s], 1); mbar_init(&bar->acc_empty[s], TC_EPI_WARPS); }
		mbar_fence_init();
	}
	__syncwarp();
	if (warp == 0) tmem_alloc<TC_TMEM_COLS>(&bar->tmem_base);
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = bar->tmem_base;

	const int n_items = P.a_tiles256 * P.chunks;
	if (warp == 0) {
		// ===== TMA producer =====
		if (lane == 0) {
			tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB);
			uint32_t n_item = 0, n_tile = 0;
			for (int item = blockIdx.x; item < n_items; item += gridDim.x, n_item++) {
				const int rp = item / P.chunks, ch = item - rp * P.chunks;
				const int t0 = ch * P.tiles_per_chunk, t1 = min(t0 + P.tiles_per_chunk, P.b_tiles);
				mbar_wait(&bar->a_empty, (n_item & 1) ^ 1);          // MMAs of the previous item are done with A
				mbar_expect_tx(&bar->a_full, 2 * TC_TILE_BYTES);
				for (int h = 0; h < 2; h++)
					for (int a = 0; a < TC_ATOMS; a++)
						tma_load_2d(sA + (h * TC_ATOMS + a) * TC_ATOM_BYTES, &tmA, a * 64, rp * TC_MROWS + h * TC_BM, &bar->a_full);
				for (int t = t0; t < t1; t++, n_tile++) {
					const int s = n_tile % TC_STAGES;
					const uint32_t use = n_tile / TC_STAGES;
					mbar_wait(&bar->b_empty[s], (use & 1) ^ 1);
					mbar_expect_tx(&bar->b_full[s], TC_TILE_BYTES);
					for (int a = 0; a < TC_ATOMS; a++)
						tma_load_2d(sB + (s * TC_ATOMS + a) * TC_ATOM_BYTES, &tmB, a * 64, rotated_tile(t, t0, t1, rp) * TC_BN, &bar->b_full[s]);
				}
			}
		}
	} else if (warp == 1) {
		// ===== MMA issuer =====
		if (lane == 0) {
			constexpr uint32_t idesc = umma_idesc_bf16(TC_BM, TC_BN);
			uint32_t n_item = 0, n_tile = 0;
			for (int item = blockIdx.x; item < n_items; item += gridDim.x, n_item++) {
				const int rp = item / P.chunks, ch = item - rp * P.chunks;
				const int t0 = ch * P.tiles_per_chunk, t1 = min(t0 + P.tiles_per_chunk, P.b_tiles);
				mbar_wait(&bar->a_full, n_item & 1);
				for (int t = t0; t < t1; t++, n_tile++) {
					const int s = n_tile % TC_STAGES;
					const uint32_t use = n_tile / TC_STAGES;
					const int as = n_tile & 1;
					const uint32_t acc_use = n_tile >> 1;
					mbar_wait(&bar->b_full[s], use & 1);
					mbar_wait(&bar->acc_empty[as], (acc_use & 1) ^ 1);   // epilogue drained this accumulator buffer
					tc_fence_after();
					for (int h = 0; h < 2; h++) {
						const uint32_t d = tmem + (uint32_t)(as * 2 * TC_BN + h * TC_BN);
						for (int a = 0; a < TC_ATOMS; a++) {
							const uint64_t da = umma_desc_sw128(smem_u32(sA + (h * TC_ATOMS + a) * TC_ATOM_BYTES));
							const uint64_t db = umma_desc_sw128(smem_u32(sB + (s * TC_ATOMS + a) * TC_ATOM_BYTES));
#pragma unroll
							for (int kk = 0; kk < 4; kk++)                   // 16 bf16 = 32 bytes = 2 descriptor units per step
								umma_bf16(d, da + 2 * kk, db + 2 * kk, idesc, (a | kk) != 0 ? 1u : 0u);
						}
					}
					umma_commit(&bar->b_empty[s]);                       // smem stage free once these MMAs have read it
					umma_commit(&bar->acc_full[as]);                     // accumulators complete
				}
				umma_commit(&bar->a_empty);
			}
		}
	} else {
		// ===== epilogue: warps 2..9.  TMEM lane quarter = warp % 4 (hardware rule); warps 2-5 take A tile 0, 6-9 A tile 1 =====
		const int q = warp & 3;
		const int h = (warp - 2) >> 2;
		const int e = (warp - 2) * 32 + lane;                        // 0..255 among the epilogue threads
		const int L = q * 32 + lane;                                 // TMEM lane = row inside the 128-row tile
		const float nb_max = __uint_as_float(*P.max_norm_b_bits);
		uint32_t n_tile = 0;
		for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
			const int rp = item / P.chunks, ch = item - rp * P.chunks;
			const int t0 = ch * P.tiles_per_chunk, t1 = min(t0 + P.tiles_per_chunk, P.b_tiles);
			const int row = rp * TC_MROWS + h * TC_BM + L;
			const bool live = row < P.nA;
			const float na = live ? P.norm_a[row] : 0.0f;
			const float tol = tolerance_of(na, nb_max);
			// A row's candidates are the columns within tol of its running minimum; while that minimum is loose nearly every
			// 32-column group has SOME lane of the warp with a new minimum and the whole warp pays the candidate path (a third of
			// the filter at 10k x 10k).  Where the caller knows a distance the row really attains — the second direction takes
			// it from the first one's matches — the minimum starts there: the true minimum and its ties are still within tol
			// (the estimate of a column at most bound = tol/2 above the exact value lies below start + tol; a quarter of tol is
			// added for the float roundings of the start value itself).
			float run = INFINITY;
			if (P.row_bound && live) {
				const double ub = P.row_bound[row];
				if (ub < 1e30) run = __double2float_ru(ub) - na + 0.25f * tol;
			}
			// Candidates are collected per thread (a row belongs to one thread inside an item) and appended to the row's list
			// TC_PEND at a time: ONE atomicAdd reserves the slots, the stores do not depend on each other.  ncu had 21 % of the
			// stall samples of the 10k x 10k filter on the atomic-then-store chain of one candidate at a time.
			int n_pend = 0;
			auto flush = [&]() {
				if (n_pend == 0) return;
				const int base = atomicAdd(P.cand_count + row, n_pend);
#pragma unroll
				for (int i = 0; i < TC_PEND; i++)
					if (i < n_pend && base + i < TC_CAND_CAP) P.cand[(size_t)row * TC_CAND_CAP + base + i] = s_pend[i * (32 * TC_EPI_WARPS) + e];
				n_pend = 0;
			};
			for (int t = t0; t < t1; t++, n_tile++) {
				const int as = n_tile & 1;
				const uint32_t acc_use = n_tile >> 1;
				const int col0 = rotated_tile(t, t0, t1, rp) * TC_BN;
				if (e < TC_BN) s_nb[as * TC_BN + e] = P.norm_b[col0 + e];
				asm volatile("bar.sync 1, 256;" ::: "memory");        // the 256 epilogue threads only
				mbar_wait(&bar->acc_full[as], acc_use & 1);
				tc_fence_after();
				const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * 2 * TC_BN + h * TC_BN);
				uint32_t ra[32], rb[32];
				// chunk c+1 is in flight while chunk c is reduced
				auto reduce = [&](uint32_t (&r)[32], int c) {
					const float* nb = s_nb + as * TC_BN + c * 32;
					float v[32];
					float cmin = INFINITY;
#pragma unroll
					for (int k = 0; k < 32; k++) {
						v[k] = fmaf(-2.0f, __uint_as_float(r[k]), nb[k]);
						cmin = fminf(cmin, v[k]);
					}
					if (P.dbg && live) {
#pragma unroll
						for (int k = 0; k < 32; k++)
							if (col0 + c * 32 + k < P.nB) P.dbg[(size_t)row * P.nB + col0 + c * 32 + k] = v[k] + na;
					}
					if (cmin <= run + tol && cmin < INFINITY && live) {
						run = fminf(run, cmin);
						const float thr = run + tol;
						unsigned hits = 0;
#pragma unroll
						for (int k = 0; k < 32; k++) hits |= (v[k] <= thr ? 1u : 0u) << k;
						while (hits) {
							const int k = __ffs(hits) - 1;
							hits &= hits - 1;
							s_pend[n_pend * (32 * TC_EPI_WARPS) + e] = col0 + c * 32 + k;
							if (++n_pend == TC_PEND) flush();
						}
					}
				};
				tmem_ld32_issue(tbase, ra);
				tmem_ld_wait(ra); tmem_ld32_issue(tbase + 32, rb); reduce(ra, 0);
				tmem_ld_wait(rb); tmem_ld32_issue(tbase + 64, ra); reduce(rb, 1);
				tmem_ld_wait(ra); tmem_ld32_issue(tbase + 96, rb); reduce(ra, 2);
				tmem_ld_wait(rb);
				tc_fence_before();
				__syncwarp();
				if (lane == 0) mbar_arrive(&bar->acc_empty[as]);     // TMEM buffer free: the last chunk lives in registers now
				reduce(rb, 3);
			}
			flush();
		}
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<TC_TMEM_COLS>(tmem); }
}

// ---------------------------------------------------------------- 3. exact re-score
// one warp per row of A: exact distance to every candidate column of B; mode 0 -> pairs/fit, mode 1 -> column statistics
__global__ void rescore_kernel(const double* __restrict__ A, int nA, const double* __restrict__ B, const int* __restrict__ cand_count,
                               const int* __restrict__ cand, int mode, double max_fit, int index_offset, int32_t* __restrict__ pairs,
                               double* __restrict__ fit, double* __restrict__ o_min, int32_t* __restrict__ o_first,
                               int32_t* __restrict__ o_last, int* __restrict__ overflow) {
	const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (row >= nA) return;
	int cnt = cand_count[row];
	if (cnt > TC_CAND_CAP) { if (lane == 0) atomicExch(overflow, 1); cnt = TC_CAND_CAP; }
	const double* a = A + (size_t)row * TC_DOF;
	double best = CUDART_INF;
	int first = -1, last = -1;
	for (int c = lane; c < cnt; c += 32) {
		const int j = cand[(size_t)row * TC_CAND_CAP + c];
		const double* b = B + (size_t)j * TC_DOF;
		double s = 0.0;
		// DescriptorDistance.euclideanSq :55-64, same order, no FMA.  The candidate's row is fetched 16 values at a time
		// (eight 16-byte loads in flight per lane) ahead of the serial sum; the a row is a broadcast load
		const double2* b2 = reinterpret_cast<const double2*>(b);
		const double2* a2 = reinterpret_cast<const double2*>(a);
#pragma unroll 1
		for (int k0 = 0; k0 < TC_DOF / 2; k0 += 8) {
			double2 bv[8], av[8];
#pragma unroll
			for (int u = 0; u < 8; u++) { bv[u] = __ldg(b2 + k0 + u); av[u] = __ldg(a2 + k0 + u); }
#pragma unroll
			for (int u = 0; u < 8; u++) {
				const double d0 = av[u].x - bv[u].x;
				s += d0 * d0;
				const double d1 = av[u].y - bv[u].y;
				s += d1 * d1;
			}
		}
		if (first < 0 || s < best) { best = s; first = j; last = j; }
		else if (s == best) { first = min(first, j); last = max(last, j); }
	}
	for (int d = 16; d > 0; d >>= 1) {
		const double ob = __shfl_xor_sync(0xffffffffu, best, d);
		const int of = __shfl_xor_sync(0xffffffffu, first, d), ol = __shfl_xor_sync(0xffffffffu, last, d);
		if (of >= 0) {
			if (first < 0 || ob < best) { best = ob; first = of; last = ol; }
			else if (ob == best) { first = min(first, of); last = max(last, ol); }
		}
	}
	if (lane != 0) return;
	if (mode == 0) {
		const bool hit = first >= 0 && best <= max_fit;      // AssociateGreedy.java:80-91
		pairs[row] = hit ? last : -1;
		fit[row] = hit ? best : max_fit;
	} else {
		o_min[row] = best;
		o_first[row] = first < 0 ? -1 : first + index_offset;
		o_last[row] = last < 0 ? -1 : last + index_offset;
	}
}

// ---------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
	static EncodeTiledFn fn = nullptr;
	if (!fn) {
		void* p = nullptr;
		cudaDriverEntryPointQueryResult qres;
		BS_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
		if (!p || qres != cudaDriverEntryPointSuccess) fail(BS_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
		fn = reinterpret_cast<EncodeTiledFn>(p);
	}
	return fn;
}

// operand matrix [rows_pad][192] bf16, box = 64 elements (one 128-byte swizzle atom) x 128 rows
static CUtensorMap operand_map(const void* base, int rows_pad) {
	CUtensorMap m;
	const cuuint64_t dims[2] = {(cuuint64_t)TC_K, (cuuint64_t)rows_pad};
	const cuuint64_t strides[1] = {(cuuint64_t)TC_K * 2};
	const cuuint32_t box[2] = {64, (cuuint32_t)TC_BM};
	const cuuint32_t estr[2] = {1, 1};
	const CUresult r = encode_tiled_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
	                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
	                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	if (r != CUDA_SUCCESS) fail(BS_ERR_CUDA, "cuTensorMapEncodeTiled failed");
	return m;
}

int tc_pad_rows(int n) { return ((n + TC_MROWS - 1) / TC_MROWS) * TC_MROWS; }
size_t tc_operand_bytes(int n) { return (size_t)tc_pad_rows(n) * TC_K * 2; }
int tc_cand_cap() { return TC_CAND_CAP; }
bool tc_supported(int dof) { return dof == TC_DOF; }

void launch_tc_split(const double* d_x, int n, int arrangement, void* d_operand, float* d_norm, unsigned* d_max_norm_bits, cudaStream_t st) {
	const int n_pad = tc_pad_rows(n);
	BS_CUDA(cudaMemsetAsync(d_max_norm_bits, 0, 4, st));
	BS_LAUNCH(split_kernel, div_up(n_pad, 8), 256, 0, st, d_x, n, n_pad, arrangement, static_cast<__nv_bfloat16*>(d_operand), d_norm,
	          d_max_norm_bits);
}

// rows of A against all rows of B: candidate lists for the rows of A
void launch_tc_filter(const void* d_opA, const float* d_normA, int nA, const void* d_opB, const float* d_normB,
                      const unsigned* d_max_norm_b_bits, int nB, int* d_cand_count, int* d_cand, float* d_dbg, int sm_count,
                      cudaStream_t st, const double* d_row_bound) {
	if (nA <= 0 || nB <= 0) return;
	static bool attr_set = false;
	if (!attr_set) {
		BS_CUDA(cudaFuncSetAttribute(filter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
		attr_set = true;
	}
	FilterArgs P{};
	P.nA = nA; P.nB = nB;
	P.a_tiles256 = tc_pad_rows(nA) / TC_MROWS;
	P.b_tiles = tc_pad_rows(nB) / TC_BN;
	// column chunks only when there are fewer 256-row items than SMs: a row's running minimum (and with it the number of
	// candidates it emits, ~ln(columns/32)) is per item; at least 8 B tiles per item to amortise loading A
	// — and as many chunks as keep all items in ONE wave of CTAs (a second, mostly empty wave costs a whole item time)
	int chunks = sm_count / std::max(P.a_tiles256, 1);
	chunks = std::max(1, std::min(chunks, P.b_tiles / 8 > 0 ? P.b_tiles / 8 : 1));
	static const int env_chunks = getenv("BS_TC_CHUNKS") ? atoi(getenv("BS_TC_CHUNKS")) : 0;     // experiments
	if (env_chunks > 0) chunks = std::min(env_chunks, std::max(P.b_tiles, 1));
	P.tiles_per_chunk = div_up(P.b_tiles, chunks);
	P.chunks = div_up(P.b_tiles, P.tiles_per_chunk);
	P.norm_a = d_normA; P.norm_b = d_normB; P.max_norm_b_bits = d_max_norm_b_bits;
	P.cand_count = d_cand_count; P.cand = d_cand; P.dbg = d_dbg;
	static const bool no_bound = getenv("BS_TC_NO_BOUND") != nullptr;      // experiments
	P.row_bound = no_bound ? nullptr : d_row_bound;
	const CUtensorMap tmA = operand_map(d_opA, tc_pad_rows(nA));
	const CUtensorMap tmB = operand_map(d_opB, tc_pad_rows(nB));
	BS_CUDA(cudaMemsetAsync(d_cand_count, 0, (size_t)nA * 4, st));
	const int grid = std::min(P.a_tiles256 * P.chunks, sm_count);
	BS_LAUNCH(filter_kernel, grid, TC_THREADS, TC_SMEM_BYTES, st, tmA, tmB, P);
}

// bound[j] = min over the rows i that chose column j of fit[i] (exact distances; non-negative doubles order like their bit patterns);
// columns nobody chose keep the fill value 0x7f7f... (huge).  The caller fills `bound` with cudaMemsetAsync(0x7f) first.
__global__ void row_bounds_kernel(const int32_t* __restrict__ pairs, const double* __restrict__ fit, int n, int m, unsigned long long* __restrict__ bound) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int j = pairs[i];
	const double f = fit[i];
	if (j >= 0 && j < m && f >= 0.0) atomicMin(bound + j, (unsigned long long)__double_as_longlong(f));
}
void launch_tc_row_bounds(const int32_t* d_pairs, const double* d_fit, int n, int m, double* d_bound, cudaStream_t st) {
	if (m <= 0) return;
	BS_CUDA(cudaMemsetAsync(d_bound, 0x7f, (size_t)m * 8, st));
	if (n <= 0) return;
	BS_LAUNCH(row_bounds_kernel, div_up(n, 256), 256, 0, st, d_pairs, d_fit, n, m, reinterpret_cast<unsigned long long*>(d_bound));
}

void launch_tc_rescore(const double* d_A, int nA, const double* d_B, const int* d_cand_count, const int* d_cand, int mode, double max_fit,
                       int index_offset, int32_t* pairs, double* fit, double* o_min, int32_t* o_first, int32_t* o_last, int* d_overflow,
                       cudaStream_t st) {
	if (nA <= 0) return;
	BS_LAUNCH(rescore_kernel, div_up(nA, 8), 256, 0, st, d_A, nA, d_B, d_cand_count, d_cand, mode, max_fit, index_offset, pairs, fit, o_min,
	          o_first, o_last, d_overflow);
}

}  // namespace bs
