// bs_assoc_tc.cu — tensor-core candidate filter for greedy association (SURVEY.md §8a rows 14-15 at 100k x 100k scale).
//
// Reference semantics (unchanged): AssociateGreedy.associate (F/alg/feature/associate/AssociateGreedy.java:65-118) with
// DescriptorDistance.euclideanSq (F/alg/descriptor/DescriptorDistance.java:55-64): per source row the destination with the
// smallest fp64 squared distance (later index wins ties), optional backward validation by the strict column minimum.
//
// The reference evaluates all N*M distances in fp64.  Here the N*M work runs on the 5th-generation tensor cores and only
// a handful of pairs per row is re-scored in fp64, in the reference's summation order, so the result is bit-identical:
//   1. split_kernel: every descriptor is scaled by a power of two to a norm in [1, 2) and rounded to fp16 (one K = 64 row = one 128-byte
//      swizzle atom); its squared norm and its scale are kept in fp32.
//   2. filter_kernel (tcgen05, kind::f16 with F16 inputs AND an F16 accumulator): s = a~ . b~ for 128 rows x 112 columns per accumulator;
//      d' = |b|^2 - 2 * 2^(ea+eb) s.  Each row keeps the running minimum m of d' and emits column j as a CANDIDATE whenever d'_j <= m + tol.
//      tol = 2 * (bound on |d' - exact|), so the exact minimum — and every exact tie — is always among the candidates
//      (the bound is stated at tolerance_of()).
//   3. rescore_kernel: exact fp64 distance of every candidate, sequential over k without FMA like the reference; minimum,
//      first and last index of the minimum -> pairs/fit (forward) or the column statistics used by backward validation.
// Why fp16 all the way (tools/probe/f16acc_probe.cu, alu_probe.cu, hs2_probe.cu): what bounds this kernel is not the tensor pipe
// but getting the accumulators out of TMEM and through a row maximum.  An F16 accumulator keeps one element per 32-bit TMEM column like an
// fp32 one, but tcgen05.ld .pack::16b delivers two per register: a thread holds a whole 112-column accumulator row in 56 registers (one read
// round, the slot goes back at once) and the packed maximum reduces two elements per issue slot.  The accumulator rounds to nearest after
// each of the four 16-term MMAs; with rows normalised to [1, 2) the estimate is MORE accurate than bf16 inputs with an fp32 accumulator
// (measured worst 7.0e-4 |a||b| against 1.8e-3).
// Shape: the 256 rows of a row block live in TMEM as the A operand (written there by the epilogue threads with tcgen05.st: 73 cycles per
// UMMA against 97 with A in shared memory, tools/probe/tc_probe.cu), B tiles stream through an 8-stage TMA ring together with their records
// (norms, flag, scales), four 112-column accumulators: slot = 2 * (A half) + (tile parity).
// Roles inside the CTA (608 threads): warp 0 = TMA producer, warps 1 and 2 = MMA issuers (one elected thread each, one per A half), warps
// 3-18 = epilogue in four groups of four warps (a group covers the four TMEM lane quarters): group (A half, tile parity) owns one slot and
// takes every other tile of its half — one TMEM lane = one row per thread, two threads per row (they exchange their running minima through
// shared memory).  The work of a launch is the sequence of (row block, B tile) pairs in row-block-major order, cut into one contiguous
// range per CTA.
#include "bs_common.cuh"
#include <cstdlib>
#include "bs_sm100.cuh"

#include <cuda_fp16.h>
#include <math_constants.h>
#include <cfloat>
#include <cmath>
#include <type_traits>

namespace bs {

using namespace sm100;

constexpr int TC_DOF = 64;                 // descriptor length this path is built for (SURF-64)
constexpr int TC_K = TC_DOF;               // fp16 values per operand row: one 128-byte swizzle atom
constexpr int TC_BM = 128, TC_BN = 112;    // one accumulator: 128 rows x 112 columns, four 128x112x16 UMMAs (four of them + A = 512 TMEM columns)
constexpr int TC_MROWS = 2 * TC_BM;        // rows per row block (two A halves resident in TMEM)
constexpr int TC_STAGES = 8;               // B ring
constexpr int TC_TILE_BYTES = TC_BN * 128; // 14 KB: 112 rows x 128 B (a multiple of the 1024-byte swizzle atom)
constexpr int TC_NB = 2 * TC_BN + 4;       // the record of a 112-row tile: 112 squared norms, {all equal?, 0, 0, 0}, 112 scales (-2 * 2^e); see norm_index()
constexpr int TC_SC = TC_BN + 4;           // float offset of the scales inside a tile's record
constexpr int TC_SLOTS = 4;                // accumulators in flight: slot = 2 * (A half) + (tile parity)
constexpr int TC_A_COL = TC_SLOTS * TC_BN; // TMEM columns [448, 512): the two A halves, 32 columns each
constexpr int TC_EPI_WARPS = 16;           // four groups: (A half) x (tile parity), each the only reader of one slot
constexpr int TC_ISSUER2 = 2;                 // warp index of the second MMA issuer.  The role warps take the LOWEST warp indices: with them at the
                                              // highest (producer 16, issuers 17 / 18) the kernel was 10 % slower (2.72 against 2.45 ms per 100k x 100k)
constexpr int TC_THREADS = 96 + 32 * TC_EPI_WARPS;
constexpr int TC_CAND_CAP = 128;           // candidates kept per row; more -> the caller falls back to the exact kernel.  Real SURF
                                           // descriptors of a 1080p frame (10k x 10k, ~8 running-minimum chains per row: 4 CTAs x 2
                                           // column halves) average 44 per row and reach 90: tools/dbg_stereo.py
constexpr int TC_TMEM_COLS = 512;

// Norms are stored tile by tile, 228 floats per 112 rows: the 112 squared norms, a flag (non-zero: the 112 are one and the same finite
// float) and three zeros, then the 112 scales -2 * 2^e of the rows (see split_kernel), so that one bulk copy brings all a tile's columns
// need into shared memory.
__host__ __device__ __forceinline__ size_t norm_index(int row) { return (size_t)row + (size_t)(row / TC_BN) * (TC_NB - TC_BN); }
static_assert(TC_BN % 16 == 0 && TC_TILE_BYTES % 1024 == 0 && (TC_NB * 4) % 16 == 0 && TC_A_COL + TC_K <= 512, "tile shape");

// ---------------------------------------------------------------- 1. split
// Every row is scaled by a power of two, 2^-e with 4^e <= |x|^2 < 4^(e+1) (of the norm as rounded to float), before it is rounded to fp16: the operand rows have norms in
// [1, 2) whatever the descriptors' magnitude (fp16 has 5 exponent bits), the scaling is exact, and the product of two rows is
// a.b = 2^(ea+eb) * (a' . b').  The record keeps |x|^2 and -2 * 2^e per row.  Rows >= n are zero (the operand is padded to a multiple of
// 256 rows) with norm = +inf (the records to whole 112-row tiles).
__global__ void split_kernel(const double* __restrict__ x, int n, int n_pad, int n_norm, __half* __restrict__ out, float* __restrict__ norm) {
	const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (row >= n_norm) return;
	double2 v = make_double2(0.0, 0.0);
	if (row < n) v = reinterpret_cast<const double2*>(x + (size_t)row * TC_DOF)[lane];
	double sq = v.x * v.x + v.y * v.y;
	for (int d = 16; d > 0; d >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, d);
	int e = 0;
	const float sqf = (float)sq;
	if (sqf > 0.0f && sqf < INFINITY) e = ilogbf(sqf) >> 1;              // floor(log4 |x|^2) of the ROUNDED norm: rows with equal float norms get equal
	                                                                     // scales (the fast path relies on it); the shift of a negative int is arithmetic
	// Norms beyond 1e-15 .. 1e15 (squared: 1e-30 .. 1e30) are outside what the estimate's error bound covers (the product of two scales must
	// stay a normal float, a scaled row must stay in fp16's range): such a row is given the norm +inf, which makes tol infinite, every column a
	// candidate of every row, the lists overflow and the caller redoes the whole problem with the exact kernel.  All-zero rows are exact.
	const bool covered = sqf == 0.0f || (sqf >= 1e-30f && sqf <= 1e30f);
	e = max(-60, min(60, e));
	if (row < n_pad) reinterpret_cast<__half2*>(out + (size_t)row * TC_K)[lane] = __halves2half2(__double2half(scalbn(v.x, -e)), __double2half(scalbn(v.y, -e)));
	if (lane == 0) {
		const size_t i = norm_index(row);
		norm[i] = row < n && covered ? sqf : INFINITY;
		norm[i + TC_SC] = -2.0f * scalbnf(1.0f, e);
	}
}
// one warp per 112-row tile (28 lanes x 4 norms): flag = the tile's norms are all the same finite value; the largest norm of the real rows (one atomic
// per tile: one per row in split_kernel was 100 k atomics on one address, 60 of that kernel's 70 us at 100 k rows)
__global__ void norm_flags_kernel(float* __restrict__ norm, int tiles, int n, unsigned* __restrict__ max_norm_bits) {
	const int tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int lane = threadIdx.x & 31;
	if (tile >= tiles) return;
	float* t = norm + (size_t)tile * TC_NB;
	const bool in = lane < TC_BN / 4;
	const float4 v = reinterpret_cast<const float4*>(t)[in ? lane : 0];
	const float first = __shfl_sync(0xffffffffu, v.x, 0);
	const bool same = v.x == first && v.y == first && v.z == first && v.w == first && first < INFINITY;
	const bool all = __all_sync(0xffffffffu, same);
	if (lane == 0) reinterpret_cast<float4*>(t)[TC_BN / 4] = make_float4(all ? 1.0f : 0.0f, 0.0f, 0.0f, 0.0f);
	const int r0 = tile * TC_BN + lane * 4;
	unsigned mx = 0;                                                 // non-negative floats order like their bit patterns
	if (in && r0 < n) mx = __float_as_uint(v.x);
	if (in && r0 + 1 < n) mx = max(mx, __float_as_uint(v.y));
	if (in && r0 + 2 < n) mx = max(mx, __float_as_uint(v.z));
	if (in && r0 + 3 < n) mx = max(mx, __float_as_uint(v.w));
	mx = __reduce_max_sync(0xffffffffu, mx);
	if (lane == 0 && mx) atomicMax(max_norm_bits, mx);
}

// Bound on |d'_computed - d'_exact| for one pair, d' = |b|^2 - 2 a.b, s = the tensor core's a' . b' of the scaled fp16 rows (F16 inputs,
// F16 accumulator: tools/probe/f16acc_probe.cu — the accumulator rounds to nearest after each of the four 16-term MMAs):
//   rounding to fp16 (11-bit significand, rows of norm 1..2: a component below the normal range costs at most 2^-25 absolute):
//       |a~.b~ - a'.b'| <= (2 * 2^-11 + 2^-22) sum |a'_k b'_k| + 2^-21 |a'||b'| <= 9.8e-4 |a'||b'|         (Cauchy-Schwarz)
//   four roundings of partial sums of magnitude <= |a~||b~|: <= 4 * 2^-11 (1 + 1e-3) |a'||b'| = 1.96e-3 |a'||b'|
//   fp32-or-better accumulation of 16 exact products inside an MMA: <= 16 * 2^-23 per MMA, 7.7e-6 |a'||b'| in all
//   => |s - a'.b'| <= 2.95e-3 |a'||b'|, budgeted 3.2e-3; undoing the (exact) scaling and doubling: 6.4e-3 |a||b|
//   fp32 rounding of |b|^2 and of the final fma: <= 2.4e-7 (|a|^2 + |b|^2); budgeted 1e-6
// tol = 2 * bound (tests/test_gpu_assoc_tc.py measures the real error and asserts it stays below three quarters of the budget).
// Non-finite values make the norms infinite, tol infinite, every column a candidate: the lists overflow and the caller redoes the problem
// with the exact kernel.
__device__ __forceinline__ float tolerance_of(float na, float nb_max) {
	return 1.28e-2f * sqrtf(na * nb_max) + 2e-6f * (na + nb_max);
}

// ---------------------------------------------------------------- 2. filter
struct FilterArgs {
	int nA, nB;                 // real rows / columns
	int a_blocks;               // ceil(nA / 256)
	int b_tiles;                // ceil(nB / 128)
	const __half* opA;          // [a_blocks * 256][64], rows scaled to norm 1..2
	const float* norm_a;        // [nA_pad]
	const float* norm_b;        // [nB_pad] (+inf in the padding)
	const unsigned* max_norm_b_bits;
	int* cand_count;            // [nA]
	int* cand;                  // [nA][TC_CAND_CAP][2]: column, estimate d' (float bits)
	float* dbg;                 // optional [nA][nB] dump of d' + |a|^2 (tests)
	const double* row_bound;    // optional [nA]: an exact |a-b_j|^2 of some column j per row (an upper bound of the row's minimum), else huge
};

struct __align__(8) FilterBarriers {
	uint64_t a_full;
	uint64_t b_full[TC_STAGES], b_empty[TC_STAGES];
	uint64_t acc_full[TC_SLOTS], acc_empty[TC_SLOTS];        // one issuer thread and one epilogue group per slot
	uint32_t tmem_base;
	uint32_t pad;
};

constexpr int TC_PEND = 8;                // candidates an epilogue thread collects before it reserves slots in its row's list
constexpr int TC_SMEM_BYTES = 1024 /*alignment slack*/ + TC_STAGES * TC_TILE_BYTES + TC_STAGES * TC_NB * 4 + (int)sizeof(FilterBarriers) +
                              2 * TC_PEND * 32 * TC_EPI_WARPS * 4 + TC_MROWS * 2 * 8;

#ifdef BS_TC_TRACE
// timing experiment (tools/trace_filter.py; build this file with -DBS_TC_TRACE): clock64 stamps of the hand-off events of tiles
// [TR_N0, TR_N0 + 32) of CTA 0.  Result of round 2 (profiles/r2d_filter_trace.txt): a tile takes ~800 cycles per A half; the thread that
// issues the MMAs needs ~360 cycles from the first MMA to the commit (its two barrier probes sit in there), the epilogue group sees the
// commit ~300 cycles later, reads the accumulator in ~135 and releases the slot 25 later — ~830 cycles after the tile's first MMA, long
// before the issuer comes back to that slot (~1600): the slots are not what it waits for, its own per-tile instruction stream is.
constexpr int TR_N0 = 200;
__device__ long long g_tc_trace[32 * 2 * 8];
#define BS_TR(nn, hh, k) do { if (blockIdx.x == 0 && (nn) >= TR_N0 && (nn) < TR_N0 + 32) g_tc_trace[(((nn) - TR_N0) * 2 + (hh)) * 8 + (k)] = clock64(); } while (0)
#else
#define BS_TR(nn, hh, k) do { } while (0)
#endif

template <bool DBG>
__global__ void __launch_bounds__(TC_THREADS, 1)
filter_kernel(const __grid_constant__ CUtensorMap tmB, FilterArgs P) {
	extern __shared__ unsigned char smem_raw[];
	unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
	unsigned char* sB = smem;                                                       // [stages][128 rows][128 B]
	float* s_nb = reinterpret_cast<float*>(sB + TC_STAGES * TC_TILE_BYTES);         // [stages][128]: column norms of the stage's tile
	FilterBarriers* bar = reinterpret_cast<FilterBarriers*>(s_nb + TC_STAGES * TC_NB);
	int* s_pend = reinterpret_cast<int*>(bar + 1);                   // [TC_PEND][512]: pending candidates, one column per epilogue thread
	unsigned long long* s_run = reinterpret_cast<unsigned long long*>(s_pend + 2 * TC_PEND * 32 * TC_EPI_WARPS);   // [256 rows][2]: {running minimum, row block}

	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0) {
		mbar_init(&bar->a_full, TC_EPI_WARPS / 2);                   // the warps that write A: column half 0 of both A halves
		for (int s = 0; s < TC_STAGES; s++) { mbar_init(&bar->b_full[s], 1); mbar_init(&bar->b_empty[s], TC_EPI_WARPS / 2); }
		for (int s = 0; s < TC_SLOTS; s++) { mbar_init(&bar->acc_full[s], 1); mbar_init(&bar->acc_empty[s], TC_EPI_WARPS / 4); }
		mbar_fence_init();
	}
	// no row block yet in the running-minimum exchange (shared memory keeps what the previous kernel left there)
	for (int i = threadIdx.x; i < TC_MROWS * 2; i += TC_THREADS) s_run[i] = ~0ull;
	__syncwarp();
	if (warp == 0) tmem_alloc<TC_TMEM_COLS>(&bar->tmem_base);
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = bar->tmem_base;

	// this CTA's share of the (row block, B tile) sequence
	const long long total = (long long)P.a_blocks * P.b_tiles;
	const long long g0 = total * blockIdx.x / gridDim.x, g1 = total * (blockIdx.x + 1) / gridDim.x;
	const int rp0 = (int)(g0 / P.b_tiles), t0 = (int)(g0 - (long long)rp0 * P.b_tiles);
	const int n_tiles = (int)(g1 - g0);

	if (warp == 0) {
		// ===== TMA producer: B tile + its 128 column norms per stage =====
		if (elect_one()) {
			tma_prefetch_desc(&tmB);
			int t = t0;
			for (int n = 0; n < n_tiles; n++) {
				const int s = n % TC_STAGES;
				const uint32_t use = n / TC_STAGES;
				mbar_wait(&bar->b_empty[s], (use & 1) ^ 1);          // the epilogue warps are done with the stage's norms (and the MMAs with its tile)
				mbar_expect_tx(&bar->b_full[s], TC_TILE_BYTES + TC_NB * 4);
				tma_load_2d(sB + s * TC_TILE_BYTES, &tmB, 0, t * TC_BN, &bar->b_full[s]);
				bulk_load_1d(s_nb + s * TC_NB, P.norm_b + (size_t)t * TC_NB, TC_NB * 4, &bar->b_full[s]);
				if (++t == P.b_tiles) t = 0;
			}
		}
	} else if (warp == 1 || warp == TC_ISSUER2) {
		// ===== MMA issuers: warp 1 takes A half 0, warp 2 A half 1 (one elected thread each), two accumulator slots per half.  An accumulator is four MMAs
		// (K = 64) between a wait and a commit: one thread's instruction stream for both halves (~45 dependent instructions per
		// accumulator, sharing its scheduler with four epilogue warps) took longer than the MMAs themselves (measured with
		// tools/probe/hs_probe.cu: 513 cycles per accumulator whatever the number of slots, 292 without the hand-off). =====
		// The tensor pipe takes MMAs from the issuing thread almost synchronously (tools/probe/hs2_probe.cu: a 200-cycle pause of the
		// thread between two groups of four MMAs costs 134 cycles of pipe time; the thread stalls at each tcgen05.mma until the one
		// before it has been dispatched), so whatever the thread does between the last MMA of one accumulator and the first of the
		// next — barrier waits, descriptor arithmetic — leaves the pipe idle beyond the ~95 cycles of that last MMA.  The loop
		// therefore does the next accumulator's bookkeeping BETWEEN the MMAs of the current one, where the thread would stall anyway.
		if (elect_one()) {
			constexpr uint32_t idesc = umma_idesc_f16_f16acc(TC_BM, TC_BN);
			const uint32_t h = warp == 1 ? 0u : 1u;
			const uint32_t a = tmem + (uint32_t)(TC_A_COL) + h * (TC_K / 2);
			const uint32_t d0 = tmem + 2u * h * TC_BN;                  // this half's two accumulators: tile parity 0 / 1
			const uint32_t full0 = smem_u32(&bar->acc_full[2 * h]), empty0 = smem_u32(&bar->acc_empty[2 * h]);
			const uint32_t bfull0 = smem_u32(&bar->b_full[0]);
			const uint64_t db0 = umma_desc_sw128(smem_u32(sB));
			if (n_tiles > 0) {
				mbar_wait(&bar->a_full, 0);                             // the epilogue threads have written the first row block's A into TMEM
				mbar_wait_addr(bfull0, 0);
				tc_fence_after();
			}
			int t = t0;
			uint32_t seg = 1;                                        // row-block segments started
			uint32_t st = 0, sph = 0;                                // B stage of tile n and the parity of its use
			uint64_t db = db0;
			for (int n = 0; n < n_tiles; n++) {
				const uint32_t par = (uint32_t)n & 1u;
				const uint32_t d = d0 + par * TC_BN;
				BS_TR(n, h, 0);
				umma_f16_ts(d, a, db, idesc, 0u);
				// tile n + 1: stage, descriptor, row block
				const bool more = n + 1 < n_tiles;
				uint32_t st1 = st + 1, sph1 = sph;
				uint64_t db1 = db + (TC_TILE_BYTES >> 4);
				if (st1 == (uint32_t)TC_STAGES) { st1 = 0; sph1 ^= 1u; db1 = db0; }
				int t1 = t + 1;
				if (t1 == P.b_tiles) t1 = 0;
				umma_f16_ts(d, a + 8, db + 2, idesc, 1u);
				if (more) mbar_wait_addr(bfull0 + 8u * st1, sph1);      // landed long ago as a rule: the producer runs TC_STAGES tiles ahead
				umma_f16_ts(d, a + 16, db + 4, idesc, 1u);
				// the other slot's epilogue group has read the accumulator of tile n - 1 (use (n - 1) / 2 of that slot): awaited HERE, where the
				// thread would stall on the last MMA anyway, not between this accumulator and the next
				// (one probe: if the group is late, this accumulator is committed first and the wait follows it)
				const bool need = more && n >= 1;
				const uint32_t e_addr = empty0 + 8u * (par ^ 1u), e_par = (((uint32_t)n - 1u) >> 1) & 1u;
				const bool got = need && mbar_try_wait(e_addr, e_par);
				umma_f16_ts(d, a + 24, db + 6, idesc, 1u);
				asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(full0 + 8u * par) : "memory");
				BS_TR(n, h, 1);
				if (need && !got) mbar_wait_addr(e_addr, e_par);
				BS_TR(n, h, 2);
#ifdef BS_TC_TRACE
				if (blockIdx.x == 0 && n >= TR_N0 && n < TR_N0 + 32) g_tc_trace[((n - TR_N0) * 2 + h) * 8 + 6] = got ? 1 : 0;
#endif
				if (more) {
					if (t1 == 0) {                                       // next row block: its A is written once the MMAs above have completed
						mbar_wait(&bar->a_full, seg & 1);
						seg++;
					}
					tc_fence_after();
				}
				st = st1; sph = sph1; db = db1; t = t1;
			}
		}
	} else {
		// ===== epilogue: warps 3..18.  TMEM lane quarter = warp % 4 (hardware rule); group = (warp - 3) / 4 = (A half, tile parity):
		// a group takes every OTHER tile of its half, reads the whole accumulator row in one round and hands the slot back before
		// any arithmetic. =====
		const int q = warp & 3;
		const int g = (warp - 3) >> 2;
		const int h = g >> 1;                                        // A half (rows h*128 .. h*128+127 of the row block)
		const int par = g & 1;                                       // tiles n = par, par + 2, ...
		const int e = (warp - 3) * 32 + lane;                        // 0..511 among the epilogue threads
		const int L = q * 32 + lane;                                 // TMEM lane = row inside the 128-row half
		const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
		const uint32_t run_mine = smem_u32(s_run + (h * TC_BM + L) * 2 + par), run_other = smem_u32(s_run + (h * TC_BM + L) * 2 + (par ^ 1));
		const float nb_max = __uint_as_float(*P.max_norm_b_bits);
		int row = 0;
		bool live = false;
		float na = 0.0f, rs = 1.0f, tol = 0.0f, run = INFINITY;
		// Candidates are collected per thread (a row belongs to two threads inside a segment) and appended to the row's list
		// TC_PEND at a time: ONE atomicAdd reserves the slots, the stores do not depend on each other.
		int n_pend = 0;
		auto flush = [&]() {
			if (n_pend == 0) return;
			const int base = atomicAdd(P.cand_count + row, n_pend);
#pragma unroll
			for (int i = 0; i < TC_PEND; i++)
				if (i < n_pend && base + i < TC_CAND_CAP)
					reinterpret_cast<int2*>(P.cand)[(size_t)row * TC_CAND_CAP + base + i] =
					    make_int2(s_pend[i * (32 * TC_EPI_WARPS) + e], s_pend[(TC_PEND + i) * (32 * TC_EPI_WARPS) + e]);
			n_pend = 0;
		};
		// this thread's row of block `blk` goes into TMEM as a row of the A operand (lane = row, column j = k 2j | 2j+1)
		auto write_a = [&](int blk) {
			uint32_t ar[32];
			const uint4* src = reinterpret_cast<const uint4*>(P.opA + (size_t)(blk * TC_MROWS + h * TC_BM + L) * TC_K);   // padded to whole row blocks
#pragma unroll
			for (int i = 0; i < 8; i++) {
				const uint4 v = __ldg(src + i);
				ar[4 * i] = v.x; ar[4 * i + 1] = v.y; ar[4 * i + 2] = v.z; ar[4 * i + 3] = v.w;
			}
			tmem_st32(lane_base + (uint32_t)(TC_A_COL + h * (TC_K / 2)), ar);
			tc_fence_before();
			__syncwarp();
			if (lane == 0) mbar_arrive(&bar->a_full);
		};
		static_assert(TC_SLOTS == 4 && TC_STAGES % 2 == 0, "slot = 2 h + tile parity; the stage stepping below is written for an even ring");
		const uint32_t slot = (uint32_t)(2 * h + par);               // the accumulators this group reads: one TMEM slot, its k-th use = tile par + 2k
		const uint32_t bfull_addr = smem_u32(&bar->b_full[0]), bempty_addr = smem_u32(&bar->b_empty[0]);
		const uint32_t accfull_addr = smem_u32(&bar->acc_full[slot]), accempty_addr = smem_u32(&bar->acc_empty[slot]);
		const uint32_t snb_addr = smem_u32(s_nb);
		if (par == 0 && n_tiles > 0) write_a(rp0);
		int rp = rp0, t = t0 + par;
		while (t >= P.b_tiles) { t -= P.b_tiles; rp++; }
		int rp_loaded = -1;
		uint32_t st = (uint32_t)par, sph = 0;                        // B stage of tile n and the parity of its use
		uint32_t k = 0;                                              // this group's tile count
		for (int n = par; n < n_tiles; n += 2, k++) {
			if (rp != rp_loaded) {
				// ---- new row block
				flush();
				rp_loaded = rp;
				row = rp * TC_MROWS + h * TC_BM + L;
				live = row < P.nA;
				na = live ? P.norm_a[norm_index(row)] : 0.0f;
				rs = live ? -0.5f * P.norm_a[norm_index(row) + TC_SC] : 1.0f;   // 2^e of this row (split_kernel)
				tol = tolerance_of(na, nb_max);
				// A row's candidates are the columns within tol of its running minimum.  Where the caller knows a distance the row
				// really attains — the second direction takes it from the first one's matches — the minimum starts there: the true
				// minimum and its ties are still within tol (the estimate of a column at most bound = tol/2 above the exact value lies
				// below start + tol; a quarter of tol is added for the float roundings of the start value itself).
				run = live ? INFINITY : -INFINITY;                   // padding rows never pass the candidate test
				if (P.row_bound && live) {
					const double ub = P.row_bound[row];
					if (ub < 1e30) run = __double2float_ru(ub) - na + 0.25f * tol;
				}
			}
			// The row's other thread (the tiles in between) shares its running minimum: a value tagged with the row block it belongs
			// to, one 64-bit shared-memory word per thread.  A stale value is still a valid (looser) bound, so every other tile is
			// enough after the first eight (where the minimum still moves fast and short shares of a CTA end).
			if (k < 8 || (k & 1) == 0) {
				const unsigned long long mine = ((unsigned long long)(uint32_t)rp << 32) | __float_as_uint(run);
				unsigned long long other;
				asm volatile("st.shared.b64 [%0], %1;" ::"r"(run_mine), "l"(mine) : "memory");
				asm volatile("ld.shared.b64 %0, [%1];" : "=l"(other) : "r"(run_other) : "memory");
				if ((int)(other >> 32) == rp) run = fminf(run, __uint_as_float((uint32_t)other));
			}
			mbar_wait_addr(bfull_addr + 8u * st, sph);               // the stage's norms have landed (long since: the MMAs needed its tile)
			mbar_wait_addr(accfull_addr, k & 1u);
			if (q == 0 && lane == 0) BS_TR(n, h, 3);
			tc_fence_after();
			// The whole accumulator row in one round: 112 fp16 values, two per register (tcgen05.ld .pack::16b), and the slot goes back to the
			// MMA issuer before any arithmetic.
			const int col0 = t * TC_BN;
			const uint32_t tbase = lane_base + slot * TC_BN;
			uint32_t r0[32], r1[16], r2[8];                              // columns [0, 64), [64, 96), [96, 112)
			tmem_ld64p_issue(tbase, r0); tmem_ld32p_issue(tbase + 64, r1); tmem_ld16p_issue(tbase + 96, r2);
			tmem_ld_wait(r0); tmem_ld_wait(r1); tmem_ld_wait(r2);
			if (q == 0 && lane == 0) BS_TR(n, h, 4);
			tc_fence_before();
			__syncwarp();
			if (lane == 0) mbar_arrive_addr(accempty_addr);
			if (q == 0 && lane == 0) BS_TR(n, h, 5);
			const uint32_t nb_addr = snb_addr + (uint32_t)(TC_NB * 4) * st;     // the tile's record: 112 norms, flag, 112 scales
			// Candidate path of one 32-column chunk (16 registers; the last chunk of a tile has 16 columns) whose best column came within tol
			// of the row's running minimum (per row ~1.3 / columns-seen per column; a 32-lane warp takes it whenever ONE of its rows does).
			// Kept small on purpose (two larger variants lost more to instruction fetch and serialisation than they saved).  A hit goes to the
			// thread's pending list as (column, cmin): the chunk's best estimate stands in for the estimate of each of its hits — never above
			// the hit's own, and the estimate of a column that is in the list, which is all the re-score's pruning rule needs.
			// val(k): estimate-side value of column k of the chunk; hits are val <= lim (LE) or val >= lim (!LE)
			auto scan = [&](bool want, float cmin, int cbase, int ncols, float lim, auto val, auto le) {
				constexpr bool LE = decltype(le)::value;
				if (want) {
					unsigned hits = 0;
#pragma unroll
					for (int kk = 0; kk < 32; kk++)
						if (kk < ncols) hits |= ((LE ? val(kk) <= lim : val(kk) >= lim) ? 1u : 0u) << kk;
					while (hits) {
						const int kk = __ffs(hits) - 1;
						hits &= hits - 1;
						s_pend[n_pend * (32 * TC_EPI_WARPS) + e] = col0 + cbase + kk;
						s_pend[(TC_PEND + n_pend) * (32 * TC_EPI_WARPS) + e] = __float_as_int(cmin);
						if (++n_pend == TC_PEND) flush();
					}
				}
			};
			// pair maximum of `cnt` registers starting at r[first] (cnt a power of two), both halves
			auto pmax = [&](const uint32_t* r, int first, int cnt) {
				uint32_t m[16];
#pragma unroll
				for (int i = 0; i < 16; i++) m[i] = i < cnt ? r[first + i] : 0u;
#pragma unroll
				for (int w = cnt / 2; w >= 1; w >>= 1)
#pragma unroll
					for (int i = 0; i < 16; i++)
						if (i < w) m[i] = hmax2(m[i], m[i + w]);
				return m[0];
			};
			// SURF descriptors are unit vectors: where the 112 column norms of a tile are one and the same float (flagged by
			// norm_flags_kernel; equal norms have equal scales) min_j (nb - 2 s_j) = nb - 2 max_j s_j (fma is monotonic), and the row maximum
			// of the raw accumulators is all the fast path needs: a quarter of an instruction per element (packed fp16 maximum).
			if (!DBG && lds32(nb_addr + (uint32_t)(TC_BN * 4)) != 0u) {
				const float nbc = __uint_as_float(lds32(nb_addr));
				const float sc = __uint_as_float(lds32(nb_addr + (uint32_t)(TC_SC * 4))) * rs;      // -2 * 2^(e_row + e_columns): exact
				const uint32_t m0 = pmax(r0, 0, 16), m1 = pmax(r0, 16, 16), m2 = pmax(r1, 0, 16), m3 = pmax(r2, 0, 8);
				const uint32_t mt = hmax2(hmax2(m0, m1), hmax2(m2, m3));
				// one test for the tile: nothing to do while its best column stays more than tol above the running minimum
				if (fmaf(sc, fmaxf(h2lo(mt), h2hi(mt)), nbc) <= run + tol) {
					// s >= lim is implied by nb + sc s <= thr (sc < 0): the difference over |sc|, less a margin for the roundings of both sides
					// The hits of a wanted chunk with packed compares: lim rounded DOWN to fp16 (a superset of s >= lim: the values are fp16),
					// set.ge per register pair, the two result bits moved to bit i (even column 2i) and bit 16 + i (odd column 2i + 1) of one
					// word by a multiply — three instructions per two columns instead of two conversions, two compares and two bit moves.
					auto chunk = [&](uint32_t mp, int cbase, int ncols, const uint32_t* r, int first) {
						const float cmin = fmaf(sc, fmaxf(h2lo(mp), h2hi(mp)), nbc);
						const bool want = cmin <= run + tol && cmin < INFINITY;
						if (want) {
							run = fminf(run, cmin);
							const float thr = run + tol;
							const float lim = (nbc - thr) / -sc - 1e-6f * (fabsf(nbc) + fabsf(thr)) / -sc;
							uint32_t lim2;
							asm("{\n\t.reg .b16 l;\n\tcvt.rm.f16.f32 l, %1;\n\tmov.b32 %0, {l, l};\n\t}" : "=r"(lim2) : "f"(fmaxf(lim, -65504.0f)));
							unsigned hits = 0;
#pragma unroll
							for (int i = 0; i < 16; i++) {
								if (2 * i < ncols) {
									uint32_t ge;
									asm("set.ge.u32.f16x2 %0, %1, %2;" : "=r"(ge) : "r"(r[first + i]), "r"(lim2));
									hits += (ge & 0x00010001u) * (1u << i);
								}
							}
							while (hits) {
								const int b = __ffs(hits) - 1;
								hits &= hits - 1;
								s_pend[n_pend * (32 * TC_EPI_WARPS) + e] = col0 + cbase + 2 * (b & 15) + (b >> 4);
								s_pend[(TC_PEND + n_pend) * (32 * TC_EPI_WARPS) + e] = __float_as_int(cmin);
								if (++n_pend == TC_PEND) flush();
							}
						}
					};
					chunk(m0, 0, 32, r0, 0);
					chunk(m1, 32, 32, r0, 16);
					chunk(m2, 64, 32, r1, 0);
					chunk(m3, 96, 16, r2, 0);
				}
			} else {
				// general norms: d'_j = |b_j|^2 + (-2 * 2^e_j * 2^e_row) s_j column by column in fp32
				auto chunk = [&](int cbase, int ncols, const uint32_t* r, int first) {
					float d[32];
					float cmin = INFINITY;
#pragma unroll
					for (int kk = 0; kk < 32; kk++) {
						if (kk < ncols) {
							const float sv = (kk & 1) ? h2hi(r[first + (kk >> 1)]) : h2lo(r[first + (kk >> 1)]);
							const float nbj = __uint_as_float(lds32(nb_addr + (uint32_t)((cbase + kk) * 4)));
							const float scj = __uint_as_float(lds32(nb_addr + (uint32_t)((TC_SC + cbase + kk) * 4))) * rs;
							d[kk] = fmaf(scj, sv, nbj);
							cmin = fminf(cmin, d[kk]);
						} else d[kk] = INFINITY;
					}
					if (DBG) {
						if (P.dbg && live) {
#pragma unroll
							for (int kk = 0; kk < 32; kk++)
								if (kk < ncols && col0 + cbase + kk < P.nB) P.dbg[(size_t)row * P.nB + col0 + cbase + kk] = d[kk] + na;
						}
					}
					const bool want = cmin <= run + tol && cmin < INFINITY;
					if (want) run = fminf(run, cmin);
					scan(want, cmin, cbase, ncols, run + tol, [&](int kk) { return d[kk]; }, std::true_type());
				};
				chunk(0, 32, r0, 0);
				chunk(32, 32, r0, 16);
				chunk(64, 32, r1, 0);
				chunk(96, 16, r2, 0);
			}
			__syncwarp();
			if (lane == 0) mbar_arrive_addr(bempty_addr + 8u * st);  // this warp is done with the stage's norms
			// The last tile of a row block: the MMAs that read this half of A have completed (this warp waited for the last of them
			// above), so the next block's rows can go in; the other group is still on the tile before and never touches A.
			if (t == P.b_tiles - 1 && n + 1 < n_tiles) write_a(rp + 1);
			// this group's next tile (n + 2): stage and its phase, column tile and row block
			st += 2;
			if (st >= (uint32_t)TC_STAGES) { st -= TC_STAGES; sph ^= 1u; }
			t += 2;
			while (t >= P.b_tiles) { t -= P.b_tiles; rp++; }
		}
		flush();
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) { tc_fence_after(); tmem_dealloc<TC_TMEM_COLS>(tmem); }
}

// ---------------------------------------------------------------- 3. exact re-score
// one THREAD per row of A: exact distance to the candidate columns of B that survive the pruning; mode 0 -> pairs/fit, mode 1 ->
// column statistics.
// A candidate carries the filter's estimate d' of its column: only those within tol of the smallest estimate of the list can be the
// exact minimum or tie with it (each estimate is within tol / 2 of the exact value), so the early entries of a list — emitted while the
// row's running minimum was still loose — are dropped without touching their descriptors: ~17 -> ~1.5 exact distances per row.
// With so few survivors a warp per row left 30 of 32 lanes idle through 192 fp64 instructions per candidate (the fp64 pipe was the
// limit: 116 us per 100k rows); with a row per lane the lanes walk their survivors in lockstep (first survivor of every row, then
// the second, ...), so a warp's fp64 instructions serve 32 rows.
constexpr int RS_KEEP = 4;                 // survivors held in registers; a row with more takes the slow loop below
__device__ __forceinline__ double exact_sq64(const double* __restrict__ a, const double* __restrict__ b) {
	// DescriptorDistance.euclideanSq :55-64, same order, no FMA; 16 values of each row in flight ahead of the serial sum
	const double2* a2 = reinterpret_cast<const double2*>(a);
	const double2* b2 = reinterpret_cast<const double2*>(b);
	double s = 0.0;
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
	return s;
}
__global__ void __launch_bounds__(128)
rescore_kernel(const double* __restrict__ A, int nA, const double* __restrict__ B, const int* __restrict__ cand_count,
               const int* __restrict__ cand, const float* __restrict__ norm_a, const unsigned* __restrict__ max_norm_b_bits,
               int mode, double max_fit, int index_offset, int32_t* __restrict__ pairs,
               double* __restrict__ fit, double* __restrict__ o_min, int32_t* __restrict__ o_first,
               int32_t* __restrict__ o_last, int* __restrict__ overflow) {
	const int row = blockIdx.x * blockDim.x + threadIdx.x;
	const bool live = row < nA;
	int cnt = live ? cand_count[row] : 0;
	if (cnt > TC_CAND_CAP) { atomicExch(overflow, 1); cnt = TC_CAND_CAP; }
	const int2* list = reinterpret_cast<const int2*>(cand) + (size_t)(live ? row : 0) * TC_CAND_CAP;
	float est_min = INFINITY;
	for (int c = 0; c < cnt; c++) est_min = fminf(est_min, __int_as_float(list[c].y));
	const float keep = live ? est_min + tolerance_of(norm_a[norm_index(row)], __uint_as_float(*max_norm_b_bits)) : 0.0f;
	int kept[RS_KEEP];
	int nk = 0;
	for (int c = 0; c < cnt; c++) {
		const int2 cj = list[c];
		if (__int_as_float(cj.y) <= keep) {
#pragma unroll
			for (int t = 0; t < RS_KEEP; t++) if (nk == t) kept[t] = cj.x;
			nk++;
		}
	}
	const double* a = A + (size_t)(live ? row : 0) * TC_DOF;
	double best = CUDART_INF;
	int first = -1, last = -1;
	auto take = [&](int j) {
		const double s = exact_sq64(a, B + (size_t)j * TC_DOF);
		if (first < 0 || s < best) { best = s; first = j; last = j; }
		else if (s == best) { first = min(first, j); last = max(last, j); }
	};
	const int rounds = __reduce_max_sync(0xffffffffu, min(nk, RS_KEEP));
	for (int t = 0; t < rounds; t++) {
		int j = -1;
#pragma unroll
		for (int u = 0; u < RS_KEEP; u++) if (t == u && u < nk) j = kept[u];
		if (j >= 0) take(j);
	}
	if (nk > RS_KEEP) {                     // many near-ties (duplicates): the remaining survivors, one by one
		int seen = 0;
		for (int c = 0; c < cnt; c++) {
			const int2 cj = list[c];
			if (__int_as_float(cj.y) <= keep) { if (seen >= RS_KEEP) take(cj.x); seen++; }
		}
	}
	if (!live) return;
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

// operand matrix [rows_pad][64] fp16, box = 64 elements (one 128-byte swizzle atom) x 112 rows
static CUtensorMap operand_map(const void* base, int rows_pad) {
	CUtensorMap m;
	const cuuint64_t dims[2] = {(cuuint64_t)TC_K, (cuuint64_t)rows_pad};
	const cuuint64_t strides[1] = {(cuuint64_t)TC_K * 2};
	const cuuint32_t box[2] = {64, (cuuint32_t)TC_BN};
	const cuuint32_t estr[2] = {1, 1};
	const CUresult r = encode_tiled_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
	                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
	                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	if (r != CUDA_SUCCESS) fail(BS_ERR_CUDA, "cuTensorMapEncodeTiled failed");
	return m;
}

int tc_pad_rows(int n) { return ((n + TC_MROWS - 1) / TC_MROWS) * TC_MROWS; }
size_t tc_operand_bytes(int n) { return (size_t)tc_pad_rows(n) * TC_K * 2; }
static int tc_norm_tiles(int n) { return (tc_pad_rows(n) + TC_BN - 1) / TC_BN; }   // covers the rows a row block reads (A side) and whole tiles (B side)
size_t tc_norm_bytes(int n) { return (size_t)tc_norm_tiles(n) * TC_NB * 4; }
int tc_cand_cap() { return TC_CAND_CAP; }
size_t tc_cand_bytes(size_t rows) { return rows * TC_CAND_CAP * 8; }   // (column, estimate) per candidate
bool tc_supported(int dof) { return dof == TC_DOF; }

void launch_tc_split(const double* d_x, int n, void* d_operand, float* d_norm, unsigned* d_max_norm_bits, cudaStream_t st) {
	const int n_pad = tc_pad_rows(n);
	BS_CUDA(cudaMemsetAsync(d_max_norm_bits, 0, 4, st));
	const int tiles = tc_norm_tiles(n);
	BS_LAUNCH(split_kernel, div_up(tiles * TC_BN, 8), 256, 0, st, d_x, n, n_pad, tiles * TC_BN, static_cast<__half*>(d_operand), d_norm);
	BS_LAUNCH(norm_flags_kernel, div_up(tiles, 8), 256, 0, st, d_norm, tiles, n, d_max_norm_bits);
}

// rows of A against all rows of B: candidate lists for the rows of A
void launch_tc_filter(const void* d_opA, const float* d_normA, int nA, const void* d_opB, const float* d_normB,
                      const unsigned* d_max_norm_b_bits, int nB, int* d_cand_count, int* d_cand, float* d_dbg, int sm_count,
                      cudaStream_t st, const double* d_row_bound) {
	if (nA <= 0 || nB <= 0) return;
	// the attribute belongs to the device: set it on every launch (a handle may live on any device of the process)
	BS_CUDA(cudaFuncSetAttribute(filter_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
	BS_CUDA(cudaFuncSetAttribute(filter_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
	FilterArgs P{};
	P.nA = nA; P.nB = nB;
	P.a_blocks = tc_pad_rows(nA) / TC_MROWS;
	P.b_tiles = div_up(nB, TC_BN);
	P.opA = static_cast<const __half*>(d_opA);
	P.norm_a = d_normA; P.norm_b = d_normB; P.max_norm_b_bits = d_max_norm_b_bits;
	P.cand_count = d_cand_count; P.cand = d_cand; P.dbg = d_dbg;
	P.row_bound = d_row_bound;
	const CUtensorMap tmB = operand_map(d_opB, tc_pad_rows(nB));
	BS_CUDA(cudaMemsetAsync(d_cand_count, 0, (size_t)nA * 4, st));
	// one contiguous share of the (row block, tile) sequence per CTA; a CTA that takes a share loads a row block's A, so at least
	// four tiles per CTA when the problem is small
	const long long total = (long long)P.a_blocks * P.b_tiles;
	const int grid = (int)std::max<long long>(1, std::min<long long>(sm_count, total / 4));
	if (d_dbg) BS_LAUNCH(filter_kernel<true>, grid, TC_THREADS, TC_SMEM_BYTES, st, tmB, P);
	else BS_LAUNCH(filter_kernel<false>, grid, TC_THREADS, TC_SMEM_BYTES, st, tmB, P);
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

void launch_tc_rescore(const double* d_A, int nA, const double* d_B, const int* d_cand_count, const int* d_cand, const float* d_norm_a,
                       const unsigned* d_max_norm_b_bits, int mode, double max_fit,
                       int index_offset, int32_t* pairs, double* fit, double* o_min, int32_t* o_first, int32_t* o_last, int* d_overflow,
                       cudaStream_t st) {
	if (nA <= 0) return;
	BS_LAUNCH(rescore_kernel, div_up(nA, 128), 128, 0, st, d_A, nA, d_B, d_cand_count, d_cand, d_norm_a, d_max_norm_b_bits, mode, max_fit, index_offset, pairs, fit, o_min,
	          o_first, o_last, d_overflow);
}

}  // namespace bs
#ifdef BS_TC_TRACE
extern "C" int bs_debug_tc_trace(long long* out) { return (int)cudaMemcpyFromSymbol(out, bs::g_tc_trace, sizeof(bs::g_tc_trace)); }
#endif
