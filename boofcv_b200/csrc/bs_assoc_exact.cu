// bs_assoc_exact.cu — greedy brute-force association with exact fp64 scores.
// Reference: AssociateGreedy.associate (F/alg/feature/associate/AssociateGreedy.java:65-118) with
// ScoreAssociateEuclideanSq_F64 -> DescriptorDistance.euclideanSq (F/alg/descriptor/DescriptorDistance.java:55-64):
//   score(a,b) = sum_k (a_k-b_k)^2, accumulated sequentially in k from 0, no FMA.
//   forward : per src row, the LAST dst index attaining the row minimum, accepted iff min <= maxFitError (:76-95)
//   backward: a match (i,m) survives iff W[i][m] is the strict unique minimum of column m (:98-117)
// The reference stores the N x M score matrix; here the matrix is never materialised: one tiled kernel computes, for every
// row of A against all rows of B, (min score, first index, last index attaining it).  Run as (src,dst) it is the forward
// pass; run as (dst,src) it gives the column minima for the backward test ((a-b)^2 == (b-a)^2 exactly, so both runs see
// bit-identical scores).
#include "bs_common.cuh"

#include <cfloat>
#include <math_constants.h>

namespace bs {

constexpr int AT = 64;            // tile: AT rows of A x AT rows of B
constexpr int AK = 64;            // k-chunk of B held in shared memory
constexpr int APAD = AT + 1;      // row stride (doubles) of the [k][row] tiles
constexpr int ATHREADS = 256;     // 16 x 16 threads, 4 x 4 pairs each
constexpr int AMAX_DOF = 256;

struct RowBest {
	double best; int first, last;
	__device__ __forceinline__ void init() { best = CUDART_INF; first = -1; last = -1; }
	// scores arrive in ascending j
	__device__ __forceinline__ void push(double f, int j) {
		if (f < best || (first < 0 && f <= best)) { best = f; first = j; last = j; }
		else if (f == best) last = j;
	}
	// merge a partial result that covers LARGER indices than this one
	__device__ __forceinline__ void merge_later(double f, int fi, int la) {
		if (fi < 0) return;
		if (first < 0 || f < best) { best = f; first = fi; last = la; }
		else if (f == best) last = la;
	}
};

// partial[(split*nA + row)] over the B rows [split*b_per_split, (split+1)*b_per_split)
__global__ void __launch_bounds__(ATHREADS)
rowmin_tile_kernel(const double* __restrict__ A, int nA, const double* __restrict__ B, int nB, int dof, int b_per_split,
                   double* __restrict__ p_min, int* __restrict__ p_first, int* __restrict__ p_last) {
	extern __shared__ double sm[];
	double* sA = sm;                       // [dof][APAD]
	double* sB = sm + (long)dof * APAD;    // [AK][APAD]
	const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
	const int row0 = blockIdx.x * AT;
	const int b_begin = blockIdx.y * b_per_split;
	const int b_end = min(nB, b_begin + b_per_split);

	// A tile, transposed to [k][row]; rows beyond nA are zero (their results are discarded)
	for (int idx = tid; idx < AT * dof; idx += ATHREADS) {
		const int r = idx / dof, k = idx - r * dof;
		sA[k * APAD + r] = (row0 + r < nA) ? A[(long)(row0 + r) * dof + k] : 0.0;
	}
	RowBest best[4];
#pragma unroll
	for (int r = 0; r < 4; r++) best[r].init();

	for (int col0 = b_begin; col0 < b_end; col0 += AT) {
		double acc[4][4];
#pragma unroll
		for (int r = 0; r < 4; r++)
#pragma unroll
			for (int c = 0; c < 4; c++) acc[r][c] = 0.0;
		for (int k0 = 0; k0 < dof; k0 += AK) {
			const int kn = min(AK, dof - k0);
			__syncthreads();       // previous chunk consumed (and sA visible on the first pass)
			for (int idx = tid; idx < AT * kn; idx += ATHREADS) {
				const int c = idx / kn, k = idx - c * kn;
				sB[k * APAD + c] = (col0 + c < b_end) ? B[(long)(col0 + c) * dof + k0 + k] : 0.0;
			}
			__syncthreads();
			for (int k = 0; k < kn; k++) {
				double a[4], b[4];
#pragma unroll
				for (int r = 0; r < 4; r++) a[r] = sA[(k0 + k) * APAD + ty + 16 * r];
#pragma unroll
				for (int c = 0; c < 4; c++) b[c] = sB[k * APAD + tx + 16 * c];
#pragma unroll
				for (int r = 0; r < 4; r++)
#pragma unroll
					for (int c = 0; c < 4; c++) {
						const double d = a[r] - b[c];
						acc[r][c] += d * d;      // -fmad=false: a separate multiply and add, like the Java
					}
			}
		}
#pragma unroll
		for (int c = 0; c < 4; c++) {
			const int j = col0 + tx + 16 * c;
			if (j < b_end) {
#pragma unroll
				for (int r = 0; r < 4; r++) best[r].push(acc[r][c], j);
			}
		}
	}
	// combine the 16 threads (tx) that share a row.  A thread's columns are tx+16c (+tile): not contiguous across threads,
	// so merge on (score, index) directly: smaller score wins; equal -> first = min, last = max.
	__syncthreads();
	double* r_min = sm;                                  // [AT][16]
	int* r_first = reinterpret_cast<int*>(sm + AT * 16);
	int* r_last = r_first + AT * 16;
#pragma unroll
	for (int r = 0; r < 4; r++) {
		const int row = ty + 16 * r;
		r_min[row * 16 + tx] = best[r].best; r_first[row * 16 + tx] = best[r].first; r_last[row * 16 + tx] = best[r].last;
	}
	__syncthreads();
	if (tid < AT && row0 + tid < nA) {
		double m = CUDART_INF; int fi = -1, la = -1;
		for (int t = 0; t < 16; t++) {
			const double f = r_min[tid * 16 + t];
			const int a = r_first[tid * 16 + t], b = r_last[tid * 16 + t];
			if (a < 0) continue;
			if (fi < 0 || f < m) { m = f; fi = a; la = b; }
			else if (f == m) { fi = min(fi, a); la = max(la, b); }
		}
		const long o = (long)blockIdx.y * nA + row0 + tid;
		p_min[o] = m; p_first[o] = fi; p_last[o] = la;
	}
}

// combine the column splits (ascending index ranges).  mode 0: forward -> pairs/fit (threshold applied);
// mode 1: column statistics with indices shifted by index_offset (global row numbering of a sharded run)
__global__ void rowmin_finalize_kernel(const double* __restrict__ p_min, const int* __restrict__ p_first, const int* __restrict__ p_last,
                                       int nA, int splits, int mode, double max_fit, int index_offset, int32_t* __restrict__ pairs,
                                       double* __restrict__ fit, double* __restrict__ o_min, int32_t* __restrict__ o_first,
                                       int32_t* __restrict__ o_last) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nA) return;
	RowBest b; b.init();
	for (int s = 0; s < splits; s++) {
		const long o = (long)s * nA + i;
		b.merge_later(p_min[o], p_first[o], p_last[o]);
	}
	if (mode == 0) {
		// bestScore starts at maxFitError and `fit <= bestScore` replaces it (:80-91)
		const bool hit = b.first >= 0 && b.best <= max_fit;
		pairs[i] = hit ? b.last : -1;
		fit[i] = hit ? b.best : max_fit;
	} else {
		o_min[i] = b.best;
		o_first[i] = b.first < 0 ? -1 : b.first + index_offset;
		o_last[i] = b.last < 0 ? -1 : b.last + index_offset;
	}
}

// :98-117 — forward match (i,m) is kept iff no other row j has W[j][m] <= W[i][m]
__global__ void validate_kernel(int n, int row_offset, int32_t* __restrict__ pairs, double* __restrict__ fit,
                                const double* __restrict__ col_min, const int32_t* __restrict__ col_first,
                                const int32_t* __restrict__ col_last) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int m = pairs[i];
	if (m < 0) return;
	const int gi = row_offset + i;
	const bool unique_min = col_min[m] == fit[i] && col_first[m] == gi && col_last[m] == gi;
	if (!unique_min) { pairs[i] = -1; fit[i] = DBL_MAX; }
}

// merge the column partials of `ranks` shards (all-gathered): min score; first = min, last = max over the shards attaining it
// (rank r's arrays start r * stride_min doubles / r * stride_idx ints after rank 0's: [ranks][m] arrays or one packed record per rank)
__global__ void merge_columns_kernel(int ranks, int m, long stride_min, long stride_idx, const double* __restrict__ a_min,
                                     const int32_t* __restrict__ a_first, const int32_t* __restrict__ a_last, double* __restrict__ o_min,
                                     int32_t* __restrict__ o_first, int32_t* __restrict__ o_last) {
	const int j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= m) return;
	double best = CUDART_INF; int fi = -1, la = -1;
	for (int r = 0; r < ranks; r++) {
		const long oi = (long)r * stride_idx + j;
		const int a = a_first[oi], b = a_last[oi];
		if (a < 0) continue;
		const double f = a_min[(long)r * stride_min + j];
		if (fi < 0 || f < best) { best = f; fi = a; la = b; }
		else if (f == best) { fi = min(fi, a); la = max(la, b); }
	}
	o_min[j] = best; o_first[j] = fi; o_last[j] = la;
}

// ---------------------------------------------------------------- host launchers

int rowmin_splits(int nA, int nB, int sm_count) {
	const int row_ctas = div_up(nA, AT);
	int splits = div_up(2 * sm_count, row_ctas > 0 ? row_ctas : 1);
	const int max_splits = div_up(nB, AT);
	if (splits > max_splits) splits = max_splits;
	if (splits < 1) splits = 1;
	return splits;
}

void launch_rowmin(const double* A, int nA, const double* B, int nB, int dof, int splits, const RowminScratch& s, cudaStream_t st) {
	if (dof < 1 || dof > AMAX_DOF) fail(BS_ERR_ARG, "descriptor length must be in [1,256]");
	int per = div_up(div_up(nB, splits), AT) * AT;
	if (per < AT) per = AT;
	const size_t smem = ((size_t)dof * APAD + (size_t)AK * APAD) * sizeof(double);
	// per-device attribute: set on every launch (a handle may live on any device of the process)
	BS_CUDA(cudaFuncSetAttribute(rowmin_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
	                             (int)(((size_t)AMAX_DOF * APAD + (size_t)AK * APAD) * sizeof(double))));
	BS_LAUNCH(rowmin_tile_kernel, dim3(div_up(nA, AT), splits), ATHREADS, smem, st, A, nA, B, nB, dof, per, s.p_min, s.p_first,
	          s.p_last);
}

void launch_rowmin_finalize(const RowminScratch& s, int nA, int splits, int mode, double max_fit, int index_offset, int32_t* pairs,
                            double* fit, double* o_min, int32_t* o_first, int32_t* o_last, cudaStream_t st) {
	BS_LAUNCH(rowmin_finalize_kernel, div_up(nA, 256), 256, 0, st, s.p_min, s.p_first, s.p_last, nA, splits, mode, max_fit,
	          index_offset, pairs, fit, o_min, o_first, o_last);
}

void launch_validate(int n, int row_offset, int32_t* pairs, double* fit, const double* col_min, const int32_t* col_first,
                     const int32_t* col_last, cudaStream_t st) {
	if (n <= 0) return;
	BS_LAUNCH(validate_kernel, div_up(n, 256), 256, 0, st, n, row_offset, pairs, fit, col_min, col_first, col_last);
}

void launch_merge_columns(int ranks, int m, const double* a_min, const int32_t* a_first, const int32_t* a_last, double* o_min,
                          int32_t* o_first, int32_t* o_last, cudaStream_t st, long stride_min, long stride_idx) {
	if (m <= 0) return;
	if (stride_min <= 0) stride_min = m;
	if (stride_idx <= 0) stride_idx = m;
	BS_LAUNCH(merge_columns_kernel, div_up(m, 256), 256, 0, st, ranks, m, stride_min, stride_idx, a_min, a_first, a_last, o_min, o_first, o_last);
}

}  // namespace bs

// ====================================================================================================================
// bs_assoc_forward_static — forward greedy association that allocates NOTHING.  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops, harnesses that forbid it).  Same kernels and the same results as the exact path of bs_assoc_run_device
// without backward validation; the split partials live in a file-scope device array instead of the handle's workspace.
//
// The array exists once per GPU that loads the library (4 MiB, held until exit).  Its NAME is a host-side stand-in in
// host code, not a device address: the address is resolved once per device with cudaGetSymbolAddress (that device
// current) and passed to the kernels.  Every call shares the array: calls that overlap on the GPU (two streams, two host
// threads, two processes' worth of handles on one context) race on it — serialise them.  Nothing in it has to start at
// zero: every entry the finalize pass reads is written by the tile pass of the same call, so the previous call's contents
// are never observed.
// ====================================================================================================================
#include <mutex>

namespace {

constexpr long kStaticEntries = 2L * BS_ASSOC_STATIC_MAX_ROWS;           // split partials: (min, first, last) per (split, row)
constexpr size_t kStaticBytes = (size_t)kStaticEntries * 16;            // 8 + 4 + 4 bytes per entry = 4 MiB

}  // namespace

static __device__ __align__(256) unsigned char g_assoc_static_scratch[kStaticBytes];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
unsigned char* g_static_ptr[kMaxDevices];     // device address of g_assoc_static_scratch on each device (nullptr: not set up yet)
int g_static_sms[kMaxDevices];

struct CurrentDevice {                         // make `dev` current for the scope of the call
	int prev = -1;
	explicit CurrentDevice(int dev) {
		BS_CUDA(cudaGetDevice(&prev));
		if (prev != dev) BS_CUDA(cudaSetDevice(dev)); else prev = -1;
	}
	~CurrentDevice() { if (prev >= 0) cudaSetDevice(prev); }
};

// once per device, with that device current: the array's device address, the kernel's shared-memory opt-in, the SM count
unsigned char* static_scratch_of(int device, int* sms) {
	std::lock_guard<std::mutex> lock(g_static_mu);
	if (!g_static_ptr[device]) {
		unsigned char* p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_assoc_static_scratch);
		if (e != cudaSuccess || !p) bs::fail(BS_ERR_CUDA, std::string("cudaGetSymbolAddress(static association scratch): ") + cudaGetErrorString(e));
		BS_CUDA(cudaFuncSetAttribute(bs::rowmin_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
		                             (int)(((size_t)bs::AMAX_DOF * bs::APAD + (size_t)bs::AK * bs::APAD) * sizeof(double))));
		BS_CUDA(cudaDeviceGetAttribute(&g_static_sms[device], cudaDevAttrMultiProcessorCount, device));
		g_static_ptr[device] = p;
	}
	*sms = g_static_sms[device];
	return g_static_ptr[device];
}

}  // namespace

extern "C" int bs_assoc_forward_static(int device, const double* d_src, int n, const double* d_dst, int m, int dof, double max_fit_error,
                                       int32_t* d_pairs, double* d_fit, void* stream) {
	using namespace bs;
	try {
		if (n < 0 || m < 0) fail(BS_ERR_ARG, "bad sizes");
		if (dof < 1 || dof > AMAX_DOF) fail(BS_ERR_ARG, "descriptor length must be in [1,256]");
		if (n > BS_ASSOC_STATIC_MAX_ROWS)
			fail(BS_ERR_ARG, "bs_assoc_forward_static: " + std::to_string(n) + " source rows exceed the static scratch (BS_ASSOC_STATIC_MAX_ROWS = " +
			                          std::to_string(BS_ASSOC_STATIC_MAX_ROWS) + "); use bs_assoc_run_device");
		if (n == 0) return BS_OK;
		if (!d_src || !d_pairs || !d_fit || (m > 0 && !d_dst)) fail(BS_ERR_ARG, "null argument");
		int ndev = 0;
		BS_CUDA(cudaGetDeviceCount(&ndev));
		if (device < 0 || device >= ndev || device >= kMaxDevices) fail(BS_ERR_ARG, "no such CUDA device");
		CurrentDevice cur(device);
		int sms = 0;
		unsigned char* scratch = static_scratch_of(device, &sms);
		RowminScratch s{reinterpret_cast<double*>(scratch), reinterpret_cast<int*>(scratch + (size_t)kStaticEntries * 8),
		                reinterpret_cast<int*>(scratch + (size_t)kStaticEntries * 12), kStaticEntries};
		cudaStream_t st = static_cast<cudaStream_t>(stream);
		int splits = 0;                               // m == 0: nothing to score, finalize alone writes -1 / max_fit_error
		if (m > 0) {
			splits = rowmin_splits(n, m, sms);
			if ((long)splits * n > kStaticEntries) splits = (int)(kStaticEntries / n);      // >= 2: n <= kStaticEntries / 2
			int per = div_up(div_up(m, splits), AT) * AT;
			if (per < AT) per = AT;
			splits = div_up(m, per);                  // no empty trailing split
			const size_t smem = ((size_t)dof * APAD + (size_t)AK * APAD) * sizeof(double);
			BS_LAUNCH(rowmin_tile_kernel, dim3(div_up(n, AT), splits), ATHREADS, smem, st, d_src, n, d_dst, m, dof, per, s.p_min, s.p_first, s.p_last);
		}
		BS_LAUNCH(rowmin_finalize_kernel, div_up(n, 256), 256, 0, st, s.p_min, s.p_first, s.p_last, n, splits, 0, max_fit_error, 0, d_pairs, d_fit,
		          (double*)nullptr, (int32_t*)nullptr, (int32_t*)nullptr);
		return BS_OK;
	} catch (const bs::Error& e) {
		bs::set_last_error(e.what());
		return e.code;
	} catch (const std::exception& e) {
		bs::set_last_error(e.what());
		return BS_ERR_CUDA;
	}
}
