// bs_integral.cu — integral image, bit-exact with ImplIntegralImageOps.transform
// (IP/alg/transform/ii/impl/ImplIntegralImageOps.java:42-66 F32, :94-118 U8->S32, :120-144 S32->S32).
//
// The reference computes, per row, a serial running sum `total += in[x]` and then `out[y][x] = out[y-1][x] + total`.
// For float images the association order of those adds is observable (sums exceed 2^24 at 1080p), and the detected
// point set depends on the last bits of the integral, so the scan is kept in the reference's order:
//   pass 1 (rows):    one lane owns one row; 32x32 tiles go through shared memory so that global loads/stores are
//                     coalesced (lane = column) while the adds run lane = row, left to right, carry in a register.
//   pass 2 (columns): one thread owns one column and adds the row sums top to bottom (coalesced across the warp).
// Parallelism comes from rows x frames and columns x frames.  Integer images use the same kernels with uint32
// wrap-around arithmetic (== Java int).
#include "bs_common.cuh"
#include <type_traits>

namespace bs {

constexpr int ROWS_WARPS = 2;        // warps per CTA in the row pass (32 rows each)
constexpr int COL_THREADS = 64;      // threads per CTA in the column pass
#ifndef BS_COL_UNROLL
#define BS_COL_UNROLL 32
#endif
constexpr int COL_UNROLL = BS_COL_UNROLL;       // rows per step in the column pass (one step ahead is in flight)

template <typename TI, typename TO, typename ACC>
__global__ void __launch_bounds__(ROWS_WARPS * 32)
integral_rows_kernel(const TI* __restrict__ in, long in_frame_stride, int in_stride, TO* __restrict__ out, int W, int H) {
	__shared__ ACC tile[ROWS_WARPS][32][33];
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int y0 = (blockIdx.x * ROWS_WARPS + warp) * 32;
	if (y0 >= H) return;
	const TI* src = in + (long)blockIdx.y * in_frame_stride;
	TO* dst = out + (long)blockIdx.y * W * H;
	ACC(*t)[33] = tile[warp];
	const int rows = min(32, H - y0);
	ACC total = 0;
	// the tile after the current one is already in flight (registers) while the current one is scanned: with only
	// rows x frames / 32 warps in the grid nothing else hides the load latency
	ACC nxt[32];
	auto fetch = [&](int x0) {
		const int x = x0 + lane;
#pragma unroll
		for (int r = 0; r < 32; r++) {
			ACC v = 0;
			if (x < W && r < rows) v = (ACC)src[(long)(y0 + r) * in_stride + x];
			nxt[r] = v;
		}
	};
	fetch(0);
	for (int x0 = 0; x0 < W; x0 += 32) {
		const int x = x0 + lane;
		const bool xin = x < W;
#pragma unroll
		for (int r = 0; r < 32; r++) t[r][lane] = nxt[r];
		if (x0 + 32 < W) fetch(x0 + 32);
		__syncwarp();
		// lane = row: the reference's serial `total += in[x]`
#pragma unroll
		for (int k = 0; k < 32; k++) {
			total = (ACC)(total + t[lane][k]);
			t[lane][k] = total;
		}
		__syncwarp();
#pragma unroll
		for (int r = 0; r < 32; r++)
			if (xin && r < rows) dst[(long)(y0 + r) * W + x] = (TO)t[r][lane];
		__syncwarp();
	}
}

template <typename TO, typename ACC>
__global__ void __launch_bounds__(COL_THREADS)
integral_cols_kernel(TO* __restrict__ out, int W, int H) {
	const int x = blockIdx.x * COL_THREADS + threadIdx.x;
	if (x >= W) return;
	TO* p = out + (long)blockIdx.y * W * H + x;
	ACC acc = (ACC)p[0];      // first row is the plain running sum
	int y = 1;
	// software pipeline: the next COL_UNROLL rows are loaded while the current ones are accumulated and stored
	ACC cur[COL_UNROLL], nxt[COL_UNROLL];
	if (y + COL_UNROLL <= H) {
#pragma unroll
		for (int k = 0; k < COL_UNROLL; k++) cur[k] = (ACC)p[(long)(y + k) * W];
	}
	for (; y + COL_UNROLL <= H; y += COL_UNROLL) {
		const bool more = y + 2 * COL_UNROLL <= H;
		if (more) {
#pragma unroll
			for (int k = 0; k < COL_UNROLL; k++) nxt[k] = (ACC)p[(long)(y + COL_UNROLL + k) * W];
		}
#pragma unroll
		for (int k = 0; k < COL_UNROLL; k++) {
			acc = (ACC)(acc + cur[k]);      // out[y-1][x] + total
			p[(long)(y + k) * W] = (TO)acc;
		}
		if (more) {
#pragma unroll
			for (int k = 0; k < COL_UNROLL; k++) cur[k] = nxt[k];
		}
	}
	for (; y < H; y++) {
		acc = (ACC)(acc + (ACC)p[(long)y * W]);
		p[(long)y * W] = (TO)acc;
	}
}

template <typename T> struct AccOf { typedef uint32_t type; };
template <> struct AccOf<float> { typedef float type; };

// ---------------------------------------------------------------- parallel scan (integer images)
// Integer sums are exact under any association (Java int arithmetic wraps, uint32 here), so GrayU8 -> S32 and S32 -> S32 need
// not run the reference's serial order: the image is cut into bands of SCAN_ROWS rows, a CTA scans a band row by row —
// every thread owns V consecutive columns: local scan, warp scan of the thread totals by shuffles, warp totals through shared
// memory — and keeps the running column sums of its columns in registers.
//   pass A: each band's contribution to the rows below it = prefix over x of the band's column sums -> partial[frame][band][W]
//           (the rows are only added up: no scan, no barrier per row; reads the input once)
//   pass B: exclusive scan of the partials over the bands of a frame (tiny: 1 / SCAN_ROWS of the image)
//   pass C: the band scan, four rows per step, starting from the band's prefix, writes the integral   (input again + output)
// Traffic per pixel: 2 x input + 4 bytes out (+ 1/16 for the partials) against 16 bytes for the two serial passes above;
// the second read goes to DRAM again (launches that would keep it in L2 are too small to fill the GPU).
// Float images do NOT take this path: the same kernels instantiated for float ran at 6.25 vs 7.05 us per 1080p frame, but another
// association is another realisation of the float integral's rounding noise (sums reach 5e8 at 1080p: ulp 32) and 85 % of the
// descriptors moved by more than 1e-4 against the reference (tools/bench_integral.py history in DESIGN.md).
__device__ __forceinline__ int word_of(int32_t v) { return v; }
__device__ __forceinline__ int word_of(float v) { return __float_as_int(v); }
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ROWS = 32;

template <typename TI, typename ACC, int V>
__device__ __forceinline__ void scan_load_row(const TI* __restrict__ row, int W, int x0, ACC (&v)[V]) {
	// V consecutive pixels starting at x0; vector loads when the row pointer allows it, zero beyond W
	if (x0 + V <= W) {
		if (sizeof(TI) == 1 && V % 8 == 0 && (reinterpret_cast<uintptr_t>(row + x0) & 7) == 0) {
#pragma unroll
			for (int i = 0; i < V; i += 8) {
				const uint2 q = __ldg(reinterpret_cast<const uint2*>(row + x0 + i));
#pragma unroll
				for (int k = 0; k < 4; k++) { v[i + k] = (ACC)((q.x >> (8 * k)) & 0xffu); v[i + 4 + k] = (ACC)((q.y >> (8 * k)) & 0xffu); }
			}
			return;
		}
		if (sizeof(TI) == 4 && V % 4 == 0 && (reinterpret_cast<uintptr_t>(row + x0) & 15) == 0) {
#pragma unroll
			for (int i = 0; i < V; i += 4) {
				const int4 q = __ldg(reinterpret_cast<const int4*>(row + x0 + i));
				const TI* e = reinterpret_cast<const TI*>(&q);
#pragma unroll
				for (int k = 0; k < 4; k++) v[i + k] = (ACC)e[k];
			}
			return;
		}
	}
#pragma unroll
	for (int i = 0; i < V; i++) v[i] = x0 + i < W ? (ACC)__ldg(row + x0 + i) : (ACC)0;
}

// eight GrayU8 pixels as two words: one 8-byte load where the address allows it, byte loads (zero beyond W) otherwise; unpack8 widens them
__device__ __forceinline__ uint2 fetch8(const uint8_t* __restrict__ row, int W, int x0) {
	if (x0 + 8 <= W && (reinterpret_cast<uintptr_t>(row + x0) & 7) == 0) return __ldg(reinterpret_cast<const uint2*>(row + x0));
	uint2 q = make_uint2(0u, 0u);
#pragma unroll
	for (int k = 0; k < 4; k++) {
		if (x0 + k < W) q.x |= (uint32_t)__ldg(row + x0 + k) << (8 * k);
		if (x0 + 4 + k < W) q.y |= (uint32_t)__ldg(row + x0 + 4 + k) << (8 * k);
	}
	return q;
}
template <typename ACC>
__device__ __forceinline__ void unpack8(uint2 q, ACC (&v)[8]) {
#pragma unroll
	for (int k = 0; k < 4; k++) { v[k] = (ACC)((q.x >> (8 * k)) & 0xffu); v[4 + k] = (ACC)((q.y >> (8 * k)) & 0xffu); }
}

// CTA-wide inclusive scan of the thread totals t (thread = V consecutive columns): returns the exclusive prefix of this thread.
// `tot` = one row of shared memory of SCAN_THREADS / 32 words per concurrent scan; the caller separates uses by __syncthreads.
template <typename ACC>
__device__ __forceinline__ ACC scan_warp_inclusive(ACC t, int lane) {
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const ACC o = __shfl_up_sync(0xffffffffu, t, d);
		if (lane >= d) t = (ACC)(t + o);
	}
	return t;
}

// pass A: partial[frame][band][x] = sum over the band's rows y of the row prefix sum_{x' <= x} in[y][x'] — which is the prefix over x
// of the band's plain COLUMN sums: the rows are only added up (independent loads, no barrier), one scan across x per band.
template <typename TI, typename ACC, int V>
__global__ void __launch_bounds__(SCAN_THREADS)
integral_band_sums_kernel(const TI* __restrict__ in, long in_frame_stride, int in_stride, ACC* __restrict__ partial, int W, int H, int bands) {
	__shared__ ACC warp_tot[SCAN_THREADS / 32];
	const int band = blockIdx.x, frame = blockIdx.y;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int x0 = threadIdx.x * V;
	const int y0 = band * SCAN_ROWS, y1 = min(H, y0 + SCAN_ROWS);
	const TI* src = in + (long)frame * in_frame_stride;
	ACC sum[V];
#pragma unroll
	for (int i = 0; i < V; i++) sum[i] = (ACC)0;
	int y = y0;
	if constexpr (sizeof(TI) == 1 && V == 8) {
		// bytes: sixteen rows in flight as raw words (two registers per row), widened only when they are added
		for (; y + 16 <= y1; y += 16) {
			uint2 q[16];
#pragma unroll
			for (int k = 0; k < 16; k++) q[k] = fetch8(reinterpret_cast<const uint8_t*>(src) + (long)(y + k) * in_stride, W, x0);
#pragma unroll
			for (int k = 0; k < 16; k++) {
				ACC r[8];
				unpack8<ACC>(q[k], r);
#pragma unroll
				for (int i = 0; i < 8; i++) sum[i] = (ACC)(sum[i] + r[i]);
			}
		}
	}
	constexpr int RF = V <= 8 ? 4 : (V <= 16 ? 2 : 1);      // rows in flight (registers)
	for (; y + RF <= y1; y += RF) {
		ACC r[RF][V];
#pragma unroll
		for (int k = 0; k < RF; k++) scan_load_row<TI, ACC, V>(src + (long)(y + k) * in_stride, W, x0, r[k]);
#pragma unroll
		for (int k = 0; k < RF; k++)
#pragma unroll
			for (int i = 0; i < V; i++) sum[i] = (ACC)(sum[i] + r[k][i]);
	}
	for (; y < y1; y++) {
		ACC r0[V];
		scan_load_row<TI, ACC, V>(src + (long)y * in_stride, W, x0, r0);
#pragma unroll
		for (int i = 0; i < V; i++) sum[i] = (ACC)(sum[i] + r0[i]);
	}
#pragma unroll
	for (int i = 1; i < V; i++) sum[i] = (ACC)(sum[i] + sum[i - 1]);
	const ACC t = scan_warp_inclusive<ACC>(sum[V - 1], lane);
	if (lane == 31) warp_tot[warp] = t;
	__syncthreads();
	ACC base = (ACC)(t - sum[V - 1]);
#pragma unroll
	for (int w = 0; w < SCAN_THREADS / 32; w++) if (w < warp) base = (ACC)(base + warp_tot[w]);
	ACC* part = partial + ((long)frame * bands + band) * W;
#pragma unroll
	for (int i = 0; i < V; i++) if (x0 + i < W) part[x0 + i] = (ACC)(sum[i] + base);
}

// pass C: out = the band's prefix (pass B) + running sums of the row scans, R rows per step: R independent scans share one barrier
// and interleave their shuffles (one row per step left the CTA waiting on a 5-shuffle chain and a barrier per row).
template <typename TI, typename TO, typename ACC, int V, int R>
__global__ void __launch_bounds__(SCAN_THREADS)
integral_scan_kernel(const TI* __restrict__ in, long in_frame_stride, int in_stride, TO* __restrict__ out, const ACC* __restrict__ partial, int W, int H,
                     int bands) {
	__shared__ ACC warp_tot[2][R][SCAN_THREADS / 32];
	const int band = blockIdx.x, frame = blockIdx.y;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const int x0 = threadIdx.x * V;
	const int y0 = band * SCAN_ROWS, y1 = min(H, y0 + SCAN_ROWS);
	const TI* src = in + (long)frame * in_frame_stride;
	const ACC* part = partial + ((long)frame * bands + band) * W;
	ACC col[V];
#pragma unroll
	for (int i = 0; i < V; i++) col[i] = x0 + i < W ? part[x0 + i] : (ACC)0;
	int step = 0;
	constexpr bool BYTES = sizeof(TI) == 1 && V == 8;    // bytes: the next step's rows are fetched as raw words (two registers per row) before this step's scan
	uint2 nq[BYTES ? R : 1];
	if constexpr (BYTES) {
#pragma unroll
		for (int r = 0; r < R; r++) nq[r] = y0 + r < y1 ? fetch8(reinterpret_cast<const uint8_t*>(src) + (long)(y0 + r) * in_stride, W, x0) : make_uint2(0u, 0u);
	}
	for (int y = y0; y < y1; y += R, step++) {
		ACC cur[R][V];
		if constexpr (BYTES) {
			uint2 cq[R];
#pragma unroll
			for (int r = 0; r < R; r++) cq[r] = nq[r];
#pragma unroll
			for (int r = 0; r < R; r++)
				nq[r] = y + R + r < y1 ? fetch8(reinterpret_cast<const uint8_t*>(src) + (long)(y + R + r) * in_stride, W, x0) : make_uint2(0u, 0u);
#pragma unroll
			for (int r = 0; r < R; r++) {
				ACC w8[8];
				unpack8<ACC>(cq[r], w8);
#pragma unroll
				for (int i = 0; i < V; i++) cur[r][i] = w8[i < 8 ? i : 0];
			}
		} else {
#pragma unroll
			for (int r = 0; r < R; r++) {
				if (y + r < y1) scan_load_row<TI, ACC, V>(src + (long)(y + r) * in_stride, W, x0, cur[r]);
				else {
#pragma unroll
					for (int i = 0; i < V; i++) cur[r][i] = (ACC)0;
				}
			}
		}
		ACC t[R];
#pragma unroll
		for (int r = 0; r < R; r++) {
#pragma unroll
			for (int i = 1; i < V; i++) cur[r][i] = (ACC)(cur[r][i] + cur[r][i - 1]);
			t[r] = cur[r][V - 1];
		}
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
#pragma unroll
			for (int r = 0; r < R; r++) {
				const ACC o = __shfl_up_sync(0xffffffffu, t[r], d);
				if (lane >= d) t[r] = (ACC)(t[r] + o);
			}
		}
		const int b = step & 1;
		if (lane == 31) {
#pragma unroll
			for (int r = 0; r < R; r++) warp_tot[b][r][warp] = t[r];
		}
		__syncthreads();                               // one barrier per R rows: the totals alternate between two buffers
#pragma unroll
		for (int r = 0; r < R; r++) {
			ACC base = (ACC)(t[r] - cur[r][V - 1]);        // exclusive prefix inside the warp
#pragma unroll
			for (int w = 0; w < SCAN_THREADS / 32; w++) if (w < warp) base = (ACC)(base + warp_tot[b][r][w]);
#pragma unroll
			for (int i = 0; i < V; i++) col[i] = (ACC)(col[i] + (ACC)(cur[r][i] + base));
			if (y + r < y1) {
				TO* dst = out + ((long)frame * H + (y + r)) * W + x0;
				if (x0 + V <= W && V % 4 == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
					for (int i = 0; i < V; i += 4) {
						*reinterpret_cast<int4*>(dst + i) = make_int4(word_of((TO)col[i]), word_of((TO)col[i + 1]), word_of((TO)col[i + 2]), word_of((TO)col[i + 3]));
					}
				} else {
#pragma unroll
					for (int i = 0; i < V; i++) if (x0 + i < W) dst[i] = (TO)col[i];
				}
			}
		}
	}
}

// pass B: partial[frame][band][x] -> exclusive prefix over the bands (thread = column).  The loads of a column do not depend on each
// other: sixteen are issued before the first store (as a read-modify-write loop the compiler could not hoist them over the stores, and
// 34 dependent L2 round trips made this tiny pass as long as pass A: 60 us per 32 frames of 1080p).
template <typename ACC>
__global__ void integral_scan_bands_kernel(ACC* __restrict__ partial, int W, int bands) {
	const int x = blockIdx.x * blockDim.x + threadIdx.x;
	if (x >= W) return;
	ACC* p = partial + (long)blockIdx.y * bands * W + x;
	ACC run = 0;
	for (int b0 = 0; b0 < bands; b0 += 16) {
		ACC v[16];
#pragma unroll
		for (int k = 0; k < 16; k++) v[k] = b0 + k < bands ? p[(long)(b0 + k) * W] : (ACC)0;
#pragma unroll
		for (int k = 0; k < 16; k++) {
			if (b0 + k < bands) p[(long)(b0 + k) * W] = run;
			run = (ACC)(run + v[k]);
		}
	}
}

size_t integral_scan_scratch_bytes(int W, int H, int batch) { return (size_t)batch * div_up(H, SCAN_ROWS) * W * 4; }

// false if the shape is outside what the scan kernels are built for (W > 8192): the caller uses the serial passes
template <typename TI, typename TO>
bool launch_integral_scan(const TI* d_in, long in_frame_stride, int in_stride, TO* d_out, void* d_scratch, int W, int H, int batch, cudaStream_t st) {
	typedef typename AccOf<TO>::type ACC;
	if (W <= 0 || H <= 0 || batch <= 0) return true;
	if (W > SCAN_THREADS * 32) return false;
	const int bands = div_up(H, SCAN_ROWS);
	ACC* part = static_cast<ACC*>(d_scratch);
	// One launch per pass over the whole batch: bands x frames CTAs (1088 for 32 frames of 1080p).  Sub-batches small enough
	// for the input to stay in L2 between pass A and pass C (5 frames) left 170 CTAs per launch and ran at half the speed.
	const int sub = batch;
	for (int f0 = 0; f0 < batch; f0 += sub) {
		const int n = std::min(sub, batch - f0);
		const TI* in = d_in + (long)f0 * in_frame_stride;
		TO* out = d_out + (long)f0 * W * H;
		ACC* pp = part + (long)f0 * bands * W;
		dim3 g(bands, n);
		auto go = [&](auto vtag, auto rtag) {
			constexpr int V = decltype(vtag)::value, R = decltype(rtag)::value;
			BS_LAUNCH((integral_band_sums_kernel<TI, ACC, V>), g, SCAN_THREADS, 0, st, in, in_frame_stride, in_stride, pp, W, H, bands);
			BS_LAUNCH((integral_scan_bands_kernel<ACC>), dim3(div_up(W, 64), n), 64, 0, st, pp, W, bands);
			BS_LAUNCH((integral_scan_kernel<TI, TO, ACC, V, R>), g, SCAN_THREADS, 0, st, in, in_frame_stride, in_stride, out, pp, W, H, bands);
		};
		const int per = div_up(W, SCAN_THREADS);
		if (per <= 4) go(std::integral_constant<int, 4>(), std::integral_constant<int, 4>());
		else if (per <= 8) go(std::integral_constant<int, 8>(), std::integral_constant<int, 4>());
		else if (per <= 16) go(std::integral_constant<int, 16>(), std::integral_constant<int, 2>());
		else go(std::integral_constant<int, 32>(), std::integral_constant<int, 1>());
	}
	return true;
}


template <typename TI, typename TO>
void launch_integral(const TI* d_in, long in_frame_stride, int in_stride, TO* d_out, int W, int H, int batch, cudaStream_t st) {
	if (W <= 0 || H <= 0 || batch <= 0) return;
	typedef typename AccOf<TO>::type ACC;
	dim3 g1(div_up(H, 32 * ROWS_WARPS), batch);
	BS_LAUNCH((integral_rows_kernel<TI, TO, ACC>), g1, ROWS_WARPS * 32, 0, st, d_in, in_frame_stride, in_stride, d_out, W, H);
	dim3 g2(div_up(W, COL_THREADS), batch);
	BS_LAUNCH((integral_cols_kernel<TO, ACC>), g2, COL_THREADS, 0, st, d_out, W, H);
}

template bool launch_integral_scan<float, float>(const float*, long, int, float*, void*, int, int, int, cudaStream_t);
template bool launch_integral_scan<uint8_t, int32_t>(const uint8_t*, long, int, int32_t*, void*, int, int, int, cudaStream_t);
template bool launch_integral_scan<int32_t, int32_t>(const int32_t*, long, int, int32_t*, void*, int, int, int, cudaStream_t);
template void launch_integral<float, float>(const float*, long, int, float*, int, int, int, cudaStream_t);
template void launch_integral<uint8_t, int32_t>(const uint8_t*, long, int, int32_t*, int, int, int, cudaStream_t);
template void launch_integral<int32_t, int32_t>(const int32_t*, long, int, int32_t*, int, int, int, cudaStream_t);

}  // namespace bs
