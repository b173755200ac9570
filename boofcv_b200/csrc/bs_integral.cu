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

template <typename TI, typename TO>
void launch_integral(const TI* d_in, long in_frame_stride, int in_stride, TO* d_out, int W, int H, int batch, cudaStream_t st) {
	if (W <= 0 || H <= 0 || batch <= 0) return;
	typedef typename AccOf<TO>::type ACC;
	dim3 g1(div_up(H, 32 * ROWS_WARPS), batch);
	BS_LAUNCH((integral_rows_kernel<TI, TO, ACC>), g1, ROWS_WARPS * 32, 0, st, d_in, in_frame_stride, in_stride, d_out, W, H);
	dim3 g2(div_up(W, COL_THREADS), batch);
	BS_LAUNCH((integral_cols_kernel<TO, ACC>), g2, COL_THREADS, 0, st, d_out, W, H);
}

template void launch_integral<float, float>(const float*, long, int, float*, int, int, int, cudaStream_t);
template void launch_integral<uint8_t, int32_t>(const uint8_t*, long, int, int32_t*, int, int, int, cudaStream_t);
template void launch_integral<int32_t, int32_t>(const int32_t*, long, int, int32_t*, int, int, int, cudaStream_t);

}  // namespace bs
