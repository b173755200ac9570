// bs_describe.cu — orientation estimate and SURF descriptor, one warp per feature.
// Reference (per feature, WrapDetectDescribeSurf.computeDescriptors, F/abst/feature/detdesc/WrapDetectDescribeSurf.java:98-110):
//   orientation.setObjectRadius(scale*2); angle = orientation.compute(x,y)
//     OrientationIntegralBase (F/alg/feature/orientation/OrientationIntegralBase.java:75-103)
//     ImplOrientationAverageGradientIntegral.compute (…/impl/ImplOrientationAverageGradientIntegral.java:54-127)
//     ImplOrientationSlidingWindowIntegral.compute (…/impl/ImplOrientationSlidingWindowIntegral.java:81-188)
//   describe.describe(x,y,angle,scale,feature)
//     DescribePointSurf.describe/features/computeLaplaceSign (F/alg/feature/describe/DescribePointSurf.java:169-313)
//     DescribePointSurfMod.features (F/alg/feature/describe/DescribePointSurfMod.java:121-192)
//     SurfDescribeOps.isInside (F/alg/feature/describe/SurfDescribeOps.java:119-159,183-204)
//     UtilFeature.normalizeL2 (F/alg/descriptor/UtilFeature.java:101-114)
//
// Work split inside a warp: the gather-heavy part (12 integral reads per sample, 169..576 samples) is spread over the 32
// lanes with no ordering constraint; every floating-point SUM the reference makes is then redone by a single lane in the
// reference's order (fp64 adds are not associative and the angle feeds (int) truncations downstream), so the results
// are bit-identical to a serial evaluation.
#include "bs_common.cuh"
#include "bs_box.cuh"
#include "bs_math.cuh"
#include <climits>
#include <cstdlib>

namespace bs {

constexpr int DESC_WARPS = 8;

// SurfDescribeOps.isInside (:119-159,183-204) only selects between the reference's unchecked gradient operator and its
// bounds-checked wrapper (which yields (0,0) outside).  Whenever isInside holds every sample passes the bounds check, so
// always using the checked operator (bs_box.cuh sparse_gradient) gives the reference's values on both branches.

// georegression UtilAngle.dist(a,b): shortest angular distance
__device__ __forceinline__ double angle_dist(double a, double b) {
	const double diff = fabs(a - b);
	if (diff > 3.14159265358979323846) return 2.0 * 3.14159265358979323846 - diff;
	return diff;
}

// ---------------------------------------------------------------- staged integral window
// ncu on the one-warp-per-feature version showed the L1 data pipe as the limiter: 12 scattered 4-byte reads per sample, ~21
// sectors and ~6.7 wavefronts per warp-level request.  The samples of one feature all fall in a window of (18..30)*scale
// pixels, so the CTA copies that window of the integral image into shared memory with coalesced cp.async transfers and
// the scattered reads become LDS.  Windows that do not fit in the tile (large scales) keep reading global memory.
struct Window {                    // pixels [x0, x0+bw) x [y0, y0+bh) of the frame, staged with row pitch `pitch`
	int x0, y0, bw, bh, pitch;     // bw == 0: not staged
};

struct Slot {                      // one feature being processed by warp `w` of the CTA (shared memory)
	double x, y, scale, angle, c, s;
	long f;                        // pool index, -1: none
	int frame;                     // frame index relative to the first frame of the launch
	Window ori, desc;
};

__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
	asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Clamp [lo, hi] (inclusive pixel range, any sign) to the frame and lay it out; `vec` = rows are copied 16 bytes at a time
// (needs W % 4 == 0, then x0 and bw are multiples of 4 as well).  Returns a window with bw == 0 if it does not fit.
__device__ __forceinline__ Window make_window(int lox, int loy, int hix, int hiy, int W, int H, bool vec, int tile_words) {
	Window w{0, 0, 0, 0, 0};
	lox = max(lox, 0); loy = max(loy, 0); hix = min(hix, W - 1); hiy = min(hiy, H - 1);
	if (hix < lox || hiy < loy) return w;
	if (vec) { lox &= ~3; hix |= 3; }
	const int bw = hix - lox + 1, bh = hiy - loy + 1;
	// pitch: multiple of 4 words with pitch/4 odd (vec) or odd (scalar), so that rows start in different banks
	const int pitch = vec ? (bw + (((bw >> 2) & 1) ? 0 : 4)) : (bw | 1);
	if ((long)pitch * bh > tile_words) return w;
	w.x0 = lox; w.y0 = loy; w.bw = bw; w.bh = bh; w.pitch = pitch;
	return w;
}

// all `nt` threads of the CTA: start the copy of window `w` of frame `ii` into `tile`
template <typename T>
__device__ __forceinline__ void stage_window(T* tile, const T* __restrict__ ii, int W, const Window& w, bool vec, int tid, int nt) {
	if (w.bw == 0) return;
	if (vec) {
		const int bw4 = w.bw >> 2, n = bw4 * w.bh;
		const float inv = 1.0f / (float)bw4;
		for (int idx = tid; idx < n; idx += nt) {
			const int row = (int)(((float)idx + 0.5f) * inv), col = (idx - row * bw4) << 2;
			cp_async16(tile + row * w.pitch + col, ii + (long)(w.y0 + row) * W + w.x0 + col);
		}
	} else {
		const int n = w.bw * w.bh;
		const float inv = 1.0f / (float)w.bw;
		for (int idx = tid; idx < n; idx += nt) {
			const int row = (int)(((float)idx + 0.5f) * inv), col = idx - row * w.bw;
			cp_async4(tile + row * w.pitch + col, ii + (long)(w.y0 + row) * W + w.x0 + col);
		}
	}
}

// SparseIntegralGradient_NoBorder.compute with `a` = address of pixel (x-r-1, y-r-1) in a buffer of row pitch `pitch`
// (bs_box.cuh sparse_gradient is the same arithmetic on the global image)
template <typename T, bool GLOBAL>
__device__ __forceinline__ void gradient_from(const T* __restrict__ a, int pitch, int r, T& dx, T& dy) {
	const int w = 2 * r + 1;
	const T* b = a + r * pitch;
	const T* c = b + pitch;
	const T* d = c + r * pitch;
#define BS_RD(p) (GLOBAL ? __ldg(p) : *(p))
	const T p0 = BS_RD(a), p1 = BS_RD(a + r), p2 = BS_RD(a + r + 1), p3 = BS_RD(a + w);
	const T p11 = BS_RD(b), p4 = BS_RD(b + w);
	const T p10 = BS_RD(c), p5 = BS_RD(c + w);
	const T p9 = BS_RD(d), p8 = BS_RD(d + r), p7 = BS_RD(d + r + 1), p6 = BS_RD(d + w);
#undef BS_RD
	const T left = Arith<T>::sum4(p8, p9, p1, p0);
	const T right = Arith<T>::sum4(p6, p7, p3, p2);
	const T top = Arith<T>::sum4(p4, p11, p3, p0);
	const T bottom = Arith<T>::sum4(p6, p9, p5, p10);
	dx = Arith<T>::sub(right, left);
	dy = Arith<T>::sub(bottom, top);
}

// gradient at (x,y), kernel radius r: (0,0) when the kernel leaves the image (SparseGradientSafe), else from the staged
// window if there is one, else from global memory
template <typename T>
__device__ __forceinline__ void gradient_at(const T* __restrict__ ii, int W, int H, const T* tile, const Window& win, int r, int x, int y,
                                            T& dx, T& dy) {
	if (x - r - 1 < 0 || y - r - 1 < 0 || x + r >= W || y + r >= H) { dx = 0; dy = 0; return; }
	if (win.bw) gradient_from<T, false>(tile + (y - r - 1 - win.y0) * win.pitch + (x - r - 1 - win.x0), win.pitch, r, dx, dy);
	else gradient_from<T, true>(ii + (long)(y - r - 1) * W + (x - r - 1), W, r, dx, dy);
}

// ---------------------------------------------------------------- orientation
// geometry shared by the window computation and the sampling loop (same expressions => same pixels)
struct OriGeo {
	int r, sw;
	double period, tl_x, tl_y;
};
__device__ __forceinline__ OriGeo ori_geometry(const DescribeParams& P, double c_x, double c_y, double ptScale) {
	OriGeo g;
	g.sw = 2 * P.ori_radius + 1;
	// setObjectRadius(scale*BoofDefaults.SURF_SCALE_TO_RADIUS): scale = radius*objectRadiusToScale; kernel width = scale*sampleWidth
	const double sc = (ptScale * 2.0) * P.ori_radius_to_scale;
	g.r = gradient_radius(sc * P.ori_sample_width);
	g.period = sc * P.ori_period;
	g.tl_x = c_x - P.ori_radius * g.period;
	g.tl_y = c_y - P.ori_radius * g.period;
	g.tl_x += 0.5; g.tl_y += 0.5;
	return g;
}
// pixels the orientation samples can touch: sample coordinates are monotone in the grid index, so the first and last
// index bound them (one pixel of slack on each side)
__device__ __forceinline__ Window ori_window(const DescribeParams& P, double x, double y, double scale, int W, int H, bool vec, int tile_words) {
	const OriGeo g = ori_geometry(P, x, y, scale);
	const int lox = (int)(g.tl_x + 0 * g.period), hix = (int)(g.tl_x + (g.sw - 1) * g.period);
	const int loy = (int)(g.tl_y + 0 * g.period), hiy = (int)(g.tl_y + (g.sw - 1) * g.period);
	return make_window(lox - g.r - 2, loy - g.r - 2, hix + g.r + 1, hiy + g.r + 1, W, H, vec, tile_words);
}

// sampling, all `nt` threads of the CTA for one feature.  scratch: dX[n], dY[n] (+ ang[n] for the sliding window)
template <typename T>
__device__ __forceinline__ void orientation_sample(const T* __restrict__ ii, int W, int H, const T* tile, const Window& win,
                                                   const DescribeParams& P, const double* __restrict__ wts, double c_x, double c_y,
                                                   double ptScale, double* dX, double* dY, double* ang, int tid, int nt) {
	const OriGeo g = ori_geometry(P, c_x, c_y, ptScale);
	const int n = g.sw * g.sw;
	for (int k = tid; k < n; k += nt) {
		const int y = k / g.sw, x = k - y * g.sw;
		const int py = (int)(g.tl_y + y * g.period);
		const int px = (int)(g.tl_x + x * g.period);
		T gdx, gdy;
		gradient_at(ii, W, H, tile, win, g.r, px, py, gdx, gdy);
		double dx = (double)gdx, dy = (double)gdy;
		if (P.ori_weighted) { const double w = wts[k]; dx = w * dx; dy = w * dy; }
		dX[k] = dx; dY[k] = dy;
		if (P.ori_variant == 1) ang[k] = sm_atan2(dy, dx);
	}
}

// reduction by the feature's own warp, in the reference's order
__device__ double orientation_reduce(const DescribeParams& P, const double* dX, const double* dY, const double* ang, int* order) {
	const int lane = threadIdx.x & 31;
	const int sw = 2 * P.ori_radius + 1, n = sw * sw;
	double angle = 0.0;
	if (P.ori_variant == 0) {
		// computeUnweighted/computeWeighted: Dx += w*dx in sample order; lane 0 sums x, lane 1 sums y
		double total = 0.0;
		if (lane < 2) {
			const double* src = lane == 0 ? dX : dY;
			for (int k = 0; k < n; k++) total += src[k];
		}
		const double Dx = __shfl_sync(0xffffffffu, total, 0);
		const double Dy = __shfl_sync(0xffffffffu, total, 1);
		angle = sm_atan2(Dy, Dx);
	} else {
		// ascending argsort, ties towards the lower index (rank by counting: the key (angle,index) is a total order)
		for (int i = lane; i < n; i += 32) {
			const double a = ang[i];
			int rank = 0;
			for (int j = 0; j < n; j++) {
				const double b = ang[j];
				rank += (b < a || (b == a && j < i)) ? 1 : 0;
			}
			order[rank] = i;
		}
		__syncwarp();
		if (lane == 0) {
			// estimateAngle :139-188
			int start = 0, end = 1;
			int startIndex = order[start], endIndex = order[end];
			double sumX = dX[startIndex], sumY = dY[startIndex];
			double best = sumX * sumX + sumY * sumY, bestX = sumX, bestY = sumY;
			double endAngle = ang[endIndex];
			while (start != n) {
				startIndex = order[start];
				const double startAngle = ang[startIndex];
				while (angle_dist(startAngle, endAngle) <= P.ori_window) {
					sumX += dX[endIndex];
					sumY += dY[endIndex];
					const double mag = sumX * sumX + sumY * sumY;
					if (mag > best) { best = mag; bestX = sumX; bestY = sumY; }
					end++;
					if (end >= n) end = 0;
					endIndex = order[end];
					endAngle = ang[endIndex];
					if (endIndex == startIndex) break;
				}
				sumX -= dX[startIndex];
				sumY -= dY[startIndex];
				start++;
			}
			angle = sm_atan2(bestY, bestX);
		}
		angle = __shfl_sync(0xffffffffu, angle, 0);
	}
	__syncwarp();
	return angle;
}

// ---------------------------------------------------------------- descriptor
// DerivativeIntegralImage.kernelDerivXX/YY geometry for a kernel of `size`
template <typename T> struct LapAcc;
template <> struct LapAcc<float> {
	// convolveSparse F32: ret = 0; ret += block_zero*1; ret += block_zero*(-3)   (float)
	static __device__ __forceinline__ double conv(const float* ii, int W, int H, int x, int y, int rx_big, int rx_small, int ry, bool transpose) {
		float ret = 0.0f;
		if (!transpose) {
			ret += box_zero(ii, W, H, x - rx_big - 1, y - ry - 1, x + rx_big, y + ry) * 1.0f;
			ret += box_zero(ii, W, H, x - rx_small - 1, y - ry - 1, x + rx_small, y + ry) * -3.0f;
		} else {
			ret += box_zero(ii, W, H, x - ry - 1, y - rx_big - 1, x + ry, y + rx_big) * 1.0f;
			ret += box_zero(ii, W, H, x - ry - 1, y - rx_small - 1, x + ry, y + rx_small) * -3.0f;
		}
		return (double)ret;
	}
};
template <> struct LapAcc<int32_t> {
	static __device__ __forceinline__ double conv(const int32_t* ii, int W, int H, int x, int y, int rx_big, int rx_small, int ry, bool transpose) {
		uint32_t ret = 0;
		if (!transpose) {
			ret += (uint32_t)box_zero(ii, W, H, x - rx_big - 1, y - ry - 1, x + rx_big, y + ry);
			ret += (uint32_t)box_zero(ii, W, H, x - rx_small - 1, y - ry - 1, x + rx_small, y + ry) * (uint32_t)(-3);
		} else {
			ret += (uint32_t)box_zero(ii, W, H, x - ry - 1, y - rx_big - 1, x + ry, y + rx_big);
			ret += (uint32_t)box_zero(ii, W, H, x - ry - 1, y - rx_small - 1, x + ry, y + rx_small) * (uint32_t)(-3);
		}
		return (double)(int32_t)ret;
	}
};

// computeLaplaceSign (DescribePointSurf.java:305-313)
template <typename T>
__device__ __forceinline__ bool laplace_sign(const T* __restrict__ ii, int W, int H, int x, int y, double scale) {
	const int size = 9 * (int)ceil(scale);
	const int blockW = size / 3, blockH = size - blockW - 1;
	const int r1 = blockW / 2, r2 = blockW + r1, r3 = blockH / 2;
	double lap = LapAcc<T>::conv(ii, W, H, x, y, r2, r1, r3, false);
	lap += LapAcc<T>::conv(ii, W, H, x, y, r2, r1, r3, true);
	return lap > 0;
}

// sample grid of the descriptor (DescribePointSurfMod samples the overlap once)
struct DescGeo {
	int r, regionR, ov, G;
	double c_x, c_y;
};
__device__ __forceinline__ DescGeo desc_geometry(const DescribeParams& P, double x, double y, double scale) {
	DescGeo g;
	g.r = gradient_radius(P.sample_width * scale);
	const int regionSize = P.grid * P.sub;
	g.regionR = regionSize / 2;
	g.ov = P.variant == 1 ? P.overlap : 0;
	g.G = regionSize + 2 * g.ov;
	g.c_x = x + 0.5; g.c_y = y + 0.5;
	return g;
}
__device__ __forceinline__ void desc_pixel(const DescGeo& g, double c, double s, double scale, int gxi, int gyi, int& pixelX, int& pixelY) {
	const double regionY = (gyi - g.regionR - g.ov) * scale;
	const double regionX = (gxi - g.regionR - g.ov) * scale;
	pixelX = (int)(g.c_x + c * regionX - s * regionY);
	pixelY = (int)(g.c_y + s * regionX + c * regionY);
}
// pixels the descriptor samples can touch: desc_pixel is monotone in gxi for fixed gyi and in gyi for fixed gxi, so the four
// corners of the grid bound it (one pixel of slack on each side)
__device__ __forceinline__ Window desc_window(const DescribeParams& P, double x, double y, double scale, double c, double s, int W, int H,
                                              bool vec, int tile_words) {
	const DescGeo g = desc_geometry(P, x, y, scale);
	int lox = INT_MAX, loy = INT_MAX, hix = INT_MIN, hiy = INT_MIN;
#pragma unroll
	for (int q = 0; q < 4; q++) {
		int px, py;
		desc_pixel(g, c, s, scale, (q & 1) ? g.G - 1 : 0, (q & 2) ? g.G - 1 : 0, px, py);
		lox = min(lox, px); hix = max(hix, px); loy = min(loy, py); hiy = max(hiy, py);
	}
	return make_window(lox - g.r - 2, loy - g.r - 2, hix + g.r + 1, hiy + g.r + 1, W, H, vec, tile_words);
}

// sampling, all `nt` threads of the CTA for one feature.  gx/gy: G*G raw gradients in the integral type (widened to double
// where the reference does)
template <typename T>
__device__ __forceinline__ void describe_sample(const T* __restrict__ ii, int W, int H, const T* tile, const Window& win,
                                                const DescribeParams& P, double x, double y, double c, double s, double scale, T* gx, T* gy,
                                                int tid, int nt) {
	const DescGeo g = desc_geometry(P, x, y, scale);
	const int n = g.G * g.G;
	for (int k = tid; k < n; k += nt) {
		const int gyi = k / g.G, gxi = k - gyi * g.G;
		int pixelX, pixelY;
		desc_pixel(g, c, s, scale, gxi, gyi, pixelX, pixelY);
		T dx, dy;
		gradient_at(ii, W, H, tile, win, g.r, pixelX, pixelY, dx, dy);
		gx[k] = dx; gy[k] = dy;
	}
}

// sums, normalisation and Laplacian sign by the feature's own warp.  feat: dof doubles
template <typename T>
__device__ void describe_reduce(const T* __restrict__ ii, int W, int H, const DescribeParams& P, const double* __restrict__ wt,
                                const double* __restrict__ wgrid, double x, double y, double c, double s, double scale, const T* gx,
                                const T* gy, double* feat, double* out_desc, uint8_t* out_white) {
	const int lane = threadIdx.x & 31;
	const int regionSize = P.grid * P.sub;
	const int ov = P.variant == 1 ? P.overlap : 0;
	const int G = regionSize + 2 * ov;
	// one sub-region per lane pair, summed in the reference's (i,j) order: even lane -> (sum_dx, sum_adx), odd -> (sum_dy, sum_ady)
	const int ww = P.sub + 2 * ov;               // window summed per sub-region
	const int nsub = P.grid * P.grid;
	for (int q = lane; q < 2 * nsub; q += 32) {
		const int region = q >> 1, which = q & 1;
		const int ry = region / P.grid, rx = region - ry * P.grid;
		double sum = 0.0, asum = 0.0;
		for (int i = 0; i < ww; i++) {
			const int row = ry * P.sub + i;
			for (int j = 0; j < ww; j++) {
				const int col = rx * P.sub + j;
				const int k = row * G + col;
				const double w = P.variant == 0 ? wt[row * regionSize + col] : wt[i * ww + j];
				const double dx = w * (double)gx[k], dy = w * (double)gy[k];
				const double pd = which == 0 ? (c * dx + s * dy) : (-s * dx + c * dy);
				sum += pd;
				asum += fabs(pd);
			}
		}
		if (P.variant == 1) { const double w = wgrid[region]; sum = w * sum; asum = w * asum; }
		feat[region * 4 + which * 2] = sum;
		feat[region * 4 + which * 2 + 1] = asum;
	}
	__syncwarp();
	// UtilFeature.normalizeL2: every lane repeats the same serial sum (shared-memory broadcast reads)
	double norm = 0.0;
	for (int k = 0; k < P.dof; k++) { const double t = feat[k]; norm += t * t; }
	const bool scale_it = norm != 0.0;
	norm = sqrt(norm);
	for (int k = lane; k < P.dof; k += 32) out_desc[k] = scale_it ? feat[k] / norm : feat[k];
	if (lane == 0) *out_white = laplace_sign(ii, W, H, (int)(x + 0.5), (int)(y + 0.5), scale) ? 1 : 0;
	__syncwarp();
}

// ---------------------------------------------------------------- kernel
// A CTA of DESC_WARPS warps takes DESC_WARPS consecutive features per round.  Sampling is done feature after feature by the
// whole CTA out of a double-buffered shared-memory tile (the next window is in flight while the current one is sampled);
// the order-sensitive reductions are done by one warp per feature, all DESC_WARPS features at once.
struct DescribeSmem {
	int ori_n, desc_g2;
	int tab_ori, tab_w, tab_wgrid;     // table sizes in doubles
	int per_warp;                      // bytes of scratch per feature slot
	int tile_words;                    // capacity of one tile
	int total;                         // bytes
};
constexpr int DESC_TILES = 2;

static DescribeSmem describe_smem(const DescribeParams& p, int tile_words) {
	DescribeSmem s{};
	const int sw = 2 * p.ori_radius + 1;
	s.ori_n = sw * sw;
	const int region = p.grid * p.sub, ov = p.variant == 1 ? p.overlap : 0;
	const int G = region + 2 * ov;
	s.desc_g2 = G * G;
	s.tab_ori = p.ori_weighted ? s.ori_n : 0;
	s.tab_w = p.variant == 0 ? region * region : (p.sub + 2 * ov) * (p.sub + 2 * ov);
	s.tab_wgrid = p.variant == 1 ? p.grid * p.grid : 0;
	const int ori_bytes = s.ori_n * (p.ori_variant == 1 ? 8 * 3 + 4 : 8 * 2);
	const int desc_bytes = p.dof * 8 + s.desc_g2 * 4 * 2;
	s.per_warp = ((ori_bytes > desc_bytes ? ori_bytes : desc_bytes) + 15) & ~15;
	s.tile_words = tile_words;
	s.total = (s.tab_ori + s.tab_w + s.tab_wgrid) * 8 + DESC_WARPS * ((int)sizeof(Slot) + s.per_warp) + DESC_TILES * tile_words * 4;
	s.total = (s.total + 15) & ~15;
	return s;
}

template <typename T>
__global__ void __launch_bounds__(DESC_WARPS * 32)
describe_kernel(const T* __restrict__ ii_all, int W, int H, DescribeParams P, DescribeTables tabs, DescribeSmem S, Pool pool,
                const int* __restrict__ d_begin, const int* __restrict__ d_end, int frame0, int do_orientation, int do_describe) {
	extern __shared__ __align__(16) unsigned char smem[];
	constexpr int NT = DESC_WARPS * 32;
	// layout: tiles (16-byte aligned) | tables | slots | per-slot scratch
	T* tiles = reinterpret_cast<T*>(smem);
	double* t_ori = reinterpret_cast<double*>(smem + (size_t)DESC_TILES * S.tile_words * 4);
	double* t_w = t_ori + S.tab_ori;
	double* t_wgrid = t_w + S.tab_w;
	Slot* slots = reinterpret_cast<Slot*>(t_wgrid + S.tab_wgrid);
	unsigned char* scratch = reinterpret_cast<unsigned char*>(slots + DESC_WARPS);
	const int tid = threadIdx.x;
	for (int k = tid; k < S.tab_ori; k += NT) t_ori[k] = tabs.d_ori_w[k];
	for (int k = tid; k < S.tab_w; k += NT) t_w[k] = tabs.d_w[k];
	for (int k = tid; k < S.tab_wgrid; k += NT) t_wgrid[k] = tabs.d_wgrid[k];
	const int warp = tid >> 5, lane = tid & 31;
	const bool vec = (W & 3) == 0;
	Slot& me = slots[warp];
	unsigned char* mine = scratch + (size_t)warp * S.per_warp;

	long end = *d_end;
	if (end > pool.cap) end = pool.cap;
	for (long base = *d_begin + (long)blockIdx.x * DESC_WARPS; base < end; base += (long)gridDim.x * DESC_WARPS) {
		__syncthreads();                       // previous round (and the table fill) done with the slots
		if (lane == 0) {
			const long f = base + warp;
			me.f = f < end ? f : -1;
			if (f < end) {
				me.x = pool.xy[2 * f]; me.y = pool.xy[2 * f + 1]; me.scale = pool.scale[f];
				me.frame = pool.frame[f] - frame0;
				if (do_orientation) me.ori = ori_window(P, me.x, me.y, me.scale, W, H, vec, S.tile_words);
				else me.angle = pool.angle[f];
			}
		}
		__syncthreads();
		if (do_orientation) {
			if (slots[0].f >= 0) stage_window(tiles, ii_all + (long)slots[0].frame * W * H, W, slots[0].ori, vec, tid, NT);
			for (int j = 0; j < DESC_WARPS; j++) {
				cp_async_wait_all();
				__syncthreads();               // window j landed; everybody is done sampling window j-1
				if (j + 1 < DESC_WARPS && slots[j + 1].f >= 0)
					stage_window(tiles + (size_t)((j + 1) & 1) * S.tile_words, ii_all + (long)slots[j + 1].frame * W * H, W, slots[j + 1].ori,
					             vec, tid, NT);
				const Slot& sj = slots[j];
				if (sj.f >= 0) {
					double* dX = reinterpret_cast<double*>(scratch + (size_t)j * S.per_warp);
					double* dY = dX + S.ori_n;
					orientation_sample(ii_all + (long)sj.frame * W * H, W, H, tiles + (size_t)(j & 1) * S.tile_words, sj.ori, P, t_ori, sj.x, sj.y,
					                   sj.scale, dX, dY, dY + S.ori_n, tid, NT);
				}
			}
			__syncthreads();
			if (me.f >= 0) {
				double* dX = reinterpret_cast<double*>(mine);
				double* dY = dX + S.ori_n;
				double* ang = dY + S.ori_n;
				const double angle = orientation_reduce(P, dX, dY, ang, reinterpret_cast<int*>(ang + S.ori_n));
				if (lane == 0) { me.angle = angle; pool.angle[me.f] = angle; }
			}
			__syncwarp();
		}
		if (do_describe) {
			if (lane == 0 && me.f >= 0) {
				me.c = sm_cos(me.angle); me.s = sm_sin(me.angle);
				me.desc = desc_window(P, me.x, me.y, me.scale, me.c, me.s, W, H, vec, S.tile_words);
			}
			__syncthreads();                   // every slot has its angle and window; orientation scratch is free
			if (slots[0].f >= 0) stage_window(tiles, ii_all + (long)slots[0].frame * W * H, W, slots[0].desc, vec, tid, NT);
			for (int j = 0; j < DESC_WARPS; j++) {
				cp_async_wait_all();
				__syncthreads();
				if (j + 1 < DESC_WARPS && slots[j + 1].f >= 0)
					stage_window(tiles + (size_t)((j + 1) & 1) * S.tile_words, ii_all + (long)slots[j + 1].frame * W * H, W, slots[j + 1].desc,
					             vec, tid, NT);
				const Slot& sj = slots[j];
				if (sj.f >= 0) {
					T* gx = reinterpret_cast<T*>(scratch + (size_t)j * S.per_warp + (size_t)P.dof * 8);
					describe_sample(ii_all + (long)sj.frame * W * H, W, H, tiles + (size_t)(j & 1) * S.tile_words, sj.desc, P, sj.x, sj.y, sj.c,
					                sj.s, sj.scale, gx, gx + S.desc_g2, tid, NT);
				}
			}
			__syncthreads();
			if (me.f >= 0) {
				double* feat = reinterpret_cast<double*>(mine);
				const T* gx = reinterpret_cast<const T*>(feat + P.dof);
				describe_reduce(ii_all + (long)me.frame * W * H, W, H, P, t_w, t_wgrid, me.x, me.y, me.c, me.s, me.scale, gx, gx + S.desc_g2, feat,
				                pool.desc + me.f * P.dof, pool.white + me.f);
			}
		}
	}
}

static int describe_tile_words() {
	static int words = 0;
	if (words == 0) {
		words = 9216;                      // 36 KB per tile: a 96 x 96 pixel window
		if (const char* e = getenv("BS_DESC_TILE_WORDS")) { const int v = atoi(e); if (v >= 0 && v <= 24576) words = v > 0 ? (v & ~3) : 4; }
	}
	return words;
}

template <typename T>
void launch_describe(const T* d_ii, int W, int H, const DescribeParams& p, const DescribeTables& t, const Pool& pool,
                     const int* d_begin, const int* d_end, int frame0, bool do_orientation, bool do_describe, cudaStream_t st) {
	static int sm_count = 0;
	if (sm_count == 0) {
		int dev = 0;
		BS_CUDA(cudaGetDevice(&dev));
		BS_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
	}
	DescribeSmem S = describe_smem(p, describe_tile_words());
	if (S.total > 220 * 1024) S = describe_smem(p, 4);      // huge sampling grids: no staging, scratch only
	if (S.total > 220 * 1024) fail(BS_ERR_ARG, "orientation/descriptor configuration needs too much shared memory");
	BS_CUDA(cudaFuncSetAttribute(describe_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, S.total));
	int ctas_per_sm = (224 * 1024) / (S.total + 1024);
	ctas_per_sm = ctas_per_sm > 4 ? 4 : (ctas_per_sm < 1 ? 1 : ctas_per_sm);
	BS_LAUNCH((describe_kernel<T>), sm_count * ctas_per_sm, DESC_WARPS * 32, S.total, st, d_ii, W, H, p, t, S, pool, d_begin, d_end, frame0,
	          do_orientation ? 1 : 0, do_describe ? 1 : 0);
}
template void launch_describe<float>(const float*, int, int, const DescribeParams&, const DescribeTables&, const Pool&, const int*, const int*,
                                     int, bool, bool, cudaStream_t);
template void launch_describe<int32_t>(const int32_t*, int, int, const DescribeParams&, const DescribeTables&, const Pool&, const int*,
                                       const int*, int, bool, bool, cudaStream_t);

}  // namespace bs
