// bs_describe.cu — orientation estimate and SURF descriptor.
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
// Work split: the gather-heavy part (12 integral reads per sample, 169..576 samples) is spread over the threads of the CTA
// with no ordering constraint; every floating-point SUM the reference makes is then redone by a single lane in the
// reference's order (fp64 adds are not associative and the angle feeds (int) truncations downstream), so the results
// are bit-identical to a serial evaluation.
#include "bs_common.cuh"
#include "bs_box.cuh"
#include "bs_math.cuh"
#include <climits>
#include <math_constants.h>
#include <cstdlib>

namespace bs {

#ifndef BS_DESC_WARPS
#define BS_DESC_WARPS 4
#endif
// features per CTA round (<= 16: the batched orientation sums use two lanes per feature).  Measured per 256 frames (1080p):
// 4 warps x 8 CTAs/SM 24.0 ms, 8 x 4 24.8 ms, 16 x 2 34.0 ms, 2 x 16 26.4 ms; fewer registers for more resident warps (48
// registers, 40 warps) 43 ms: the kernel is bound by the L1 data pipe and more warps only thrash it.
constexpr int DESC_WARPS = BS_DESC_WARPS;

// SurfDescribeOps.isInside (:119-159,183-204) only selects between the reference's unchecked gradient operator and its
// bounds-checked wrapper (which yields (0,0) outside).  Whenever isInside holds every sample passes the bounds check, so
// always using the checked operator (bs_box.cuh sparse_gradient) gives the reference's values on both branches.

// georegression UtilAngle.dist(a,b): shortest angular distance
__device__ __forceinline__ double angle_dist(double a, double b) {
	const double diff = fabs(a - b);
	if (diff > 3.14159265358979323846) return 2.0 * 3.14159265358979323846 - diff;
	return diff;
}

// ---------------------------------------------------------------- feature slots
// Measured and dropped (profiles/r1_describe_staged_window_sweep.txt): copying each feature's integral window into shared
// memory with cp.async before sampling.  The windows are (18..30)*scale pixels wide, i.e. about as many words as there are
// reads, so the copy costs as much L1 traffic as it saves and the tile halves the occupancy: 37..59 ms against 29 ms per
// 256 frames.  The samples therefore read global memory (L1/L2 resident) directly.
// ---------------------------------------------------------------- orientation
// sampling geometry of one feature, computed once per feature
struct OriGeo {
	int r, sw;
	float inv_sw;                  // (k + 0.5f) * inv_sw truncates to k / sw for every k of the grid (never within 1/(2 sw) of an integer)
	double period, tl_x, tl_y;
};
__device__ __forceinline__ OriGeo ori_geometry(const DescribeParams& P, double c_x, double c_y, double ptScale) {
	OriGeo g;
	g.sw = 2 * P.ori_radius + 1;
	g.inv_sw = 1.0f / (float)g.sw;
	// setObjectRadius(scale*BoofDefaults.SURF_SCALE_TO_RADIUS): scale = radius*objectRadiusToScale; kernel width = scale*sampleWidth
	const double sc = (ptScale * P.ori_object_radius_mult) * P.ori_radius_to_scale;
	g.r = gradient_radius(sc * P.ori_sample_width);
	g.period = sc * P.ori_period;
	g.tl_x = c_x - P.ori_radius * g.period;
	g.tl_y = c_y - P.ori_radius * g.period;
	g.tl_x += 0.5; g.tl_y += 0.5;
	return g;
}

// sample k of the (2*radius+1)^2 grid.  scratch: dX[n], dY[n] (+ ang[n] for the sliding window)
template <typename T>
__device__ __forceinline__ void orientation_sample(const T* __restrict__ ii, int W, int H, const DescribeParams& P,
                                                   const double* __restrict__ wts, const OriGeo& g, int k, double* dX, double* dY,
                                                   double* ang) {
	const int y = (int)(((float)k + 0.5f) * g.inv_sw), x = k - y * g.sw;
	const int py = (int)(g.tl_y + y * g.period);
	const int px = (int)(g.tl_x + x * g.period);
	T gdx, gdy;
	sparse_gradient(ii, W, H, g.r, px, py, gdx, gdy);
	double dx = (double)gdx, dy = (double)gdy;
	if (P.ori_weighted) { const double w = wts[k]; dx = w * dx; dy = w * dy; }
	dX[k] = dx; dY[k] = dy;
#ifdef BS_EXP_NOATAN
	if (P.ori_variant == 1) ang[k] = dy + dx;
#else
	if (P.ori_variant == 1) ang[k] = sm_atan2(dy, dx);
#endif
}

// ImplOrientationSlidingWindowIntegral.compute :104 sorts the sample angles ascending (ddogleg QuickSort_F64 index sort; ties
// are unpinned there — here towards the lower index, like the oracle's stable argsort).
// Stable ascending argsort of ang[0..n) by one warp: sorted runs of 16 (insertion sort, lane = run), then merge levels where
// every element finds its place in the sibling run by binary search.  (angle, index) is a total order, so "rank among the
// smaller ones" is well defined and equal to the stable order.  ordA receives the result; ordB is scratch.
__device__ __forceinline__ bool key_less(const double* ang, int a, int b) {
	const double ka = ang[a], kb = ang[b];
	return ka < kb || (ka == kb && a < b);
}
__device__ void warp_merge_argsort(const double* __restrict__ ang, int n, int* ordA, int* ordB) {
	const int lane = threadIdx.x & 31;
	for (int r0 = lane * 16; r0 < n; r0 += 32 * 16) {                // runs of 16
		const int r1 = min(r0 + 16, n);
		for (int i = r0; i < r1; i++) {
			int j = i;
			while (j > r0 && key_less(ang, i, ordA[j - 1])) { ordA[j] = ordA[j - 1]; j--; }
			ordA[j] = i;
		}
	}
	__syncwarp();
	for (int len = 16; len < n; len <<= 1) {
		for (int e = lane; e < n; e += 32) {
			const int run0 = e & ~(len - 1);                          // start of my run
			const bool right = (run0 & len) != 0;
			const int pair0 = run0 & ~len;
			const int s0 = min(right ? pair0 : pair0 + len, n), s1 = min(s0 + len, n);   // sibling run
			const int me = ordA[e];
			int lo = s0, hi = s1;                                     // number of sibling elements smaller than me
			while (lo < hi) {
				const int mid = (lo + hi) >> 1;
				if (key_less(ang, ordA[mid], me)) lo = mid + 1; else hi = mid;
			}
			ordB[pair0 + (e - run0) + (lo - s0)] = me;
		}
		__syncwarp();
		int* t = ordA; ordA = ordB; ordB = t;
	}
	// an odd number of levels leaves the result in the caller's ordB: copy it home
	int levels = 0;
	for (int len = 16; len < n; len <<= 1) levels++;
	if (levels & 1) {
		for (int e = lane; e < n; e += 32) ordB[e] = ordA[e];         // (ordA/ordB were swapped: ordB is the caller's ordA)
		__syncwarp();
	}
}

// out[k] = in[order[k]] by one warp
__device__ __forceinline__ void warp_gather(double* __restrict__ out, const double* __restrict__ in, const int* __restrict__ order, int n) {
	for (int k = threadIdx.x & 31; k < n; k += 32) out[k] = in[order[k]];
	__syncwarp();
}

// estimateAngle :139-188 — the circular two-pointer sweep over the samples in sorted order (sA, sX, sY = angle, dx, dy of
// the k-th smallest angle; the reference's startIndex == endIndex test becomes start == end), one lane per feature: the
// running sums are order-sensitive fp64, so the walk itself is serial; several features walk side by side in one warp.
__device__ double sliding_window_sweep(const DescribeParams& P, const double* sX, const double* sY, const double* sA, int n) {
	// The reference's nested loops as one flat loop (a step either grows the window at `end` or drops `start`), so that the
	// lanes of the warp, each on its own feature, stay converged.  `may_grow` is false right after the inner loop's
	// "wrapped around to start" break, which sends the reference to the subtraction as well.
	int start = 0, end = 1;
	double sumX = sX[0], sumY = sY[0];
	double best = sumX * sumX + sumY * sumY, bestX = sumX, bestY = sumY;
	double startAngle = sA[0], endAngle = sA[1];
	bool may_grow = true;
	while (true) {
		if (may_grow && angle_dist(startAngle, endAngle) <= P.ori_window) {
			sumX += sX[end];
			sumY += sY[end];
			const double mag = sumX * sumX + sumY * sumY;
			if (mag > best) { best = mag; bestX = sumX; bestY = sumY; }
			end++;
			if (end >= n) end = 0;
			endAngle = sA[end];
			may_grow = end != start;
		} else {
			sumX -= sX[start];
			sumY -= sY[start];
			start++;
			if (start == n) break;
			startAngle = sA[start];
			may_grow = true;
		}
	}
	return sm_atan2(bestY, bestX);
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
	int r, regionR, ov, G, haar;
	float inv_G;                   // (k + 0.5f) * inv_G truncates to k / G
	double c_x, c_y;
};
__device__ __forceinline__ DescGeo desc_geometry(const DescribeParams& P, double x, double y, double scale) {
	DescGeo g;
	g.r = gradient_radius(P.sample_width * scale);
	g.haar = P.use_haar;
	const int regionSize = P.grid * P.sub;
	g.regionR = regionSize / 2;
	g.ov = P.variant == 1 ? P.overlap : 0;
	g.G = regionSize + 2 * g.ov;
	g.inv_G = 1.0f / (float)g.G;
	g.c_x = x + 0.5; g.c_y = y + 0.5;
	return g;
}
__device__ __forceinline__ void desc_pixel(const DescGeo& g, double c, double s, double scale, int gxi, int gyi, int& pixelX, int& pixelY) {
	const double regionY = (gyi - g.regionR - g.ov) * scale;
	const double regionX = (gxi - g.regionR - g.ov) * scale;
	pixelX = (int)(g.c_x + c * regionX - s * regionY);
	pixelY = (int)(g.c_y + s * regionX + c * regionY);
}
// sample k of the G x G grid.  gx/gy: raw gradients in the integral type (widened to double where the reference does)
template <typename T>
__device__ __forceinline__ void describe_sample(const T* __restrict__ ii, int W, int H, const DescGeo& g, double c, double s, double scale,
                                                int k, T* gx, T* gy) {
	const int gyi = (int)(((float)k + 0.5f) * g.inv_G), gxi = k - gyi * g.G;
	int pixelX, pixelY;
	desc_pixel(g, c, s, scale, gxi, gyi, pixelX, pixelY);
	T dx, dy;
	if (g.haar) sparse_haar(ii, W, H, g.r, pixelX, pixelY, dx, dy);
	else sparse_gradient(ii, W, H, g.r, pixelX, pixelY, dx, dy);
	gx[k] = dx; gy[k] = dy;
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
	if (!P.finish) {                       // one band of a colour descriptor: the raw sums, finished by planar_finish_kernel
		for (int k = lane; k < P.dof; k += 32) out_desc[k] = feat[k];
		__syncwarp();
		return;
	}
	for (int k = lane; k < P.dof; k += 32) out_desc[k] = scale_it ? feat[k] / norm : feat[k];
	if (lane == 0) *out_white = laplace_sign(ii, W, H, (int)(x + 0.5), (int)(y + 0.5), scale) ? 1 : 0;
	__syncwarp();
}

// ---------------------------------------------------------------- kernel
// A CTA of DESC_WARPS warps takes DESC_WARPS consecutive features per round.  The samples of all of them form one index
// space that the whole CTA walks (no idle lanes when the grid size is not a multiple of 32); the order-sensitive
// reductions are then done by one warp per feature, all DESC_WARPS features at once.
//
// Measured and dropped (profiles/r1_describe_staged_window_sweep.txt): copying each feature's integral window into shared
// memory with cp.async before sampling.  The windows are (18..30)*scale pixels wide, i.e. about as many words as there are
// reads, so the copy costs as much L1 traffic as it saves and the tile halves the occupancy: 37..59 ms against 29 ms per
// 256 frames.  The samples therefore read global memory (L1/L2 resident) directly.
struct Slot {                      // one feature of the round (shared memory)
	double x, y, scale, angle, c, s;
	long f;                        // pool index, -1: none
	int frame;                     // frame index relative to the first frame of the launch
	const void* ii;                // that frame's integral image
	int bin;                       // orientation bin of the descriptor's sample visiting order
	OriGeo og;
	DescGeo dg;
};

struct DescribeSmem {
	int ori_n, desc_g2;
	int tab_ori, tab_w, tab_wgrid;     // table sizes in doubles
	int per_warp;                      // bytes of scratch per feature slot
	int total;                         // bytes
};

static DescribeSmem describe_smem(const DescribeParams& p) {
	DescribeSmem s{};
	const int sw = 2 * p.ori_radius + 1;
	s.ori_n = sw * sw;
	const int region = p.grid * p.sub, ov = p.variant == 1 ? p.overlap : 0;
	const int G = region + 2 * ov;
	s.desc_g2 = G * G;
	s.tab_ori = p.ori_weighted ? s.ori_n : 0;
	s.tab_w = p.variant == 0 ? region * region : (p.sub + 2 * ov) * (p.sub + 2 * ov);
	s.tab_wgrid = p.variant == 1 ? p.grid * p.grid : 0;
	const int ori_bytes = p.ori_variant == 1 ? s.ori_n * (8 * 4 + 4 * 2) : s.ori_n * 8 * 2;   // dX, dY (, ang, spare, two index arrays)
	const int desc_bytes = p.dof * 8 + s.desc_g2 * 4 * 2;
	s.per_warp = ((ori_bytes > desc_bytes ? ori_bytes : desc_bytes) + 15) & ~15;
	s.total = (s.tab_ori + s.tab_w + s.tab_wgrid) * 8 + DESC_WARPS * ((int)sizeof(Slot) + s.per_warp);
	s.total = (s.total + 15) & ~15;
	return s;
}

#ifndef BS_DESC_MIN_BLOCKS
#define BS_DESC_MIN_BLOCKS 7
#endif
template <typename T>
__global__ void __launch_bounds__(DESC_WARPS * 32, BS_DESC_MIN_BLOCKS)
describe_kernel(const T* __restrict__ ii_all, int W, int H, DescribeParams P, DescribeTables tabs, DescribeSmem S, Pool pool,
                const int* __restrict__ d_begin, const int* __restrict__ d_end, int frame0, int do_orientation, int do_describe) {
	extern __shared__ __align__(16) unsigned char smem[];
	constexpr int NT = DESC_WARPS * 32;
	// layout: tables | slots | per-slot scratch
	double* t_ori = reinterpret_cast<double*>(smem);
	double* t_w = t_ori + S.tab_ori;
	double* t_wgrid = t_w + S.tab_w;
	Slot* slots = reinterpret_cast<Slot*>(t_wgrid + S.tab_wgrid);
	unsigned char* scratch = reinterpret_cast<unsigned char*>(slots + DESC_WARPS);
	const int tid = threadIdx.x;
	for (int k = tid; k < S.tab_ori; k += NT) t_ori[k] = tabs.d_ori_w[k];
	for (int k = tid; k < S.tab_w; k += NT) t_w[k] = tabs.d_w[k];
	for (int k = tid; k < S.tab_wgrid; k += NT) t_wgrid[k] = tabs.d_wgrid[k];
	const int warp = tid >> 5, lane = tid & 31;
	Slot& me = slots[warp];
	unsigned char* mine = scratch + (size_t)warp * S.per_warp;
	const long frame_words = (long)W * H;
	// (idx + 0.5) / n is never within 1/(2n) of an integer, far more than the float rounding error for idx < 2^16
	const float inv_ori = 1.0f / (float)S.ori_n, inv_desc = 1.0f / (float)S.desc_g2;

	long end = *d_end;
	if (end > pool.cap) end = pool.cap;
	const long begin = *d_begin;
	const bool emit_walk = do_orientation == 2;        // sliding window: leave the walks to sliding_walk_kernel
	for (long base = begin + (long)blockIdx.x * DESC_WARPS; base < end; base += (long)gridDim.x * DESC_WARPS) {
		__syncthreads();                       // previous round (and the table fill) done with slots and scratch
		const int live = (int)(end - base < DESC_WARPS ? end - base : DESC_WARPS);
		if (lane == 0) {
			const long f = base + warp;
			me.f = f < end ? f : -1;
			if (f < end) {
				me.x = pool.xy[2 * f]; me.y = pool.xy[2 * f + 1]; me.scale = pool.scale[f];
				me.frame = pool.frame[f] - frame0;
				me.ii = ii_all + me.frame * frame_words;
				if (do_orientation) me.og = ori_geometry(P, me.x, me.y, me.scale);
				else me.angle = pool.angle[f];
				if (do_describe) me.dg = desc_geometry(P, me.x, me.y, me.scale);
			}
		}
		__syncthreads();
		if (do_orientation) {
			for (int idx = tid; idx < live * S.ori_n; idx += NT) {
				const int j = (int)(((float)idx + 0.5f) * inv_ori), k = idx - j * S.ori_n;
				const Slot& sj = slots[j];
				double* dX = reinterpret_cast<double*>(scratch + (size_t)j * S.per_warp);
				double* dY = dX + S.ori_n;
				orientation_sample(static_cast<const T*>(sj.ii), W, H, P, t_ori, sj.og, k, dX, dY, dY + S.ori_n);
			}
			__syncthreads();
			if (P.ori_variant == 0) {
				// average: the serial sums of all slots at once, lane 2j sums x and lane 2j+1 sums y of slot j (sample order),
				// instead of one mostly idle warp per feature
				if (warp == 0) {
					const int j = lane >> 1;
					const bool on = lane < 2 * DESC_WARPS && slots[j < DESC_WARPS ? j : 0].f >= 0;
					double total = 0.0;
					if (on) {
						const double* src = reinterpret_cast<const double*>(scratch + (size_t)j * S.per_warp) + (lane & 1) * S.ori_n;
						for (int k = 0; k < S.ori_n; k++) total += src[k];
					}
					const double Dy = __shfl_down_sync(0xffffffffu, total, 1);
					if (on && (lane & 1) == 0) {
						const double angle = sm_atan2(Dy, total);
						slots[j].angle = angle;
						pool.angle[slots[j].f] = angle;
					}
				}
			} else {
				// sliding window: every warp sorts the sample angles of its own feature and walks the window over them
				if (me.f >= 0) {
					bool continue_walk = true;
					// scratch: dX | dY | ang | spare | order.  After the sort the three arrays are rewritten in sorted order,
					// ping-pong through the spare: sA = spare, sX = old ang, sY = old dX
					double* dX = reinterpret_cast<double*>(mine);
					double* dY = dX + S.ori_n; double* ang = dY + S.ori_n; double* spare = ang + S.ori_n;
					int* order = reinterpret_cast<int*>(spare + S.ori_n);
#ifdef BS_EXP_NOSORT
					for (int k = lane; k < S.ori_n; k += 32) order[k] = k;
					__syncwarp();
#else
					warp_merge_argsort(ang, S.ori_n, order, order + S.ori_n);
#endif
					const long f_local = me.f - begin;
					if (emit_walk && f_local < tabs.walk_cap) {
						// sorted records to the scratch; sliding_walk_kernel walks them, one thread per feature
						double2* rec = reinterpret_cast<double2*>(tabs.d_walk) + (size_t)f_local * S.ori_n * 2;
						for (int k = lane; k < S.ori_n; k += 32) {
							const int o = order[k];
							rec[2 * k] = make_double2(dX[o], dY[o]);
							rec[2 * k + 1] = make_double2(ang[o], 0.0);
						}
						continue_walk = false;
					}
					if (continue_walk) {
					warp_gather(spare, ang, order, S.ori_n);
					warp_gather(ang, dX, order, S.ori_n);
					warp_gather(dX, dY, order, S.ori_n);
					// measured: walking the windows of all slots side by side in warp 0 (4x fewer issue slots) leaves the other
					// warps at the barrier for the whole walk and is slower (767 vs 679 us per 1080p frame); each warp walks its own
					if (lane == 0) {
#ifdef BS_EXP_NOSWEEP
						const double angle = sm_atan2(dX[0], ang[0]);
#else
						const double angle = sliding_window_sweep(P, ang, dX, spare, S.ori_n);
#endif
						me.angle = angle;
						pool.angle[me.f] = angle;
					}
					}
				}
			}
			__syncthreads();                   // angles are in the slots
		}
		if (do_describe) {
			// cos/sin of all slots by the first lanes of warp 0
			if (warp == 0 && lane < DESC_WARPS && slots[lane].f >= 0) {
				const double a = slots[lane].angle;
				slots[lane].c = sm_cos(a); slots[lane].s = sm_sin(a);
				int bin = (int)((a < 0.0 ? a + 3.14159265358979323846 : a) * (DESC_PERM_BINS / 3.14159265358979323846));
				slots[lane].bin = bin < 0 ? 0 : (bin >= DESC_PERM_BINS ? DESC_PERM_BINS - 1 : bin);   // also catches NaN
			}
			__syncthreads();                   // every slot has cos/sin; the orientation scratch is free
			for (int idx = tid; idx < live * S.desc_g2; idx += NT) {
				const int j = (int)(((float)idx + 0.5f) * inv_desc);
				const Slot& sj = slots[j];
				const int k = __ldg(tabs.d_perm + sj.bin * S.desc_g2 + (idx - j * S.desc_g2));
				T* gx = reinterpret_cast<T*>(scratch + (size_t)j * S.per_warp + (size_t)P.dof * 8);
				describe_sample(static_cast<const T*>(sj.ii), W, H, sj.dg, sj.c, sj.s, sj.scale, k, gx, gx + S.desc_g2);
			}
			__syncthreads();
			if (me.f >= 0) {
				double* feat = reinterpret_cast<double*>(mine);
				const T* gx = reinterpret_cast<const T*>(feat + P.dof);
				describe_reduce(static_cast<const T*>(me.ii), W, H, P, t_w, t_wgrid, me.x, me.y, me.c, me.s, me.scale, gx, gx + S.desc_g2, feat,
				                pool.desc + me.f * P.out_stride + P.out_offset, pool.white + me.f);
			}
		}
	}
}

// estimateAngle :139-188 for the features whose sorted samples describe_kernel left in the scratch: one THREAD per feature.
// Same flat walk as sliding_window_sweep over 32-byte records {dx, dy, angle, -}; the record each pointer will need next is
// loaded one step ahead.  Every step is a dependent chain (and an L2 access), so the kernel lives on the number of walks
// in flight: 1024 per SM here against 16 when every walk owns a warp.
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
constexpr int WALK_AHEAD = 12;     // records (32 B) the L2 prefetch runs ahead of each pointer
__global__ void __launch_bounds__(128, 10) sliding_walk_kernel(const double2* __restrict__ rec_all, int n, DescribeParams P, Pool pool,
                                                           const int* __restrict__ d_begin, const int* __restrict__ d_end, long walk_cap) {
	long end_f = *d_end;
	if (end_f > pool.cap) end_f = pool.cap;
	const long begin_f = *d_begin;
	long count = end_f - begin_f;
	if (count > walk_cap) count = walk_cap;
	for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (long)gridDim.x * blockDim.x) {
		const double2* r = rec_all + (size_t)i * n * 2;
		int start = 0, end = 1;
		double2 sxy = __ldg(r), sa = __ldg(r + 1);                       // record at `start`
		double2 exy = __ldg(r + 2), ea = __ldg(r + 3);                   // record at `end`
		int ne = 2 < n ? 2 : 0;                                          // index of the record after `end`
		double2 nexy = __ldg(r + 2 * ne), nea = __ldg(r + 2 * ne + 1);
		double2 nsxy = exy, nsa = ea;                                    // record after `start` (index 1)
		double sumX = sxy.x, sumY = sxy.y;
		double best = sumX * sumX + sumY * sumY, bestX = sumX, bestY = sumY;
		bool may_grow = true;
		// the scratch of a chunk is far larger than L2: pull the lines both pointers are about to reach out of DRAM early
		for (int k = 4; k < WALK_AHEAD && k < n; k += 4) prefetch_l2(r + 2 * k);
		while (true) {
			if (may_grow && angle_dist(sa.x, ea.x) <= P.ori_window) {
				sumX += exy.x;
				sumY += exy.y;
				const double mag = sumX * sumX + sumY * sumY;
				if (mag > best) { best = mag; bestX = sumX; bestY = sumY; }
				end = ne;
				exy = nexy; ea = nea;
				ne = end + 1 < n ? end + 1 : 0;
				nexy = __ldg(r + 2 * ne); nea = __ldg(r + 2 * ne + 1);
				if ((end & 3) == 0) { const int pf = end + WALK_AHEAD; prefetch_l2(r + 2 * (pf < n ? pf : pf - n)); }   // one 128-byte line = 4 records
				may_grow = end != start;
			} else {
				sumX -= sxy.x;
				sumY -= sxy.y;
				start++;
				if (start == n) break;
				sxy = nsxy; sa = nsa;
				const int ns = start + 1 < n ? start + 1 : start;
				nsxy = __ldg(r + 2 * ns); nsa = __ldg(r + 2 * ns + 1);
				if ((start & 3) == 0 && start + WALK_AHEAD < n) prefetch_l2(r + 2 * (start + WALK_AHEAD));
				may_grow = true;
			}
		}
		pool.angle[begin_f + i] = sm_atan2(bestY, bestX);
	}
}

template <typename T>
void launch_describe(const T* d_ii, int W, int H, const DescribeParams& p, const DescribeTables& t, const Pool& pool,
                     const int* d_begin, const int* d_end, int frame0, bool do_orientation, bool do_describe, cudaStream_t st) {
	const int sm_count = device_sm_count();
	const DescribeSmem S = describe_smem(p);
	if (S.total > 220 * 1024) fail(BS_ERR_ARG, "orientation/descriptor configuration needs too much shared memory");
	BS_CUDA(cudaFuncSetAttribute(describe_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, S.total));
	int ctas_per_sm = (224 * 1024) / (S.total + 1024);
	const int cap = BS_DESC_MIN_BLOCKS > 1 ? BS_DESC_MIN_BLOCKS : 32 / DESC_WARPS;   // resident CTAs per SM the register budget allows
	static const int env_cap = getenv("BS_DESC_CTAS") ? atoi(getenv("BS_DESC_CTAS")) : 0;
	ctas_per_sm = ctas_per_sm > cap ? cap : (ctas_per_sm < 1 ? 1 : ctas_per_sm);
	if (env_cap > 0 && env_cap < ctas_per_sm) ctas_per_sm = env_cap;
	if (do_orientation && p.ori_variant == 1 && t.d_walk && t.walk_cap > 0) {
		// sliding window with scratch: samples + argsort (describe_kernel, mode 2), the walks (one thread each), then the descriptors
		BS_LAUNCH((describe_kernel<T>), sm_count * ctas_per_sm, DESC_WARPS * 32, S.total, st, d_ii, W, H, p, t, S, pool, d_begin, d_end, frame0, 2, 0);
		BS_LAUNCH(sliding_walk_kernel, sm_count * 10, 128, 0, st, reinterpret_cast<const double2*>(t.d_walk), S.ori_n, p, pool, d_begin, d_end,
		          t.walk_cap);
		if (do_describe)
			BS_LAUNCH((describe_kernel<T>), sm_count * ctas_per_sm, DESC_WARPS * 32, S.total, st, d_ii, W, H, p, t, S, pool, d_begin, d_end, frame0, 0, 1);
		return;
	}
	BS_LAUNCH((describe_kernel<T>), sm_count * ctas_per_sm, DESC_WARPS * 32, S.total, st, d_ii, W, H, p, t, S, pool, d_begin, d_end, frame0,
	          do_orientation ? 1 : 0, do_describe ? 1 : 0);
}
// DescribePointSurfPlanar.describe :106-110: UtilFeature.normalizeL2 over all bands, Laplacian sign from the grey integral.
// One warp per feature (persistent, the range is only known on the device); every lane repeats the serial sum of
// squares (broadcast loads) so the order is the reference's.
template <typename T>
__global__ void __launch_bounds__(256) planar_finish_kernel(const T* __restrict__ gray_ii, int W, int H, int total_dof, Pool pool,
                                                            const int* __restrict__ d_begin, const int* __restrict__ d_end, int frame0) {
	const int lane = threadIdx.x & 31;
	long end = *d_end;
	if (end > pool.cap) end = pool.cap;
	const long warps = (long)gridDim.x * (blockDim.x >> 5);
	for (long f = *d_begin + (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); f < end; f += warps) {
		double* d = pool.desc + f * total_dof;
		double norm = 0.0;
		for (int k = 0; k < total_dof; k++) { const double t = d[k]; norm += t * t; }
		__syncwarp();
		if (norm != 0.0) {
			norm = sqrt(norm);
			for (int k = lane; k < total_dof; k += 32) d[k] = d[k] / norm;
		}
		if (lane == 0) {
			const T* ii = gray_ii + (long)(pool.frame[f] - frame0) * W * H;
			pool.white[f] = laplace_sign(ii, W, H, (int)(pool.xy[2 * f] + 0.5), (int)(pool.xy[2 * f + 1] + 0.5), pool.scale[f]) ? 1 : 0;
		}
	}
}

template <typename T>
void launch_planar_finish(const T* d_gray_ii, int W, int H, int total_dof, const Pool& pool, const int* d_begin, const int* d_end,
                          int frame0, cudaStream_t st) {
	int dev = 0, sm_count = 0;
	BS_CUDA(cudaGetDevice(&dev));
	BS_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
	BS_LAUNCH((planar_finish_kernel<T>), sm_count * 8, 256, 0, st, d_gray_ii, W, H, total_dof, pool, d_begin, d_end, frame0);
}
template void launch_planar_finish<float>(const float*, int, int, int, const Pool&, const int*, const int*, int, cudaStream_t);
template void launch_planar_finish<int32_t>(const int32_t*, int, int, int, const Pool&, const int*, const int*, int, cudaStream_t);

// ConvertImage.average (core/image/impl/ImplConvertPlanarToGray.java:44-84 U8, :213-253 S32, :296-336 F32): bands summed left to
// right in the pixel type's arithmetic (float adds, wrapping int adds), then divided by the band count
template <typename T> struct AvgAcc;
template <> struct AvgAcc<float> { static __device__ __forceinline__ float run(const float* p, long bs, int nb) {
	if (nb == 1) return p[0];
	float sum = nb == 3 ? p[0] : 0.0f;
	for (int b = nb == 3 ? 1 : 0; b < nb; b++) sum += p[b * bs];
	return sum / nb; } };
template <> struct AvgAcc<uint8_t> { static __device__ __forceinline__ uint8_t run(const uint8_t* p, long bs, int nb) {
	int sum = 0;
	for (int b = 0; b < nb; b++) sum += p[b * bs];
	return (uint8_t)(sum / nb); } };
template <> struct AvgAcc<int32_t> { static __device__ __forceinline__ int32_t run(const int32_t* p, long bs, int nb) {
	uint32_t sum = 0;
	for (int b = 0; b < nb; b++) sum += (uint32_t)p[b * bs];
	return (int32_t)sum / nb; } };
template <typename T>
__global__ void average_bands_kernel(const T* __restrict__ in, long frame_stride, long band_stride, int stride, int bands, T* __restrict__ gray,
                                     int W, int H) {
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, f = blockIdx.z;
	if (x >= W) return;
	gray[((long)f * H + y) * W + x] = AvgAcc<T>::run(in + f * frame_stride + (long)y * stride + x, band_stride, bands);
}
template <typename T>
void launch_average_bands(const T* d_in, long frame_stride, long band_stride, int stride, int bands, T* d_gray, int W, int H, int batch,
                          cudaStream_t st) {
	BS_LAUNCH((average_bands_kernel<T>), dim3(div_up(W, 256), H, batch), 256, 0, st, d_in, frame_stride, band_stride, stride, bands, d_gray,
	          W, H);
}
template void launch_average_bands<float>(const float*, long, long, int, int, float*, int, int, int, cudaStream_t);
template void launch_average_bands<uint8_t>(const uint8_t*, long, long, int, int, uint8_t*, int, int, int, cudaStream_t);
template void launch_average_bands<int32_t>(const int32_t*, long, long, int, int, int32_t*, int, int, int, cudaStream_t);

template void launch_describe<float>(const float*, int, int, const DescribeParams&, const DescribeTables&, const Pool&, const int*, const int*,
                                     int, bool, bool, cudaStream_t);
template void launch_describe<int32_t>(const int32_t*, int, int, const DescribeParams&, const DescribeTables&, const Pool&, const int*,
                                       const int*, int, bool, bool, cudaStream_t);

}  // namespace bs
