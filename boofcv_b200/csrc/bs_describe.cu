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

// ---------------------------------------------------------------- orientation
// scratch: dX[n], dY[n] (+ ang[n], order[n] for the sliding window)
template <typename T>
__device__ double orientation_warp(const T* __restrict__ ii, int W, int H, const DescribeParams& P, const double* __restrict__ wts,
                                   double c_x, double c_y, double ptScale, double* dX, double* dY, double* ang, int* order) {
	const int lane = threadIdx.x & 31;
	const int sw = 2 * P.ori_radius + 1, n = sw * sw;
	// setObjectRadius(scale*BoofDefaults.SURF_SCALE_TO_RADIUS): scale = radius*objectRadiusToScale; kernel width = scale*sampleWidth
	const double sc = (ptScale * 2.0) * P.ori_radius_to_scale;
	const int r = gradient_radius(sc * P.ori_sample_width);
	const double period = sc * P.ori_period;
	double tl_x = c_x - P.ori_radius * period;
	double tl_y = c_y - P.ori_radius * period;
	tl_x += 0.5; tl_y += 0.5;
	for (int k = lane; k < n; k += 32) {
		const int y = k / sw, x = k - y * sw;
		const int py = (int)(tl_y + y * period);
		const int px = (int)(tl_x + x * period);
		T gdx, gdy;
		sparse_gradient(ii, W, H, r, px, py, gdx, gdy);
		double dx = (double)gdx, dy = (double)gdy;
		if (P.ori_weighted) { const double w = wts[k]; dx = w * dx; dy = w * dy; }
		dX[k] = dx; dY[k] = dy;
		if (P.ori_variant == 1) ang[k] = sm_atan2(dy, dx);
	}
	__syncwarp();
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

// gx/gy: G*G raw gradients in the integral type (widened to double where the reference does); feat: dof doubles
template <typename T>
__device__ void describe_warp(const T* __restrict__ ii, int W, int H, const DescribeParams& P, const double* __restrict__ wt,
                              const double* __restrict__ wgrid, double x, double y, double angle, double scale, T* gx, T* gy,
                              double* feat, double* out_desc, uint8_t* out_white) {
	const int lane = threadIdx.x & 31;
	const double c = sm_cos(angle), s = sm_sin(angle);
	const int r = gradient_radius(P.sample_width * scale);
	const int regionSize = P.grid * P.sub;
	const int regionR = regionSize / 2;
	const int ov = P.variant == 1 ? P.overlap : 0;
	const int G = regionSize + 2 * ov;           // sample grid (DescribePointSurfMod samples the overlap once)
	const double c_x = x + 0.5, c_y = y + 0.5;
	for (int k = lane; k < G * G; k += 32) {
		const int gyi = k / G, gxi = k - gyi * G;
		const double regionY = (gyi - regionR - ov) * scale;
		const double regionX = (gxi - regionR - ov) * scale;
		const int pixelX = (int)(c_x + c * regionX - s * regionY);
		const int pixelY = (int)(c_y + s * regionX + c * regionY);
		sparse_gradient(ii, W, H, r, pixelX, pixelY, gx[k], gy[k]);
	}
	__syncwarp();
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
struct DescribeSmem {
	int ori_n, desc_g2;
	int tab_ori, tab_w, tab_wgrid;     // table sizes in doubles
	int per_warp;                      // bytes
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
	const int ori_bytes = s.ori_n * (p.ori_variant == 1 ? 8 * 3 + 4 : 8 * 2);
	const int desc_bytes = p.dof * 8 + s.desc_g2 * 4 * 2;
	s.per_warp = ((ori_bytes > desc_bytes ? ori_bytes : desc_bytes) + 15) & ~15;
	s.total = (s.tab_ori + s.tab_w + s.tab_wgrid) * 8 + DESC_WARPS * s.per_warp;
	return s;
}

template <typename T>
__global__ void __launch_bounds__(DESC_WARPS * 32)
describe_kernel(const T* __restrict__ ii_all, int W, int H, DescribeParams P, DescribeTables tabs, DescribeSmem S, Pool pool,
                const int* __restrict__ d_begin, const int* __restrict__ d_end, int frame0, int do_orientation, int do_describe) {
	extern __shared__ __align__(16) unsigned char smem[];
	double* t_ori = reinterpret_cast<double*>(smem);
	double* t_w = t_ori + S.tab_ori;
	double* t_wgrid = t_w + S.tab_w;
	for (int k = threadIdx.x; k < S.tab_ori; k += blockDim.x) t_ori[k] = tabs.d_ori_w[k];
	for (int k = threadIdx.x; k < S.tab_w; k += blockDim.x) t_w[k] = tabs.d_w[k];
	for (int k = threadIdx.x; k < S.tab_wgrid; k += blockDim.x) t_wgrid[k] = tabs.d_wgrid[k];
	__syncthreads();
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	unsigned char* mine = reinterpret_cast<unsigned char*>(t_wgrid + S.tab_wgrid) + (long)warp * S.per_warp;
	// orientation view
	double* dX = reinterpret_cast<double*>(mine);
	double* dY = dX + S.ori_n;
	double* ang = dY + S.ori_n;
	int* order = reinterpret_cast<int*>(ang + S.ori_n);
	// descriptor view (same bytes; the two stages run one after the other)
	double* feat = reinterpret_cast<double*>(mine);
	T* gx = reinterpret_cast<T*>(feat + P.dof);
	T* gy = gx + S.desc_g2;

	long end = *d_end;
	if (end > pool.cap) end = pool.cap;
	const long nwarps = (long)gridDim.x * DESC_WARPS;
	for (long f = *d_begin + (long)blockIdx.x * DESC_WARPS + warp; f < end; f += nwarps) {
		const double x = pool.xy[2 * f], y = pool.xy[2 * f + 1], scale = pool.scale[f];
		const T* ii = ii_all + (long)(pool.frame[f] - frame0) * W * H;
		double angle;
		if (do_orientation) {
			angle = orientation_warp(ii, W, H, P, t_ori, x, y, scale, dX, dY, ang, order);
			if (lane == 0) pool.angle[f] = angle;
		} else {
			angle = pool.angle[f];
		}
		if (do_describe) describe_warp(ii, W, H, P, t_w, t_wgrid, x, y, angle, scale, gx, gy, feat, pool.desc + f * P.dof, pool.white + f);
	}
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
	const DescribeSmem S = describe_smem(p);
	if (S.total > 200 * 1024) fail(BS_ERR_ARG, "orientation/descriptor configuration needs too much shared memory");
	BS_CUDA(cudaFuncSetAttribute(describe_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, S.total));
	const int ctas_per_sm = S.total > 0 ? (200 * 1024) / S.total : 8;
	const int grid = sm_count * (ctas_per_sm > 8 ? 8 : (ctas_per_sm < 1 ? 1 : ctas_per_sm));
	BS_LAUNCH((describe_kernel<T>), grid, DESC_WARPS * 32, S.total, st, d_ii, W, H, p, t, S, pool, d_begin, d_end, frame0,
	          do_orientation ? 1 : 0, do_describe ? 1 : 0);
}
template void launch_describe<float>(const float*, int, int, const DescribeParams&, const DescribeTables&, const Pool&, const int*, const int*,
                                     int, bool, bool, cudaStream_t);
template void launch_describe<int32_t>(const int32_t*, int, int, const DescribeParams&, const DescribeTables&, const Pool&, const int*,
                                       const int*, int, bool, bool, cudaStream_t);

}  // namespace bs
