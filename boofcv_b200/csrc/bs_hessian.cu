// bs_hessian.cu — Fast-Hessian determinant response for every layer of the scale space.
// Reference: IntegralImageFeatureIntensity.hessian (F/alg/feature/detect/intensity/IntegralImageFeatureIntensity.java:43-79)
//   = hessianBorder (impl/ImplIntegralImageFeatureIntensity.java:72-113, computeHessian :115-127 through
//     ImplIntegralImageOps.convolveSparse/block_zero, IP/alg/transform/ii/impl/ImplIntegralImageOps.java:172-214)
//   + hessianInner (:132-213; S32 twins :245-386).
// Box geometry: DerivativeIntegralImage.kernelDerivXX/YY/XY (IP/alg/transform/ii/DerivativeIntegralImage.java:102-158).
//
// The float expressions are evaluated in the reference's order: every box is br-tr-bl+tl left to right,
// Dxx = A - 3*B, Dxy = TL - TR + BR - BL, each *= norm, response = Dxx*Dyy - (0.81f*Dxy)*Dxy.  No FMA (-fmad=false).
// Border samples (`inner == false`) follow convolveSparse: ret = 0; ret += box*weight — the same values for float, a
// different int/float conversion point for S32 integrals, so both forms are kept.
#include "bs_common.cuh"
#include <cstdlib>
#include "bs_box.cuh"

namespace bs {

__device__ __forceinline__ float finish(float Dxx, float Dyy, float Dxy, float norm) {
	Dxx *= norm; Dxy *= norm; Dyy *= norm;
	const float a = Dxx * Dyy;
	float b = 0.81f * Dxy;
	b = b * Dxy;
	return a - b;
}

// hessianInner at pixel (px,py) (:160-213): all 32 corners inside the image.  The reference addresses the corners from the
// top-left of the feature (radiusFeature = size/2, blockSmall = size/3, blockLarge = size-blockSmall-1,
// radiusSkinny = blockLarge/2); the same arithmetic is used here so that unusual sizes behave like the reference too.
template <typename T> struct Corners {
	const T *top, *bottom;         // Dxx rows, at column px-radiusFeature-1
	const T *yl;                   // Dyy: row py-radiusFeature-1, column px-radiusSkinny-1
	const T *y1, *y2, *y3, *y4;    // Dxy rows, at column px-blockSmall-1
	__device__ __forceinline__ Corners(const T* ii, int W, const Layer& L, int px, int py) {
		const int rf = L.size / 2, bs = L.b, bl = L.size - bs - 1, rs = bl / 2;
		top = ii + (long)(py - rs - 1) * W + (px - rf - 1);
		bottom = top + (long)bl * W;
		yl = ii + (long)(py - rf - 1) * W + (px - rs - 1);
		y1 = ii + (long)(py - bs - 1) * W + (px - bs - 1);
		y2 = y1 + (long)bs * W;
		y3 = y2 + W;
		y4 = y3 + (long)bs * W;
	}
};
__device__ __forceinline__ float response_inner(const float* __restrict__ ii, int W, const Layer& L, int px, int py) {
	const Corners<float> c(ii, W, L, px, py);
	const int bs = L.b, bl = L.size - bs - 1, w2 = 2 * bs, w3 = 3 * bs, x3 = bs + 1, x4 = x3 + bs;
	const long ro1 = (long)bs * W, ro2 = 2 * ro1, ro3 = 3 * ro1;
	typedef Arith<float> A;
	float Dxx = A::sum4(__ldg(c.bottom + w3), __ldg(c.top + w3), __ldg(c.bottom), __ldg(c.top));
	Dxx -= 3.0f * A::sum4(__ldg(c.bottom + w2), __ldg(c.top + w2), __ldg(c.bottom + bs), __ldg(c.top + bs));
	const float* yr = c.yl + bl;
	float Dyy = A::sum4(__ldg(yr + ro3), __ldg(c.yl + ro3), __ldg(yr), __ldg(c.yl));
	Dyy -= 3.0f * A::sum4(__ldg(yr + ro2), __ldg(c.yl + ro2), __ldg(yr + ro1), __ldg(c.yl + ro1));
	float Dxy = A::sum4(__ldg(c.y2 + bs), __ldg(c.y1 + bs), __ldg(c.y2), __ldg(c.y1));
	Dxy -= A::sum4(__ldg(c.y2 + x4), __ldg(c.y1 + x4), __ldg(c.y2 + x3), __ldg(c.y1 + x3));
	Dxy += A::sum4(__ldg(c.y4 + x4), __ldg(c.y3 + x4), __ldg(c.y4 + x3), __ldg(c.y3 + x3));
	Dxy -= A::sum4(__ldg(c.y4 + bs), __ldg(c.y3 + bs), __ldg(c.y4), __ldg(c.y3));
	return finish(Dxx, Dyy, Dxy, L.norm);
}
__device__ __forceinline__ float response_inner(const int32_t* __restrict__ ii, int W, const Layer& L, int px, int py) {
	// :356-368 — int box sums; "float Dxx = <int>; Dxx -= 3*(<int>)" converts each term separately
	const Corners<int32_t> c(ii, W, L, px, py);
	const int bs = L.b, bl = L.size - bs - 1, w2 = 2 * bs, w3 = 3 * bs, x3 = bs + 1, x4 = x3 + bs;
	const long ro1 = (long)bs * W, ro2 = 2 * ro1, ro3 = 3 * ro1;
	typedef Arith<int32_t> A;
	float Dxx = (float)A::sum4(__ldg(c.bottom + w3), __ldg(c.top + w3), __ldg(c.bottom), __ldg(c.top));
	Dxx -= (float)(int32_t)(3u * (uint32_t)A::sum4(__ldg(c.bottom + w2), __ldg(c.top + w2), __ldg(c.bottom + bs), __ldg(c.top + bs)));
	const int32_t* yr = c.yl + bl;
	float Dyy = (float)A::sum4(__ldg(yr + ro3), __ldg(c.yl + ro3), __ldg(yr), __ldg(c.yl));
	Dyy -= (float)(int32_t)(3u * (uint32_t)A::sum4(__ldg(yr + ro2), __ldg(c.yl + ro2), __ldg(yr + ro1), __ldg(c.yl + ro1)));
	float Dxy = (float)A::sum4(__ldg(c.y2 + bs), __ldg(c.y1 + bs), __ldg(c.y2), __ldg(c.y1));
	Dxy -= (float)A::sum4(__ldg(c.y2 + x4), __ldg(c.y1 + x4), __ldg(c.y2 + x3), __ldg(c.y1 + x3));
	Dxy += (float)A::sum4(__ldg(c.y4 + x4), __ldg(c.y3 + x4), __ldg(c.y4 + x3), __ldg(c.y3 + x3));
	Dxy -= (float)A::sum4(__ldg(c.y4 + bs), __ldg(c.y3 + bs), __ldg(c.y4), __ldg(c.y3));
	return finish(Dxx, Dyy, Dxy, L.norm);
}

// computeHessian at a border pixel: convolveSparse with block_zero reads
__device__ __forceinline__ float response_border(const float* __restrict__ ii, int W, int H, const Layer& L, int px, int py) {
	float Dxx = 0.0f;
	Dxx += box_zero(ii, W, H, px - L.r2 - 1, py - L.r3 - 1, px + L.r2, py + L.r3) * 1.0f;
	Dxx += box_zero(ii, W, H, px - L.r1 - 1, py - L.r3 - 1, px + L.r1, py + L.r3) * -3.0f;
	float Dyy = 0.0f;
	Dyy += box_zero(ii, W, H, px - L.r3 - 1, py - L.r2 - 1, px + L.r3, py + L.r2) * 1.0f;
	Dyy += box_zero(ii, W, H, px - L.r3 - 1, py - L.r1 - 1, px + L.r3, py + L.r1) * -3.0f;
	const int b = L.b;
	float Dxy = 0.0f;
	Dxy += box_zero(ii, W, H, px - b - 1, py - b - 1, px - 1, py - 1) * 1.0f;
	Dxy += box_zero(ii, W, H, px, py - b - 1, px + b, py - 1) * -1.0f;
	Dxy += box_zero(ii, W, H, px, py, px + b, py + b) * 1.0f;
	Dxy += box_zero(ii, W, H, px - b - 1, py, px - 1, py + b) * -1.0f;
	return finish(Dxx, Dyy, Dxy, L.norm);
}
__device__ __forceinline__ float response_border(const int32_t* __restrict__ ii, int W, int H, const Layer& L, int px, int py) {
	// S32 convolveSparse (:216-227) accumulates in int and converts once
	uint32_t xx = (uint32_t)box_zero(ii, W, H, px - L.r2 - 1, py - L.r3 - 1, px + L.r2, py + L.r3);
	xx += (uint32_t)box_zero(ii, W, H, px - L.r1 - 1, py - L.r3 - 1, px + L.r1, py + L.r3) * (uint32_t)(-3);
	uint32_t yy = (uint32_t)box_zero(ii, W, H, px - L.r3 - 1, py - L.r2 - 1, px + L.r3, py + L.r2);
	yy += (uint32_t)box_zero(ii, W, H, px - L.r3 - 1, py - L.r1 - 1, px + L.r3, py + L.r1) * (uint32_t)(-3);
	const int b = L.b;
	uint32_t xy = (uint32_t)box_zero(ii, W, H, px - b - 1, py - b - 1, px - 1, py - 1);
	xy -= (uint32_t)box_zero(ii, W, H, px, py - b - 1, px + b, py - 1);
	xy += (uint32_t)box_zero(ii, W, H, px, py, px + b, py + b);
	xy -= (uint32_t)box_zero(ii, W, H, px - b - 1, py, px - 1, py + b);
	return finish((float)(int32_t)xx, (float)(int32_t)yy, (float)(int32_t)xy, L.norm);
}

// The reference's inner region: x,y in [border, dim-border) with
// borderOrig = radiusFeature+1+(skip-(radiusFeature+1)%skip), border = borderOrig/skip (:140-158)
__device__ __forceinline__ int layer_border(const Layer& L) {
	const int rf = L.size / 2;
	const int border_orig = rf + 1 + (L.skip - (rf + 1) % L.skip);
	return border_orig / L.skip;
}

template <typename T>
__device__ __forceinline__ float response_at(const T* __restrict__ ii, int W, int H, const Layer& L, int x, int y) {
	const int border = layer_border(L);
	const bool inner = x >= border && x < L.w - border && y >= border && y < L.h - border;
	const int px = x * L.skip, py = y * L.skip;
	return inner ? response_inner(ii, W, L, px, py) : response_border(ii, W, H, L, px, py);
}

// ---------------------------------------------------------------- generic kernel: one thread = one sample of one octave,
// all scales of that octave.  grid.x enumerates 32x8 sample tiles of every octave, grid.y = frame.
constexpr int HT_X = 32, HT_Y = 8;

struct OctaveTiles {
	int n_oct;
	int first_layer[MAX_OCTAVES];
	int n_scales[MAX_OCTAVES];
	int tile_begin[MAX_OCTAVES + 1];
	int tiles_x[MAX_OCTAVES];
};

template <typename T>
__global__ void __launch_bounds__(HT_X * HT_Y)
hessian_generic_kernel(const T* __restrict__ ii_all, const Plan* __restrict__ plan, OctaveTiles ot, float* __restrict__ inten_all) {
	int o = 0;
	while (o + 1 < ot.n_oct && (int)blockIdx.x >= ot.tile_begin[o + 1]) o++;
	const int tile = blockIdx.x - ot.tile_begin[o];
	const int tx = tile % ot.tiles_x[o], ty = tile / ot.tiles_x[o];
	const int W = plan->W, H = plan->H;
	const T* ii = ii_all + (long)blockIdx.y * W * H;
	float* inten = inten_all + (long)blockIdx.y * plan->inten_per_frame;
	const int x = tx * HT_X + (threadIdx.x & (HT_X - 1));
	const int y = ty * HT_Y + (threadIdx.x / HT_X);
	const Layer& L0 = plan->layers[ot.first_layer[o]];
	if (x >= L0.w || y >= L0.h) return;
	for (int s = 0; s < ot.n_scales[o]; s++) {
		const Layer& L = plan->layers[ot.first_layer[o] + s];
		inten[L.offset + (long)y * L.w + x] = response_at(ii, W, H, L, x, y);
	}
}

}  // namespace bs
#include "bs_hessian_tile.cuh"
namespace bs {

// octaves the shared-memory kernel is specialised for (float integral, FH-9 default sizes); returns true if launched
template <typename T>
static bool try_tiled(const T*, const Plan&, int, int, float*, int, cudaStream_t) { return false; }
template <>
bool try_tiled<float>(const float* d_ii, const Plan& plan, int first, int ns, float* d_inten, int batch, cudaStream_t st) {
	if (ns != 4) return false;
	const Layer& L0 = plan.layers[first];
	const int step = plan.layers[first + 1].size - L0.size;
	for (int i = 1; i < ns; i++)
		if (plan.layers[first + i].size != L0.size + i * step) return false;
	static const int tma_mask = getenv("BS_HESSIAN_NO_TMA") ? 0 : (getenv("BS_HESSIAN_TMA_MASK") ? atoi(getenv("BS_HESSIAN_TMA_MASK")) : 3);
	if (L0.skip == 1 && L0.size == 9 && step == 6) {
		if (!((tma_mask & 1) && launch_hessian_tma<1, 9, 6, 4, 64, 16>(d_ii, plan, first, d_inten, batch, st)))
			launch_hessian_tile<1, 9, 6, 4, 64, 16>(d_ii, plan, first, d_inten, batch, st);
		return true;
	}
	if (L0.skip == 2 && L0.size == 15 && step == 12) {
		if (!((tma_mask & 2) && launch_hessian_tma<2, 15, 12, 4, 32, 16, 512>(d_ii, plan, first, d_inten, batch, st)))
			launch_hessian_tile<2, 15, 12, 4, 32, 16>(d_ii, plan, first, d_inten, batch, st);
		return true;
	}
	return false;
}

template <typename T>
void launch_hessian_plan(const T* d_ii, const Plan& plan, const Plan* d_plan, float* d_inten, int batch, cudaStream_t st) {
	if (batch <= 0) return;
	OctaveTiles ot{};
	int tiles = 0, o = 0;
	for (int l = 0; l < plan.n_layers;) {
		int ns = 1;
		while (l + ns < plan.n_layers && plan.layers[l + ns].index != 0) ns++;
		if (!try_tiled<T>(d_ii, plan, l, ns, d_inten, batch, st)) {
			const Layer& L = plan.layers[l];
			ot.first_layer[o] = l; ot.n_scales[o] = ns; ot.tile_begin[o] = tiles;
			ot.tiles_x[o] = div_up(L.w, HT_X);
			tiles += ot.tiles_x[o] * div_up(L.h, HT_Y);
			o++;
		}
		l += ns;
	}
	ot.n_oct = o;
	ot.tile_begin[o] = tiles;
	if (tiles == 0) return;
	BS_LAUNCH((hessian_generic_kernel<T>), dim3(tiles, batch), HT_X * HT_Y, 0, st, d_ii, d_plan, ot, d_inten);
}
template void launch_hessian_plan<float>(const float*, const Plan&, const Plan*, float*, int, cudaStream_t);
template void launch_hessian_plan<int32_t>(const int32_t*, const Plan&, const Plan*, float*, int, cudaStream_t);

// ---------------------------------------------------------------- single layer (stage-level entry point bs_hessian)
template <typename T>
__global__ void hessian_single_kernel(const T* __restrict__ ii, int W, int H, Layer L, float* __restrict__ out) {
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
	if (x >= L.w || y >= L.h) return;
	out[(long)y * L.w + x] = response_at(ii, W, H, L, x, y);
}

Layer make_layer(int size, int skip, int W, int H);

template <typename T>
void launch_hessian_single(const T* d_ii, int W, int H, int skip, int size, float* d_out, cudaStream_t st) {
	Layer L = make_layer(size, skip, W, H);
	if (L.w <= 0 || L.h <= 0) return;
	dim3 block(32, 8), grid(div_up(L.w, 32), div_up(L.h, 8));
	BS_LAUNCH((hessian_single_kernel<T>), grid, block, 0, st, d_ii, W, H, L, d_out);
}
template void launch_hessian_single<float>(const float*, int, int, int, int, float*, cudaStream_t);
template void launch_hessian_single<int32_t>(const int32_t*, int, int, int, int, float*, cudaStream_t);

}  // namespace bs
