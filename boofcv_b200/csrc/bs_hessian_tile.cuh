// bs_hessian_tile.cuh — shared-memory Fast-Hessian kernel for one octave whose box sizes are known at compile time
// (the FH-9 default schedule: 9/15/21/27 @ skip 1, 15/27/39/51 @ skip 2).  Included by bs_hessian.cu.
//
// Why: the generic kernel issues ~370 instructions per warp-response (64-bit address arithmetic, Layer fields fetched from
// memory, border tests) for 32 loads + ~40 flops of real work, and ncu shows it issue-bound (sm__throughput 78 %).  Here
//   - a CTA stages the integral-image footprint of its TX x TY output tile (all scales of the octave) in shared memory
//     once, coalesced, and every corner read becomes ONE LDS with an immediate offset (sizes are template constants);
//   - for skip > 1 the columns are de-interleaved into `skip` planes while staging, so that the 32 lanes of a warp (32
//     consecutive samples = pixels `skip` apart) read consecutive words: no bank conflicts;
//   - samples in the reference's border ring (computeHessian / block_zero semantics) take the generic path per thread.
// Arithmetic order is exactly hessianInner's (see bs_hessian.cu).
#pragma once

namespace bs {

template <int SIZE> struct BoxGeo {
	static constexpr int bs = SIZE / 3;            // blockSmall
	static constexpr int bl = SIZE - bs - 1;       // blockLarge
	static constexpr int rf = SIZE / 2;            // radiusFeature
	static constexpr int rs = bl / 2;              // radiusSkinny
};

// Geometry of the staged tile for an octave: sizes S0 + i*STEP (i < NS), sample spacing SKIP, TX x TY samples per CTA
template <int SKIP, int S0, int STEP, int NS, int TX, int TY> struct TileGeo {
	static constexpr int SMAX = S0 + (NS - 1) * STEP;
	static constexpr int HALO = SMAX / 2 + 1;                          // pixels needed left of / above a sample
	static constexpr int TW = (TX - 1) * SKIP + 2 * HALO;              // footprint width  (px-HALO .. px+HALO-1)
	static constexpr int TH = (TY - 1) * SKIP + 2 * HALO;
	static constexpr int PW = (TW + SKIP - 1) / SKIP + ((((TW + SKIP - 1) / SKIP) % 2 == 0) ? 1 : 0);   // odd plane pitch
	static constexpr int PITCH = PW * SKIP;                            // words per staged row
	static constexpr int WORDS = PITCH * TH;
	// word offset of pixel (px+dc, py+dr) relative to the thread's base (= sample (tx,ty) of the tile)
	static __host__ __device__ constexpr int off(int dr, int dc) {
		return (HALO + dr) * PITCH + ((HALO + dc) % SKIP) * PW + (HALO + dc) / SKIP;
	}
};

template <int SIZE, typename G>
__device__ __forceinline__ float tile_response(const float* __restrict__ b) {
	typedef BoxGeo<SIZE> B;
	constexpr int bs = B::bs, bl = B::bl, rf = B::rf, rs = B::rs;
	constexpr float norm = 1.0f / (float)(SIZE * SIZE);
#define BS_T(dr, dc) b[G::off((dr), (dc))]
	// Dxx: rows top = -rs-1, bottom = top+bl; columns -rf-1 + {0, bs, 2bs, 3bs}
	constexpr int t = -rs - 1, bo = t + bl, c0 = -rf - 1;
	float Dxx = BS_T(bo, c0 + 3 * bs) - BS_T(t, c0 + 3 * bs) - BS_T(bo, c0) + BS_T(t, c0);
	Dxx -= 3.0f * (BS_T(bo, c0 + 2 * bs) - BS_T(t, c0 + 2 * bs) - BS_T(bo, c0 + bs) + BS_T(t, c0 + bs));
	// Dyy: columns left = -rs-1, right = left+bl; rows -rf-1 + {0, bs, 2bs, 3bs}
	constexpr int l = -rs - 1, r = l + bl, r0 = -rf - 1;
	float Dyy = BS_T(r0 + 3 * bs, r) - BS_T(r0 + 3 * bs, l) - BS_T(r0, r) + BS_T(r0, l);
	Dyy -= 3.0f * (BS_T(r0 + 2 * bs, r) - BS_T(r0 + 2 * bs, l) - BS_T(r0 + bs, r) + BS_T(r0 + bs, l));
	// Dxy: rows y1 = -bs-1, y2 = -1, y3 = 0, y4 = bs; columns -bs-1 + {0, bs, bs+1, 2bs+1}
	constexpr int y1 = -bs - 1, y2 = -1, y3 = 0, y4 = bs, x0 = -bs - 1, x1 = x0 + bs, x3 = x0 + bs + 1, x4 = x3 + bs;
	float Dxy = BS_T(y2, x1) - BS_T(y1, x1) - BS_T(y2, x0) + BS_T(y1, x0);
	Dxy -= BS_T(y2, x4) - BS_T(y1, x4) - BS_T(y2, x3) + BS_T(y1, x3);
	Dxy += BS_T(y4, x4) - BS_T(y3, x4) - BS_T(y4, x3) + BS_T(y3, x3);
	Dxy -= BS_T(y4, x1) - BS_T(y3, x1) - BS_T(y4, x0) + BS_T(y3, x0);
#undef BS_T
	Dxx *= norm; Dxy *= norm; Dyy *= norm;
	const float a = Dxx * Dyy;
	float bb = 0.81f * Dxy;
	bb = bb * Dxy;
	return a - bb;
}

// inner-region border of a layer (same as layer_border(), compile-time)
template <int SIZE, int SKIP> struct InnerBorder {
	static constexpr int rf = SIZE / 2;
	static constexpr int value = (rf + 1 + (SKIP - (rf + 1) % SKIP)) / SKIP;
};

template <int SKIP, int S0, int STEP, int NS, int TX, int TY, int I>
struct TileLayers {
	typedef TileGeo<SKIP, S0, STEP, NS, TX, TY> G;
	static __device__ __forceinline__ void run(const float* __restrict__ base, const float* __restrict__ ii, int W, int H,
	                                           const Layer* __restrict__ layers, float* __restrict__ inten, int x, int y) {
		constexpr int SIZE = S0 + I * STEP;
		constexpr int border = InnerBorder<SIZE, SKIP>::value;
		const Layer& L = layers[I];
		const bool inner = x >= border && x < L.w - border && y >= border && y < L.h - border;
		float v;
		if (inner) v = tile_response<SIZE, G>(base);
		else v = response_border(ii, W, H, L, x * SKIP, y * SKIP);
		inten[L.offset + (long)y * L.w + x] = v;
		TileLayers<SKIP, S0, STEP, NS, TX, TY, I + 1>::run(base, ii, W, H, layers, inten, x, y);
	}
};
template <int SKIP, int S0, int STEP, int NS, int TX, int TY>
struct TileLayers<SKIP, S0, STEP, NS, TX, TY, NS> {
	static __device__ __forceinline__ void run(const float*, const float*, int, int, const Layer*, float*, int, int) {}
};

constexpr int TILE_THREADS = 256;

// grid = (tiles_x * tiles_y, frames); layers = the NS layers of this octave (kernel parameter: lives in the constant bank)
struct OctaveLayers { Layer l[MAX_SCALES]; };

template <int SKIP, int S0, int STEP, int NS, int TX, int TY>
__global__ void __launch_bounds__(TILE_THREADS)
hessian_tile_kernel(const float* __restrict__ ii_all, int W, int H, OctaveLayers oct, int tiles_x, long inten_per_frame,
                    float* __restrict__ inten_all) {
	typedef TileGeo<SKIP, S0, STEP, NS, TX, TY> G;
	__shared__ float tile[G::WORDS];
	const float* ii = ii_all + (long)blockIdx.y * W * H;
	float* inten = inten_all + (long)blockIdx.y * inten_per_frame;
	const int tyi = blockIdx.x / tiles_x, txi = blockIdx.x - tyi * tiles_x;
	const int X0 = txi * TX, Y0 = tyi * TY;                  // first sample of the tile
	const int gx0 = X0 * SKIP - G::HALO, gy0 = Y0 * SKIP - G::HALO;   // image pixel of staged (row 0, col 0)
	// stage the footprint: consecutive threads read consecutive pixels of a row; out-of-image pixels are never used by an
	// inner sample, they are filled with 0
	for (int idx = threadIdx.x; idx < G::TW * G::TH; idx += TILE_THREADS) {
		const int row = idx / G::TW, col = idx - row * G::TW;
		const int gx = gx0 + col, gy = gy0 + row;
		float v = 0.0f;
		if (gx >= 0 && gy >= 0 && gx < W && gy < H) v = __ldg(ii + (long)gy * W + gx);
		tile[row * G::PITCH + (col % SKIP) * G::PW + col / SKIP] = v;
	}
	__syncthreads();
	const int w = oct.l[0].w, h = oct.l[0].h;
	constexpr int ROWS_PER_PASS = TILE_THREADS / TX;
	const int tx = threadIdx.x % TX, ty0 = threadIdx.x / TX;
	const int x = X0 + tx;
	if (x >= w) return;
#pragma unroll
	for (int ty = ty0; ty < TY; ty += ROWS_PER_PASS) {
		const int y = Y0 + ty;
		if (y < h) {
			const float* base = tile + ty * SKIP * G::PITCH + tx;
			TileLayers<SKIP, S0, STEP, NS, TX, TY, 0>::run(base, ii, W, H, oct.l, inten, x, y);
		}
	}
}

template <int SKIP, int S0, int STEP, int NS, int TX, int TY>
void launch_hessian_tile(const float* d_ii, const Plan& plan, int first_layer, float* d_inten, int batch, cudaStream_t st) {
	OctaveLayers oct{};
	for (int i = 0; i < NS; i++) oct.l[i] = plan.layers[first_layer + i];
	const int tiles_x = div_up(oct.l[0].w, TX), tiles_y = div_up(oct.l[0].h, TY);
	if (tiles_x * tiles_y == 0) return;
	BS_LAUNCH((hessian_tile_kernel<SKIP, S0, STEP, NS, TX, TY>), dim3(tiles_x * tiles_y, batch), TILE_THREADS, 0, st, d_ii, plan.W,
	          plan.H, oct, tiles_x, plan.inten_per_frame, d_inten);
}

}  // namespace bs
