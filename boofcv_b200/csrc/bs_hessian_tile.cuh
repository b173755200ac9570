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
#include "bs_sm100.cuh"
#include <algorithm>
#include <cstdlib>

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

// RING: the sample may belong to the reference's border ring (`ring` true), whose boxes block_zero sums as br-tr-bl+tl; only the
// Dyy boxes differ from hessianInner's order (br-bl-tr+tl, ImplIntegralImageFeatureIntensity.java:185-186), so both sums are
// formed from the same eight loads and selected — no divergence between ring and inner lanes of a warp.
template <int SIZE, typename G, bool RING = false>
__device__ __forceinline__ float tile_response(const float* __restrict__ b, bool ring = false) {
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
	const float a_br = BS_T(r0 + 3 * bs, r), a_bl = BS_T(r0 + 3 * bs, l), a_tr = BS_T(r0, r), a_tl = BS_T(r0, l);
	const float b_br = BS_T(r0 + 2 * bs, r), b_bl = BS_T(r0 + 2 * bs, l), b_tr = BS_T(r0 + bs, r), b_tl = BS_T(r0 + bs, l);
	float Dyy = a_br - a_bl - a_tr + a_tl;
	Dyy -= 3.0f * (b_br - b_bl - b_tr + b_tl);
	if (RING) {
		float Dyy_ring = a_br - a_tr - a_bl + a_tl;
		Dyy_ring -= 3.0f * (b_br - b_tr - b_bl + b_tl);
		Dyy = ring ? Dyy_ring : Dyy;
	}
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

// the border ring is rare: keep ONE out-of-line copy of its (large) code instead of one per unrolled layer and row — the
// fully inlined kernels were 140-154 KB and ncu showed the warps starving on instruction fetch (stall_no_instruction 5.2)
static __device__ __noinline__ float tile_border_response(const float* __restrict__ ii, int W, int H, const Layer* __restrict__ L, int px, int py) {
	return response_border(ii, W, H, *L, px, py);
}

template <typename G, int SKIP, int S0, int STEP, int NS, int I>
struct TileLayers {
	static __device__ __forceinline__ void run(const float* __restrict__ base, const float* __restrict__ ii, int W, int H,
	                                           const Layer* __restrict__ layers, float* __restrict__ inten, int x, int y) {
		constexpr int SIZE = S0 + I * STEP;
		constexpr int border = InnerBorder<SIZE, SKIP>::value;
		const Layer& L = layers[I];
		const bool inner = x >= border && x < L.w - border && y >= border && y < L.h - border;
		float v;
		if (inner) v = tile_response<SIZE, G>(base);
		else v = tile_border_response(ii, W, H, &L, x * SKIP, y * SKIP);
		inten[L.offset + (long)y * L.w + x] = v;
		TileLayers<G, SKIP, S0, STEP, NS, I + 1>::run(base, ii, W, H, layers, inten, x, y);
	}
};
template <typename G, int SKIP, int S0, int STEP, int NS>
struct TileLayers<G, SKIP, S0, STEP, NS, NS> {
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
#pragma unroll 1
	for (int ty = ty0; ty < TY; ty += ROWS_PER_PASS) {
		const int y = Y0 + ty;
		if (y < h) {
			const float* base = tile + ty * SKIP * G::PITCH + tx;
			TileLayers<G, SKIP, S0, STEP, NS, 0>::run(base, ii, W, H, oct.l, inten, x, y);
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


// ---------------------------------------------------------------- TMA-staged variant
// ncu on the kernel above (profiles/r1_ncu_full_batch8_chunk8_tiled_hessian.txt, source page): 25 % (skip 1) and 60 % (skip 2)
// of the executed instructions are the staging loop (index division, bounds tests, de-interleaving stores), and nothing
// hides its L2 latency inside a CTA.  Here one thread issues cp.async.bulk.tensor for the tile of the NEXT work item while
// the CTA computes the current one; out-of-image pixels are zero-filled by the TMA unit.
//   skip 1: two TMA buffers, the responses are computed straight out of the one that has landed;
//   skip 2: one TMA buffer + one buffer of de-interleaved column planes (a short LDS.64 / 2 x STS pass), the next tile's TMA
//           is issued as soon as that pass has emptied the TMA buffer.
// Measured constraint (tools/probe/tma_probe.cu): the box must start at a multiple of 4 pixels (16 bytes), so the left halo
// is rounded up to 4 and the tensor map's element stride cannot be used to fetch the odd column plane.
// Needs W % 4 == 0 (row pitch multiple of 16 bytes); other widths use the kernel above.
template <int SKIP, int S0, int STEP, int NS, int TX, int TY> struct TmaGeo {
	static_assert(SKIP == 1 || SKIP == 2, "TMA variant is written for sample spacing 1 and 2");
	static_assert((TX * SKIP) % 4 == 0, "tiles must start at a multiple of 4 pixels");
	static constexpr int SMAX = S0 + (NS - 1) * STEP;
	static constexpr int HALO = SMAX / 2 + 1;
	static constexpr int HALO_X = (HALO + 3) & ~3;                                    // left halo, 16-byte aligned box start
	static constexpr int BOXW = (HALO_X + (TX - 1) * SKIP + HALO + 3) & ~3;           // box width in pixels
	static constexpr int TH = (TY - 1) * SKIP + 2 * HALO;
	static constexpr int BOX_WORDS = BOXW * TH;
	static constexpr int BUF = (BOX_WORDS + 31) & ~31;                                // 128-byte aligned buffers
	static constexpr int PW = BOXW / SKIP;                                            // plane row pitch (skip 2)
	static constexpr int PLANE = PW * TH;
	// word offset of pixel (px+dc, py+dr) relative to the thread's base (= sample (tx,ty) of the tile)
	static __host__ __device__ constexpr int off(int dr, int dc) {
		return SKIP == 1 ? (HALO + dr) * BOXW + (HALO_X + dc) : (HALO + dr) * PW + ((HALO_X + dc) % SKIP) * PLANE + (HALO_X + dc) / SKIP;
	}
};

// Border samples out of the staged tile.  computeHessian / block_zero (ImplIntegralImageFeatureIntensity.java:115-127,
// ImplIntegralImageOps.java:195-214) reads 0 where a corner coordinate is negative and clamps coordinates beyond the last
// column / row.  The TMA unit zero-fills everything outside the image, so the negative side needs nothing; the far side
// reads the tile with clamped row / column indices.  The border ring cannot simply take the compile-time path above even
// where no clamp is needed: hessianInner sums the Dyy boxes as br-bl-tr+tl (:185-186) while block_zero sums every box as
// br-tr-bl+tl, and the two round differently.  One out-of-line copy, runtime box size.  cr, cc = tile row / column of the
// sample's pixel; rmax, cmax = tile row / column of the image's last row / column.
template <typename G, int SKIP>
static __device__ __noinline__ float tile_clamped_response(const float* __restrict__ tile, int size, int cr, int cc, int rmax, int cmax) {
	const int bs = size / 3, bl = size - bs - 1, rf = size / 2, rs = bl / 2;
	const float norm = 1.0f / (float)(size * size);
	auto at = [&](int dr, int dc) -> float {
		const int r = min(cr + dr, rmax), c = min(cc + dc, cmax);
		return SKIP == 1 ? tile[r * G::BOXW + c] : tile[r * G::PW + (c % SKIP) * G::PLANE + c / SKIP];
	};
	const int t = -rs - 1, bo = t + bl, c0 = -rf - 1;
	float Dxx = at(bo, c0 + 3 * bs) - at(t, c0 + 3 * bs) - at(bo, c0) + at(t, c0);
	Dxx -= 3.0f * (at(bo, c0 + 2 * bs) - at(t, c0 + 2 * bs) - at(bo, c0 + bs) + at(t, c0 + bs));
	const int l = -rs - 1, r = l + bl, r0 = -rf - 1;
	float Dyy = at(r0 + 3 * bs, r) - at(r0, r) - at(r0 + 3 * bs, l) + at(r0, l);                     // block_zero order: br - tr - bl + tl
	Dyy -= 3.0f * (at(r0 + 2 * bs, r) - at(r0 + bs, r) - at(r0 + 2 * bs, l) + at(r0 + bs, l));
	const int y1 = -bs - 1, y2 = -1, y3 = 0, y4 = bs, x0 = -bs - 1, x1 = x0 + bs, x3 = x0 + bs + 1, x4 = x3 + bs;
	float Dxy = at(y2, x1) - at(y1, x1) - at(y2, x0) + at(y1, x0);
	Dxy -= at(y2, x4) - at(y1, x4) - at(y2, x3) + at(y1, x3);
	Dxy += at(y4, x4) - at(y3, x4) - at(y4, x3) + at(y3, x3);
	Dxy -= at(y4, x1) - at(y3, x1) - at(y4, x0) + at(y3, x0);
	Dxx *= norm; Dxy *= norm; Dyy *= norm;
	const float a = Dxx * Dyy;
	float bb = 0.81f * Dxy;
	bb = bb * Dxy;
	return a - bb;
}

// EDGE = the tile holds samples of the reference's border ring of some layer (uniform per tile)
template <typename G, int SKIP, int S0, int STEP, int NS, int I, bool EDGE>
struct TmaLayers {
	static __device__ __forceinline__ void run(const float* __restrict__ base, const float* __restrict__ tile, int cr, int cc, int rmax, int cmax,
	                                           const Layer* __restrict__ layers, float* __restrict__ inten, int x, int y, int at) {
		constexpr int SIZE = S0 + I * STEP;
		constexpr int border = InnerBorder<SIZE, SKIP>::value;
		const Layer& L = layers[I];
		float v;
		if (EDGE) {
			// ring samples whose boxes stay left of / above the last column / row need no clamp: TMA's zero fill is block_zero's
			// zero, only the summation order is the ring's; the clamped path is left to the right / bottom edge
			const bool ring = !(x >= border && x < L.w - border && y >= border && y < L.h - border);
			if (cc + SIZE / 2 > cmax || cr + SIZE / 2 > rmax) v = tile_clamped_response<G, SKIP>(tile, SIZE, cr, cc, rmax, cmax);
			else v = tile_response<SIZE, G, true>(base, ring);
		} else {
			v = tile_response<SIZE, G>(base);
		}
		(inten + L.offset)[at] = v;                // at = y * w + x: every layer of an octave has the same dimensions
		TmaLayers<G, SKIP, S0, STEP, NS, I + 1, EDGE>::run(base, tile, cr, cc, rmax, cmax, layers, inten, x, y, at);
	}
};
template <typename G, int SKIP, int S0, int STEP, int NS, bool EDGE>
struct TmaLayers<G, SKIP, S0, STEP, NS, NS, EDGE> {
	static __device__ __forceinline__ void run(const float*, const float*, int, int, int, int, const Layer*, float*, int, int, int) {}
};

template <int SKIP, int S0, int STEP, int NS, int TX, int TY, int THREADS>
__global__ void __launch_bounds__(THREADS)
hessian_tma_kernel(const __grid_constant__ CUtensorMap tm, const float* __restrict__ ii_all, int W, int H, OctaveLayers oct, int tiles_x,
                   int tiles_y, int frames, long inten_per_frame, float* __restrict__ inten_all) {
	typedef TmaGeo<SKIP, S0, STEP, NS, TX, TY> G;
	extern __shared__ unsigned char tma_smem_raw[];
	float* buf = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(tma_smem_raw) + 127) & ~(uintptr_t)127);
	uint64_t* full = reinterpret_cast<uint64_t*>(buf + 2 * G::BUF);            // [2]
	const CUtensorMap* tmap = &tm;                 // address of the kernel parameter itself (a copy is not a valid descriptor)
	const int per_frame = tiles_x * tiles_y;
	const long total = (long)per_frame * frames;
	if (threadIdx.x == 0) {
		sm100::mbar_init(&full[0], 1); sm100::mbar_init(&full[1], 1);
		sm100::mbar_fence_init();
	}
	__syncthreads();
	auto issue = [=](long item, int b) {           // thread 0 only: TMA of work item `item` into TMA buffer b
		const unsigned it = (unsigned)item;        // total < 2^31 (checked by the launcher): 32-bit divisions
		const int frame = (int)(it / (unsigned)per_frame), t = (int)(it - (unsigned)frame * (unsigned)per_frame);
		const int tyi = (int)((unsigned)t / (unsigned)tiles_x), txi = t - tyi * tiles_x;
		sm100::mbar_expect_tx(&full[b], G::BOX_WORDS * 4);
		sm100::tma_load_3d(buf + b * G::BUF, tmap, txi * TX * SKIP - G::HALO_X, tyi * TY * SKIP - G::HALO, frame, &full[b]);
	};
	if (threadIdx.x == 0 && (long)blockIdx.x < total) issue(blockIdx.x, 0);
	const int w = oct.l[0].w, h = oct.l[0].h;
	constexpr int ROWS_PER_PASS = THREADS / TX;
	const int tx = threadIdx.x % TX, ty0 = threadIdx.x / TX;
	uint32_t n = 0;
	for (long item = blockIdx.x; item < total; item += gridDim.x, n++) {
		const float* tile;
		if (SKIP == 1) {
			const int b = n & 1;
			// buffer b^1 was last read in the previous iteration, which ended with __syncthreads
			if (threadIdx.x == 0 && item + gridDim.x < total) issue(item + gridDim.x, b ^ 1);
			sm100::mbar_wait(&full[b], (n >> 1) & 1);
			tile = buf + b * G::BUF;
		} else {
			// TMA buffer = buf, planes = buf + BUF
			sm100::mbar_wait(&full[0], n & 1);
			float* planes = buf + G::BUF;
			// BOXW = 2 PW: float2 i of the TMA buffer holds columns (2j, 2j+1) of row r with i = r PW + j, which is also the
			// element's place inside either plane — the de-interleave is a flat copy
			const float2* src = reinterpret_cast<const float2*>(buf);
#pragma unroll 4
			for (int i = threadIdx.x; i < G::PLANE; i += THREADS) {
				const float2 v = src[i];
				planes[i] = v.x;
				planes[G::PLANE + i] = v.y;
			}
			__syncthreads();                       // TMA buffer drained, planes complete
			if (threadIdx.x == 0 && item + gridDim.x < total) issue(item + gridDim.x, 0);
			tile = planes;
		}
		const unsigned it = (unsigned)item;
		const int frame = (int)(it / (unsigned)per_frame), t = (int)(it - (unsigned)frame * (unsigned)per_frame);
		const int tyi = (int)((unsigned)t / (unsigned)tiles_x), txi = t - tyi * tiles_x;
		float* inten = inten_all + (long)frame * inten_per_frame;
		const int x = txi * TX + tx;
		// tile row / column of the image's last row / column; edge tile = it reaches into the widest border ring of the octave
		const int rmax = H - 1 - (tyi * TY * SKIP - G::HALO), cmax = W - 1 - (txi * TX * SKIP - G::HALO_X);
		constexpr int MB = InnerBorder<G::SMAX, SKIP>::value;
		const bool edge = txi * TX < MB || txi * TX + TX > w - MB || tyi * TY < MB || tyi * TY + TY > h - MB;
		constexpr int ROWP = SKIP == 1 ? G::BOXW : G::PW;
		if (x < w) {
			if (!edge) {
#pragma unroll 1
				for (int ty = ty0; ty < TY; ty += ROWS_PER_PASS) {
					const int y = tyi * TY + ty;
					if (y < h) TmaLayers<G, SKIP, S0, STEP, NS, 0, false>::run(tile + ty * SKIP * ROWP + tx, tile, 0, 0, 0, 0, oct.l, inten, x, y, y * w + x);
				}
			} else {
#pragma unroll 1
				for (int ty = ty0; ty < TY; ty += ROWS_PER_PASS) {
					const int y = tyi * TY + ty;
					if (y < h) TmaLayers<G, SKIP, S0, STEP, NS, 0, true>::run(tile + ty * SKIP * ROWP + tx, tile, G::HALO + ty * SKIP, G::HALO_X + tx * SKIP,
					                                                           rmax, cmax, oct.l, inten, x, y, y * w + x);
				}
			}
		}
		__syncthreads();
	}
}

typedef CUresult (*HessianEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                         const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                         CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int SKIP, int S0, int STEP, int NS, int TX, int TY, int THREADS = TILE_THREADS>
bool launch_hessian_tma(const float* d_ii, const Plan& plan, int first_layer, float* d_inten, int batch, cudaStream_t st) {
	typedef TmaGeo<SKIP, S0, STEP, NS, TX, TY> G;
	if ((plan.W & 3) != 0 || (reinterpret_cast<uintptr_t>(d_ii) & 15) != 0) return false;
	static HessianEncodeTiledFn encode = nullptr;
	if (!encode) {
		void* p = nullptr;
		cudaDriverEntryPointQueryResult q;
		if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p || q != cudaDriverEntryPointSuccess) {
			cudaGetLastError();
			return false;
		}
		encode = reinterpret_cast<HessianEncodeTiledFn>(p);
	}
	OctaveLayers oct{};
	for (int i = 0; i < NS; i++) oct.l[i] = plan.layers[first_layer + i];
	const int tiles_x = div_up(oct.l[0].w, TX), tiles_y = div_up(oct.l[0].h, TY);
	if (tiles_x * tiles_y == 0) return true;
	CUtensorMap tm;
	const cuuint64_t dims[3] = {(cuuint64_t)plan.W, (cuuint64_t)plan.H, (cuuint64_t)batch};
	const cuuint64_t strides[2] = {(cuuint64_t)plan.W * 4, (cuuint64_t)plan.W * plan.H * 4};
	const cuuint32_t box[3] = {(cuuint32_t)G::BOXW, (cuuint32_t)G::TH, 1};
	const cuuint32_t estr[3] = {1, 1, 1};
	if (encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(d_ii), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
	           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
		return false;
	constexpr int smem = 2 * G::BUF * 4 + 16 + 128;
	auto kern = hessian_tma_kernel<SKIP, S0, STEP, NS, TX, TY, THREADS>;
	const int sm_count = device_sm_count();
	BS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));   // per device: set on every launch
	int per_sm = (200 * 1024) / (smem + 1024);
	static const int per_sm_cap = getenv("BS_HT_PER_SM") ? atoi(getenv("BS_HT_PER_SM")) : 4;
	per_sm = per_sm > per_sm_cap ? per_sm_cap : (per_sm < 1 ? 1 : per_sm);
	const long total = (long)tiles_x * tiles_y * batch;
	if (total >= (1L << 31)) return false;
	const int grid = (int)std::min<long>(total, (long)sm_count * per_sm);
	BS_LAUNCH(kern, grid, THREADS, smem, st, tm, d_ii, plan.W, plan.H, oct, tiles_x, tiles_y, batch, plan.inten_per_frame, d_inten);
	return true;
}

}  // namespace bs
