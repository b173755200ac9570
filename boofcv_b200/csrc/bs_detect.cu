// bs_detect.cu — non-maximum suppression, scale-space maximum test, sub-pixel refinement and ordered compaction.
// Reference:
//   NonMaxBlock.process (F/alg/feature/detect/extract/NonMaxBlock.java:69-94) with
//   NonMaxBlockSearchStrict.Max.searchBlock / checkLocalMax (…/NonMaxBlockSearchStrict.java:56-79,196-221);
//   FastHessianFeatureDetector.findLocalScaleSpaceMax / checkMax / polyPeak
//   (F/alg/feature/detect/interest/FastHessianFeatureDetector.java:230-298,304-313,336-350).
//
// One thread owns one (radius+1)^2 NMS block — the reference emits at most one point per block, in block-raster order
// per level, levels in (octave, level) order.  A CTA owns NMS_CTA_BLOCKS consecutive blocks of one level, so CTA order ==
// emission order; points are staged per CTA (rank = prefix count inside the CTA) and a scan over the CTA counts gives every
// point its slot in the reference's single-thread order.  No atomics, no sort: the output order is deterministic.
#include "bs_common.cuh"

#include <cfloat>
#include <math_constants.h>

namespace bs {

// polyPeak(float,float,float) :336-350
__device__ __forceinline__ float poly_peak(float lower, float middle, float upper) {
	const float a = 0.5f * lower - middle + 0.5f * upper;
	const float b = 0.5f * upper - 0.5f * lower;
	if (a == 0.0f) return 0.0f;
	return -b / (2.0f * a);
}

// checkMax :304-313 on an ImageBorderValue(0) view: out-of-image reads are 0
__device__ __forceinline__ bool scale_neighbour_ok(const float* __restrict__ img, int w, int h, float best, int cx, int cy) {
	for (int y = cy - 1; y <= cy + 1; y++)
		for (int x = cx - 1; x <= cx + 1; x++) {
			const float v = (x >= 0 && y >= 0 && x < w && y < h) ? __ldg(img + (long)y * w + x) : 0.0f;
			if (v >= best) return false;
		}
	return true;
}

enum { NMS_PLAIN = 0, NMS_KEYS = 1, NMS_KEPT = 2 };

// order-preserving map float -> uint32 (larger response = larger key); 0 is never produced by a finite value
__device__ __forceinline__ unsigned response_key(float v) {
	const unsigned b = __float_as_uint(v);
	return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// 32 CTAs of 64 threads per SM (32 registers): the kernel waits on dependent loads, so resident warps are what it lives on
// (nms_refine 15.7 / 13.0 / 11.5 ms per 1024 frames at 1 / 24 / 32 CTAs per SM)
#ifndef BS_NMS_MINBLOCKS
#define BS_NMS_MINBLOCKS 32
#endif
__global__ void __launch_bounds__(NMS_CTA_BLOCKS, BS_NMS_MINBLOCKS)
nms_kernel(const Plan* __restrict__ plan, const float* __restrict__ inten_all, RawPoint* __restrict__ stage_all,
           int* __restrict__ counts_all, int mode, unsigned* __restrict__ keys_all, const unsigned char* __restrict__ keep_all) {
	__shared__ int warp_counts[NMS_CTA_BLOCKS / 32];
	const int cta = blockIdx.x, frame = blockIdx.y;
	int lv = 0;
	while (lv + 1 < plan->n_levels && cta >= plan->levels[lv + 1].cta_begin) lv++;
	const Level& V = plan->levels[lv];
	const Layer& M = plan->layers[V.mid];
	const int w = M.w, h = M.h;
	const float* inten = inten_all + (long)frame * plan->inten_per_frame;
	const float* mid = inten + M.offset;
	const int radius = plan->radius, step = radius + 1;
	const int bid = (cta - V.cta_begin) * NMS_CTA_BLOCKS + threadIdx.x;

	bool found = false;
	RawPoint pt;
	// the extractor: at most one strict maximum per block
	bool ok = false;
	int peakX = 0, peakY = 0;
	float peak = -FLT_MAX;
	if (bid < V.nbx * V.nby) {
		const int by = bid / V.nbx, bx = bid - by * V.nbx;
		const int endX = w - V.border, endY = h - V.border;
		const int x0 = V.border + bx * step, y0 = V.border + by * step;
		const int x1 = min(x0 + step, endX), y1 = min(y0 + step, endY);
		// searchBlock: first strict maximum in raster order
		if (step == 3 && x1 - x0 == 3 && y1 - y0 == 3) {
			// the common case (radius 2, full block): nine loads in flight at once instead of a dependent loop
			const float* p = mid + (long)y0 * w + x0;
			float v[9];
#pragma unroll
			for (int k = 0; k < 9; k++) v[k] = __ldg(p + (k / 3) * w + (k % 3));
#pragma unroll
			for (int k = 0; k < 9; k++)
				if (v[k] > peak) { peak = v[k]; peakX = x0 + k % 3; peakY = y0 + k / 3; }
		} else {
			for (int y = y0; y < y1; y++)
				for (int x = x0; x < x1; x++) {
					const float v = __ldg(mid + (long)y * w + x);
					if (v > peak) { peak = v; peakX = x; peakY = y; }
				}
		}
		ok = peak >= plan->threshold && peak != FLT_MAX;
		if (ok) {
			// Part of checkLocalMax, done per thread before the warp-wide window test: the four direct neighbours (most block
			// maxima that are not window maxima lose against one of them).  Every test must pass for a point to be emitted, so
			// their order is free: the two cheapest reads of checkMax (the same pixel in the neighbouring layers,
			// FastHessianFeatureDetector.java:304-313) are fetched in the same round trip — not in the n-best modes, where the
			// extractor's own output is needed.
			const long c = (long)peakY * w + peakX;
			const float NONE = -CUDART_INF_F;
			const bool plain = mode == NMS_PLAIN;
			const float nl = peakX > 0 ? __ldg(mid + c - 1) : NONE, nr = peakX + 1 < w ? __ldg(mid + c + 1) : NONE;
			const float nu = peakY > 0 ? __ldg(mid + c - w) : NONE, nd = peakY + 1 < h ? __ldg(mid + c + w) : NONE;
			const float sl = plain ? __ldg(inten + plan->layers[V.lower].offset + c) : NONE;
			const float su = plain ? __ldg(inten + plan->layers[V.upper].offset + c) : NONE;
			if (nl >= peak || nr >= peak || nu >= peak || nd >= peak || sl >= peak || su >= peak) ok = false;
		}
	}
	// From here on the warp works on one candidate at a time, all 32 lanes together: ncu showed the per-thread version at 12
	// active lanes per instruction, i.e. every warp paid the 25-pixel window, the 18 scale neighbours and the refinement of its
	// few candidates serially.
	const int lane = threadIdx.x & 31;
	{
		// checkLocalMax: strictly greater than everything else in the (2r+1)^2 window clamped to the image
		unsigned todo = __ballot_sync(0xffffffffu, ok);
		while (todo) {
			const int src = __ffs(todo) - 1;
			todo &= todo - 1;
			const int cx = __shfl_sync(0xffffffffu, peakX, src), cy = __shfl_sync(0xffffffffu, peakY, src);
			const float val = __shfl_sync(0xffffffffu, peak, src);
			const int wx0 = max(cx - radius, 0), wx1 = min(cx + radius, w - 1);
			const int wy0 = max(cy - radius, 0), wy1 = min(cy + radius, h - 1);
			const int ww = wx1 - wx0 + 1, n = ww * (wy1 - wy0 + 1);
			bool bad = false;
			for (int k = lane; k < n; k += 32) {
				const int y = wy0 + k / ww, x = wx0 + k % ww;
				const float v = __ldg(mid + (long)y * w + x);
				bad = bad || (v >= val && !(x == cx && y == cy));
			}
			if (__any_sync(0xffffffffu, bad) && lane == src) ok = false;
		}
	}
	if (mode != NMS_PLAIN) {
		// SelectNBestFeatures works on the extractor's output, i.e. on what passed the tests above (whole CTA takes this branch)
		const long slot = ((long)frame * plan->ctas_per_frame + cta) * NMS_CTA_BLOCKS + threadIdx.x;
		if (mode == NMS_KEYS) { keys_all[slot] = ok ? response_key(peak) : 0u; return; }
		ok = ok && keep_all[slot] != 0;
	}
	if (ok) {
		// findLocalScaleSpaceMax :248-294
		const int ir = V.border + radius;
		if (peakX < ir || peakX >= w - ir || peakY < ir || peakY >= h - ir) ok = false;
	}
	{
		const float* lower = inten + plan->layers[V.lower].offset;
		const float* upper = inten + plan->layers[V.upper].offset;
		unsigned todo = __ballot_sync(0xffffffffu, ok);
		while (todo) {
			const int src = __ffs(todo) - 1;
			todo &= todo - 1;
			const int cx = __shfl_sync(0xffffffffu, peakX, src), cy = __shfl_sync(0xffffffffu, peakY, src);
			const float val = __shfl_sync(0xffffffffu, peak, src);
			// checkMax :304-313 in both neighbouring layers: lanes 0-8 the lower one, 9-17 the upper one; out-of-image reads are 0
			bool bad = false;
			if (lane < 18) {
				const int k = lane < 9 ? lane : lane - 9;
				const int x = cx - 1 + k % 3, y = cy - 1 + k / 3;
				const float* img = lane < 9 ? lower : upper;
				const float v = (x >= 0 && y >= 0 && x < w && y < h) ? __ldg(img + (long)y * w + x) : 0.0f;
				bad = v >= val;
			}
			if (__any_sync(0xffffffffu, bad)) continue;
			if (lane == src) {
				const long c = (long)cy * w + cx;
				const float peakXf = poly_peak(__ldg(mid + c - 1), val, __ldg(mid + c + 1));
				const float peakYf = poly_peak(__ldg(mid + c - w), val, __ldg(mid + c + w));
				const float peakS = poly_peak(__ldg(lower + c), val, __ldg(upper + c));
				const float interpX = ((float)cx + peakXf) * (float)M.skip;
				const float interpY = ((float)cy + peakYf) * (float)M.skip;
				const float interpS = (float)V.level_size + peakS * (float)V.level_step;
				pt.x = (double)interpX; pt.y = (double)interpY;
				pt.scale = 1.2 * (double)interpS / 9.0;
				pt.response = val;
				pt.px = (int16_t)cx; pt.py = (int16_t)cy;
				pt.octave = (int16_t)V.octave; pt.level = (int16_t)V.level;
				pt.pad = 0;
				found = true;
			}
		}
	}
	// rank inside the CTA in thread (= block-raster) order
	const unsigned ballot = __ballot_sync(0xffffffffu, found);
	const int warp = threadIdx.x >> 5;
	if (lane == 0) warp_counts[warp] = __popc(ballot);
	__syncthreads();
	int base = 0, total = 0;
	for (int k = 0; k < NMS_CTA_BLOCKS / 32; k++) {
		const int c = warp_counts[k];
		if (k < warp) base += c;
		total += c;
	}
	const long slot0 = ((long)frame * plan->ctas_per_frame + cta) * NMS_CTA_BLOCKS;
	if (found) stage_all[slot0 + base + __popc(ballot & ((1u << lane) - 1u))] = pt;
	if (threadIdx.x == 0) counts_all[(long)frame * plan->ctas_per_frame + cta] = total;
}

// ---------------------------------------------------------------- SelectNBestFeatures
// Reference: SelectNBestFeatures.process (F/alg/feature/detect/extract/SelectNBestFeatures.java:51-93), applied per detect
// level to the extractor's maxima before the scale-space test (FastHessianFeatureDetector.java:257-260).  If a level has more
// than N maxima the N with the largest response survive.  Which of several EQUAL responses survives is decided by ddogleg's
// QuickSelect in the reference (unpinned); here ties go to the earlier block (the rule the parity tests check), and
// the survivors keep their block-raster order.
// One CTA per (level, frame): 4-pass radix select of the N-th largest key, then an ordered pass that hands out the quota of
// keys equal to it.
constexpr int SELECT_THREADS = 1024;

__global__ void __launch_bounds__(SELECT_THREADS)
select_nbest_kernel(const Plan* __restrict__ plan, const unsigned* __restrict__ keys_all, unsigned char* __restrict__ keep_all) {
	__shared__ unsigned hist[256];
	__shared__ unsigned s_prefix, s_remaining, s_total;
	__shared__ int warp_sums[SELECT_THREADS / 32];
	__shared__ int s_carry;
	const int lv = blockIdx.x, frame = blockIdx.y;
	const int cta0 = plan->levels[lv].cta_begin;
	const int cta1 = lv + 1 < plan->n_levels ? plan->levels[lv + 1].cta_begin : plan->ctas_per_frame;
	const long first = ((long)frame * plan->ctas_per_frame + cta0) * NMS_CTA_BLOCKS;
	const int n = (cta1 - cta0) * NMS_CTA_BLOCKS;
	const unsigned* keys = keys_all + first;
	unsigned char* keep = keep_all + first;
	const unsigned N = (unsigned)plan->max_per_level;
	const int tid = threadIdx.x;

	// number of maxima on the level
	if (tid == 0) s_total = 0;
	__syncthreads();
	unsigned mine = 0;
	for (int i = tid; i < n; i += SELECT_THREADS) mine += keys[i] != 0u;
	if (mine) atomicAdd(&s_total, mine);
	__syncthreads();
	if (s_total <= N) {
		for (int i = tid; i < n; i += SELECT_THREADS) keep[i] = keys[i] != 0u;
		return;
	}
	// radix select, most significant byte first: after the loop s_prefix is the N-th largest key and s_remaining how many
	// keys EQUAL to it are still wanted
	if (tid == 0) { s_prefix = 0; s_remaining = N; }
	for (int shift = 24; shift >= 0; shift -= 8) {
		if (tid < 256) hist[tid] = 0;
		__syncthreads();
		const unsigned prefix = s_prefix;
		const unsigned himask = shift == 24 ? 0u : (0xffffffffu << (shift + 8));
		for (int i = tid; i < n; i += SELECT_THREADS) {
			const unsigned k = keys[i];
			if (k != 0u && (k & himask) == prefix) atomicAdd(&hist[(k >> shift) & 255u], 1u);
		}
		__syncthreads();
		if (tid == 0) {
			unsigned want = s_remaining;
			int b = 255;
			for (; b > 0; b--) {
				if (hist[b] >= want) break;
				want -= hist[b];
			}
			s_prefix = prefix | ((unsigned)b << shift);
			s_remaining = want;
		}
		__syncthreads();
	}
	const unsigned T = s_prefix;
	const int quota = (int)s_remaining;
	// ordered pass: keys above T survive; of the keys equal to T the first `quota` in block order
	if (tid == 0) s_carry = 0;
	__syncthreads();
	const int lane = tid & 31, warp = tid >> 5;
	for (int i0 = 0; i0 < n; i0 += SELECT_THREADS) {
		const int i = i0 + tid;
		const unsigned k = i < n ? keys[i] : 0u;
		const bool eq = k != 0u && k == T;
		const unsigned ballot = __ballot_sync(0xffffffffu, eq);
		if (lane == 0) warp_sums[warp] = __popc(ballot);
		__syncthreads();
		int before = s_carry, chunk = 0;
		for (int w = 0; w < SELECT_THREADS / 32; w++) {
			const int c = warp_sums[w];
			if (w < warp) before += c;
			chunk += c;
		}
		before += __popc(ballot & ((1u << lane) - 1u));
		if (i < n) keep[i] = (k != 0u && k > T) || (eq && before < quota);
		__syncthreads();
		if (tid == 0) s_carry += chunk;
		__syncthreads();
	}
}

void launch_nms(const Plan& plan, const Plan* d_plan, const float* d_inten, int batch, RawPoint* d_stage, int* d_cta_counts,
                unsigned* d_keys, unsigned char* d_keep, cudaStream_t st) {
	if (plan.ctas_per_frame == 0 || batch <= 0) return;
	const dim3 grid(plan.ctas_per_frame, batch);
	if (plan.max_per_level > 0) {
		if (!d_keys || !d_keep) fail(BS_ERR_STATE, "n-best selection needs its workspace");
		BS_LAUNCH(nms_kernel, grid, NMS_CTA_BLOCKS, 0, st, d_plan, d_inten, d_stage, d_cta_counts, (int)NMS_KEYS, d_keys,
		          (const unsigned char*)d_keep);
		BS_LAUNCH(select_nbest_kernel, dim3(plan.n_levels, batch), SELECT_THREADS, 0, st, d_plan, (const unsigned*)d_keys, d_keep);
		BS_LAUNCH(nms_kernel, grid, NMS_CTA_BLOCKS, 0, st, d_plan, d_inten, d_stage, d_cta_counts, (int)NMS_KEPT, d_keys,
		          (const unsigned char*)d_keep);
	} else {
		BS_LAUNCH(nms_kernel, grid, NMS_CTA_BLOCKS, 0, st, d_plan, d_inten, d_stage, d_cta_counts, (int)NMS_PLAIN, (unsigned*)nullptr,
		          (const unsigned char*)nullptr);
	}
}

// ---------------------------------------------------------------- ordered compaction
constexpr int SCAN_THREADS = 1024;

// one CTA per frame: exclusive scan of the frame's per-CTA counts (relative offsets) and the frame total
__global__ void __launch_bounds__(SCAN_THREADS)
scan_counts_kernel(const int* __restrict__ counts, int ctas, int* __restrict__ offsets, int* __restrict__ frame_total) {
	__shared__ int warp_sums[SCAN_THREADS / 32];
	__shared__ int carry;
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int f = blockIdx.x;
	if (threadIdx.x == 0) carry = 0;
	__syncthreads();
	constexpr int ITEMS = 4;                   // consecutive counts per thread and pass
	for (int c0 = 0; c0 < ctas; c0 += SCAN_THREADS * ITEMS) {
		const int c = c0 + threadIdx.x * ITEMS;
		int v[ITEMS], mine = 0;
#pragma unroll
		for (int k = 0; k < ITEMS; k++) { v[k] = c + k < ctas ? counts[(long)f * ctas + c + k] : 0; mine += v[k]; }
		int incl = mine;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const int t = __shfl_up_sync(0xffffffffu, incl, d);
			if (lane >= d) incl += t;
		}
		if (lane == 31) warp_sums[warp] = incl;
		__syncthreads();
		int wbase = 0, chunk_total = 0;
		for (int k = 0; k < SCAN_THREADS / 32; k++) {
			const int s = warp_sums[k];
			if (k < warp) wbase += s;
			chunk_total += s;
		}
		const int base = carry;
		int run = base + wbase + incl - mine;
#pragma unroll
		for (int k = 0; k < ITEMS; k++) {
			if (c + k < ctas) offsets[(long)f * ctas + c + k] = run;
			run += v[k];
		}
		__syncthreads();
		if (threadIdx.x == 0) carry = base + chunk_total;
		__syncthreads();
	}
	if (threadIdx.x == 0) frame_total[f] = carry;
}

// frame_offset[frame0 + f + 1] = frame_offset[frame0] + sum of the totals of frames 0..f of the chunk (one warp)
__global__ void frame_offsets_kernel(const int* __restrict__ frame_total, int batch, int frame0, int* __restrict__ frame_offset) {
	const int lane = threadIdx.x;
	int carry = frame_offset[frame0];
	for (int f0 = 0; f0 < batch; f0 += 32) {
		const int f = f0 + lane;
		const int v = f < batch ? frame_total[f] : 0;
		int incl = v;
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			const int t = __shfl_up_sync(0xffffffffu, incl, d);
			if (lane >= d) incl += t;
		}
		if (f < batch) frame_offset[frame0 + f + 1] = carry + incl;
		carry += __shfl_sync(0xffffffffu, incl, 31);
	}
}

// one thread per staging slot; a CTA covers GATHER_THREADS / NMS_CTA_BLOCKS consecutive NMS CTAs of one frame
constexpr int GATHER_THREADS = NMS_CTA_BLOCKS > 256 ? NMS_CTA_BLOCKS : 256;
__global__ void __launch_bounds__(GATHER_THREADS)
gather_points_kernel(const RawPoint* __restrict__ stage, const int* __restrict__ counts, const int* __restrict__ offsets,
                     const int* __restrict__ frame_offset, int ctas, int frame0, Pool pool) {
	constexpr int PER = GATHER_THREADS / NMS_CTA_BLOCKS;
	const int sub = threadIdx.x / NMS_CTA_BLOCKS, slot = threadIdx.x - sub * NMS_CTA_BLOCKS;
	const int cta = blockIdx.x * PER + sub;
	if (cta >= ctas) return;
	const long idx = (long)blockIdx.y * ctas + cta;
	const int n = counts[idx];
	if (slot >= n) return;
	const long dst = (long)frame_offset[frame0 + blockIdx.y] + offsets[idx] + slot;
	if (dst >= pool.cap) return;      // counted, not stored: the host reports BS_ERR_CAPACITY
	const RawPoint p = stage[idx * NMS_CTA_BLOCKS + slot];
	pool.xy[2 * dst] = p.x; pool.xy[2 * dst + 1] = p.y;
	pool.scale[dst] = p.scale;
	pool.meta[4 * dst] = p.octave; pool.meta[4 * dst + 1] = p.level; pool.meta[4 * dst + 2] = p.px; pool.meta[4 * dst + 3] = p.py;
	pool.response[dst] = p.response;
	pool.frame[dst] = frame0 + blockIdx.y;
}

void launch_compact(const Plan& plan, int batch, int frame0, const RawPoint* d_stage, const int* d_cta_counts, int* d_cta_offsets,
                    int* d_frame_offset, const Pool& pool, cudaStream_t st) {
	if (batch <= 0) return;
	// the per-CTA offset array has one spare int per frame at its end: the frame totals
	int* d_frame_total = d_cta_offsets + (long)batch * plan.ctas_per_frame;
	BS_LAUNCH(scan_counts_kernel, batch, SCAN_THREADS, 0, st, d_cta_counts, plan.ctas_per_frame, d_cta_offsets, d_frame_total);
	BS_LAUNCH(frame_offsets_kernel, 1, 32, 0, st, (const int*)d_frame_total, batch, frame0, d_frame_offset);
	if (plan.ctas_per_frame == 0) return;
	BS_LAUNCH(gather_points_kernel, dim3(div_up(plan.ctas_per_frame, GATHER_THREADS / NMS_CTA_BLOCKS), batch), GATHER_THREADS, 0, st, d_stage, d_cta_counts, (const int*)d_cta_offsets,
	          (const int*)d_frame_offset, plan.ctas_per_frame, frame0, pool);
}

// ---------------------------------------------------------------- stand-alone NonMaxBlock (strict rule, maxima)
// NonMaxBlock.process :69-94 with NonMaxBlockSearchStrict.Max (searchBlock :56-79, checkLocalMax :196-221) on ONE intensity
// image: what FactoryFeatureExtractor.nonmax(ConfigExtract(radius, threshold, ignoreBorder, useStrictRule = true)) builds and
// what BOverrideFactoryFeatureExtractor.nonmax can hand to a plug-in (F/factory/feature/detect/extract/FactoryFeatureExtractor.java:
// 69-71).  One thread per (radius+1)^2 block; the survivors are compacted in block-row-major order (the reference's single-thread
// emission order) by one CTA.
__global__ void nonmax_block_kernel(const float* __restrict__ img, int stride, int W, int H, int radius, float thr, int border, int bw, int bh,
                                    int* __restrict__ cand) {
	const int b = blockIdx.x * blockDim.x + threadIdx.x;
	if (b >= bw * bh) return;
	const int by = b / bw, bx = b - by * bw;
	const int step = radius + 1;
	const int endX = W - border, endY = H - border;
	const int x = border + bx * step, y = border + by * step;
	const int x1 = min(x + step, endX), y1 = min(y + step, endY);
	int peakX = 0, peakY = 0;
	float peak = -FLT_MAX;
	for (int yy = y; yy < y1; yy++)
		for (int xx = x; xx < x1; xx++) {
			const float v = __ldg(img + (long)yy * stride + xx);
			if (v > peak) { peak = v; peakX = xx; peakY = yy; }
		}
	int res = -1;
	if (peak >= thr && peak != FLT_MAX) {
		const int x0 = max(peakX - radius, 0), xe = min(peakX + radius, W - 1);
		const int y0 = max(peakY - radius, 0), ye = min(peakY + radius, H - 1);
		bool ok = true;
		for (int yy = y0; yy <= ye && ok; yy++)
			for (int xx = x0; xx <= xe; xx++) {
				const float v = __ldg(img + (long)yy * stride + xx);
				if (v >= peak && !(xx == peakX && yy == peakY)) { ok = false; break; }
			}
		if (ok) res = (peakY << 16) | peakX;              // both < 32768 (checked by the caller)
	}
	cand[b] = res;
}
// one CTA of 1024 threads: ordered compaction of cand[0..n) (>= 0) into out (x, y as int16), *total = how many
__global__ void __launch_bounds__(1024) nonmax_compact_kernel(const int* __restrict__ cand, int n, int cap, int16_t* __restrict__ out, int* __restrict__ total) {
	__shared__ int warp_cnt[32];
	__shared__ int base;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (threadIdx.x == 0) base = 0;
	__syncthreads();
	for (int i0 = 0; i0 < n; i0 += 1024) {
		const int i = i0 + threadIdx.x;
		const int c = i < n ? cand[i] : -1;
		const unsigned m = __ballot_sync(0xffffffffu, c >= 0);
		if (lane == 0) warp_cnt[warp] = __popc(m);
		__syncthreads();
		int off = base;
		for (int w = 0; w < warp; w++) off += warp_cnt[w];
		off += __popc(m & ((1u << lane) - 1));
		if (c >= 0 && off < cap) { out[2 * off] = (int16_t)(c & 0xffff); out[2 * off + 1] = (int16_t)(c >> 16); }
		__syncthreads();
		if (threadIdx.x == 0) { int t = 0; for (int w = 0; w < 32; w++) t += warp_cnt[w]; base += t; }
		__syncthreads();
	}
	if (threadIdx.x == 0) *total = base;
}
void launch_nonmax_single(const float* d_img, int stride, int W, int H, int radius, float thr, int border, int* d_cand, int cap, int16_t* d_out,
                          int* d_total, cudaStream_t st) {
	const int step = radius + 1;
	const int endX = W - border, endY = H - border;
	BS_CUDA(cudaMemsetAsync(d_total, 0, 4, st));
	if (endX <= border || endY <= border) return;
	const int bw = (endX - border + step - 1) / step, bh = (endY - border + step - 1) / step;
	BS_LAUNCH(nonmax_block_kernel, div_up(bw * bh, 128), 128, 0, st, d_img, stride, W, H, radius, thr, border, bw, bh, d_cand);
	BS_LAUNCH(nonmax_compact_kernel, 1, 1024, 0, st, (const int*)d_cand, bw * bh, cap, d_out, d_total);
}

}  // namespace bs
