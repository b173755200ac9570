// ORACLE — TEST INFRASTRUCTURE ONLY.  See orc_surf.h for status ("parity unpinned" end to end) and citations.
#include "orc_surf.h"
#include "orc_math.h"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <functional>
#include <numeric>
#include <stdexcept>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>

namespace orc {

// BoofConcurrency.loopFor (TY/concurrency/BoofConcurrency.java:82-88, :158-177) stand-in.  The reference hands the index range to
// a persistent ForkJoinPool, which splits it into contiguous sub-ranges that idle workers steal; here a persistent pool of
// worker threads (created once, parked on a condition variable between loops) takes contiguous chunks of the range off an
// atomic counter.  Spawning and joining fresh std::threads for each of the ~30 loops of a frame — what this file did before —
// cost more than the loops themselves beyond 16 threads.  Every body(i) writes its own outputs, so results do not depend on
// which thread runs which chunk.
namespace {
class Pool {
public:
	static Pool& get() { static Pool p; return p; }
	void run(int begin, int end, int nthreads, const std::function<void(int)>& body) {
		const int n = end - begin;
		const int nt = std::min(nthreads, n);
		std::unique_lock<std::mutex> job_lock(job_mutex_);              // one loop at a time (nested or concurrent callers queue up)
		grow(nt - 1);
		{
			std::lock_guard<std::mutex> lk(m_);
			body_ = &body; begin_ = begin; end_ = end;
			chunk_ = std::max(1, n / (nt * 4));                         // ~4 chunks per thread: contiguous blocks, balanced tails
			next_.store(begin, std::memory_order_relaxed);
			want_ = nt - 1; active_ = nt - 1;
			epoch_++;
		}
		cv_.notify_all();
		work();
		std::unique_lock<std::mutex> lk(m_);
		done_cv_.wait(lk, [&] { return active_ == 0; });
		body_ = nullptr;
	}
private:
	Pool() {}
	~Pool() {
		{ std::lock_guard<std::mutex> lk(m_); stop_ = true; epoch_++; }
		cv_.notify_all();
		for (auto& t : threads_) t.join();
	}
	void grow(int n) {
		while ((int)threads_.size() < n) {
			const int id = (int)threads_.size();
			threads_.emplace_back([this, id] { loop(id); });
		}
	}
	void work() {
		for (;;) {
			const int i0 = next_.fetch_add(chunk_, std::memory_order_relaxed);
			if (i0 >= end_) break;
			const int i1 = std::min(end_, i0 + chunk_);
			for (int i = i0; i < i1; i++) (*body_)(i);
		}
	}
	void loop(int id) {
		unsigned long seen = 0;
		for (;;) {
			{
				std::unique_lock<std::mutex> lk(m_);
				cv_.wait(lk, [&] { return epoch_ != seen; });
				seen = epoch_;
				if (stop_) return;
				if (id >= want_) continue;                                  // this loop uses fewer threads than the pool holds
			}
			work();
			std::lock_guard<std::mutex> lk(m_);
			if (--active_ == 0) done_cv_.notify_one();
		}
	}
	std::mutex job_mutex_, m_;
	std::condition_variable cv_, done_cv_;
	std::vector<std::thread> threads_;
	const std::function<void(int)>* body_ = nullptr;
	std::atomic<int> next_{0};
	int begin_ = 0, end_ = 0, chunk_ = 1, want_ = 0, active_ = 0;
	unsigned long epoch_ = 0;
	bool stop_ = false;
};
}  // namespace

static void parallel_for(int begin, int end, int nthreads, const std::function<void(int)>& body) {
	if (nthreads <= 1 || end - begin <= 1) { for (int i = begin; i < end; i++) body(i); return; }
	Pool::get().run(begin, end, nthreads, body);
}

// ===================================================================== integral image
// ImplIntegralImageOps.java:42-66 (F32), :94-118 (U8->S32), :120-144 (S32).  First row: running sum.
// Other rows: total += in; out = out_prev_row + total.  Java int arithmetic wraps; done in uint32 here.
template <typename TI, typename TO, typename ACC>
static void integral_generic(Img<TI> in, ImgMut<TO> out) {
	if (in.width <= 0 || in.height <= 0) return;
	ACC total = 0;
	for (int x = 0; x < in.width; x++) { total = (ACC)(total + (ACC)in.data[x]); out.data[x] = (TO)total; }
	for (int y = 1; y < in.height; y++) {
		const TI* src = in.data + (long)y * in.stride;
		TO* dst = out.data + (long)y * out.stride;
		const TO* prev = dst - out.stride;
		total = 0;
		for (int x = 0; x < in.width; x++) {
			total = (ACC)(total + (ACC)src[x]);
			dst[x] = (TO)((ACC)prev[x] + total);
		}
	}
}
void integral_f32(Img<float> in, ImgMut<float> out) { integral_generic<float, float, float>(in, out); }
void integral_u8(Img<uint8_t> in, ImgMut<int32_t> out) { integral_generic<uint8_t, int32_t, uint32_t>(in, out); }
void integral_s32(Img<int32_t> in, ImgMut<int32_t> out) { integral_generic<int32_t, int32_t, uint32_t>(in, out); }

// ===================================================================== block sums
// ImplIntegralImageOps.java:185-193 / :195-214: br-tr-bl+tl, evaluated left to right.
template <typename T> static T block_unsafe_t(Img<T> ii, int x0, int y0, int x1, int y1) {
	T br = ii.at(x1, y1), tr = ii.at(x1, y0), bl = ii.at(x0, y1), tl = ii.at(x0, y0);
	return br - tr - bl + tl;
}
template <typename T> static T block_zero_t(Img<T> ii, int x0, int y0, int x1, int y1) {
	x0 = std::min(x0, ii.width - 1); y0 = std::min(y0, ii.height - 1);
	x1 = std::min(x1, ii.width - 1); y1 = std::min(y1, ii.height - 1);
	T br = 0, tr = 0, bl = 0, tl = 0;
	if (x1 >= 0 && y1 >= 0) br = ii.at(x1, y1);
	if (y0 >= 0 && x1 >= 0) tr = ii.at(x1, y0);
	if (x0 >= 0 && y1 >= 0) bl = ii.at(x0, y1);
	if (x0 >= 0 && y0 >= 0) tl = ii.at(x0, y0);
	return br - tr - bl + tl;
}
// int32 versions wrap like Java
static int32_t wrap_sub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }
static int32_t wrap_add(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static int32_t wrap_mul(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }
template <> int32_t block_unsafe_t<int32_t>(Img<int32_t> ii, int x0, int y0, int x1, int y1) {
	int32_t br = ii.at(x1, y1), tr = ii.at(x1, y0), bl = ii.at(x0, y1), tl = ii.at(x0, y0);
	return wrap_add(wrap_sub(wrap_sub(br, tr), bl), tl);
}
template <> int32_t block_zero_t<int32_t>(Img<int32_t> ii, int x0, int y0, int x1, int y1) {
	x0 = std::min(x0, ii.width - 1); y0 = std::min(y0, ii.height - 1);
	x1 = std::min(x1, ii.width - 1); y1 = std::min(y1, ii.height - 1);
	int32_t br = 0, tr = 0, bl = 0, tl = 0;
	if (x1 >= 0 && y1 >= 0) br = ii.at(x1, y1);
	if (y0 >= 0 && x1 >= 0) tr = ii.at(x1, y0);
	if (x0 >= 0 && y1 >= 0) bl = ii.at(x0, y1);
	if (x0 >= 0 && y0 >= 0) tl = ii.at(x0, y0);
	return wrap_add(wrap_sub(wrap_sub(br, tr), bl), tl);
}
float block_unsafe(Img<float> ii, int x0, int y0, int x1, int y1) { return block_unsafe_t(ii, x0, y0, x1, y1); }
float block_zero(Img<float> ii, int x0, int y0, int x1, int y1) { return block_zero_t(ii, x0, y0, x1, y1); }
int32_t block_unsafe(Img<int32_t> ii, int x0, int y0, int x1, int y1) { return block_unsafe_t(ii, x0, y0, x1, y1); }
int32_t block_zero(Img<int32_t> ii, int x0, int y0, int x1, int y1) { return block_zero_t(ii, x0, y0, x1, y1); }

// ===================================================================== derivative kernels
// DerivativeIntegralImage.java:102-121
IntegralKernel kernel_deriv_xx(int size) {
	IntegralKernel k; k.n = 2;
	int blockW = size / 3, blockH = size - blockW - 1;
	int r1 = blockW / 2, r2 = blockW + r1, r3 = blockH / 2;
	k.blocks[0] = Rect{-r2 - 1, -r3 - 1, r2, r3};
	k.blocks[1] = Rect{-r1 - 1, -r3 - 1, r1, r3};
	k.scales[0] = 1; k.scales[1] = -3;
	return k;
}
// :123-140
IntegralKernel kernel_deriv_yy(int size) {
	IntegralKernel k; k.n = 2;
	int blockW = size / 3, blockH = size - blockW - 1;
	int r1 = blockW / 2, r2 = blockW + r1, r3 = blockH / 2;
	k.blocks[0] = Rect{-r3 - 1, -r2 - 1, r3, r2};
	k.blocks[1] = Rect{-r3 - 1, -r1 - 1, r3, r1};
	k.scales[0] = 1; k.scales[1] = -3;
	return k;
}
// :142-158
IntegralKernel kernel_deriv_xy(int size) {
	IntegralKernel k; k.n = 4;
	int block = size / 3;
	k.blocks[0] = Rect{-block - 1, -block - 1, -1, -1};
	k.blocks[1] = Rect{0, -block - 1, block, -1};
	k.blocks[2] = Rect{0, 0, block, block};
	k.blocks[3] = Rect{-block - 1, 0, -1, block};
	k.scales[0] = 1; k.scales[1] = -1; k.scales[2] = 1; k.scales[3] = -1;
	return k;
}
// ImplIntegralImageOps.java:172-183: ret += block_zero(...)*scale, float accumulate (int scale promoted to float)
float convolve_sparse(Img<float> ii, const IntegralKernel& k, int x, int y) {
	float ret = 0;
	for (int i = 0; i < k.n; i++) {
		const Rect& r = k.blocks[i];
		ret += block_zero_t(ii, x + r.x0, y + r.y0, x + r.x1, y + r.y1) * (float)k.scales[i];
	}
	return ret;
}
// :216-227, all-int arithmetic
int32_t convolve_sparse(Img<int32_t> ii, const IntegralKernel& k, int x, int y) {
	int32_t ret = 0;
	for (int i = 0; i < k.n; i++) {
		const Rect& r = k.blocks[i];
		ret = wrap_add(ret, wrap_mul(block_zero_t(ii, x + r.x0, y + r.y0, x + r.x1, y + r.y1), k.scales[i]));
	}
	return ret;
}

// ===================================================================== Fast-Hessian intensity
// computeHessian, ImplIntegralImageFeatureIntensity.java:115-127 (F32) / :290-302 (S32: int result converted to float)
template <typename T>
static float compute_hessian(Img<T> ii, const IntegralKernel& kxx, const IntegralKernel& kyy, const IntegralKernel& kxy,
                             float norm, int xx, int yy) {
	float Dxx = (float)convolve_sparse(ii, kxx, xx, yy);
	float Dyy = (float)convolve_sparse(ii, kyy, xx, yy);
	float Dxy = (float)convolve_sparse(ii, kxy, xx, yy);
	Dxx *= norm; Dxy *= norm; Dyy *= norm;
	float a = Dxx * Dyy;
	float b = 0.81f * Dxy;
	b = b * Dxy;
	return a - b;
}

// hessianNaive :45-67
template <typename T> void hessian_naive(Img<T> ii, int skip, int size, ImgMut<float> inten) {
	IntegralKernel kxx = kernel_deriv_xx(size), kyy = kernel_deriv_yy(size), kxy = kernel_deriv_xy(size);
	float norm = 1.0f / (float)(size * size);
	for (int y = 0; y < inten.height; y++)
		for (int x = 0; x < inten.width; x++)
			inten.data[(long)y * inten.stride + x] = compute_hessian(ii, kxx, kyy, kxy, norm, x * skip, y * skip);
}

// the eight-corner sums of hessianInner :183-195.  F32: float all the way.  S32 (:356-368): int box sums,
// "float Dxx = <int>; Dxx -= 3*(<int>)" = two separate int->float conversions.
static inline float sum4(float a, float b, float c, float d) { return a - b - c + d; }
static inline int32_t sum4(int32_t a, int32_t b, int32_t c, int32_t d) { return wrap_add(wrap_sub(wrap_sub(a, b), c), d); }
static inline float times3(float v) { return 3.0f * v; }
static inline float times3(int32_t v) { return (float)wrap_mul(3, v); }

// hessianBorder :72-113 + hessianInner :132-213
template <typename T> void hessian(Img<T> ii, int skip, int size, ImgMut<float> inten, int nthreads) {
	const int w = inten.width, h = inten.height;
	IntegralKernel kxx = kernel_deriv_xx(size), kyy = kernel_deriv_yy(size), kxy = kernel_deriv_xy(size);
	const float norm = 1.0f / (float)(size * size);
	const int blockSmall = size / 3, blockLarge = size - blockSmall - 1;
	const int radiusFeature = size / 2, radiusSkinny = blockLarge / 2;
	const int blockW2 = 2 * blockSmall, blockW3 = 3 * blockSmall;
	const long rowOff1 = (long)blockSmall * ii.stride, rowOff2 = 2 * rowOff1, rowOff3 = 3 * rowOff1;
	const int borderOrig = radiusFeature + 1 + (skip - (radiusFeature + 1) % skip);
	const int border = borderOrig / skip;
	const int lostPixel = borderOrig - radiusFeature - 1;
	const int endY = h - border, endX = w - border;
	const T* d = ii.data;

	auto row = [&](int y) {
		float* out = inten.data + (long)y * inten.stride;
		int yy = y * skip;
		if (y < border || y >= endY) {   // second loop nest of hessianBorder (+ the corners from the first)
			for (int x = 0; x < w; x++) out[x] = compute_hessian(ii, kxx, kyy, kxy, norm, x * skip, yy);
			return;
		}
		for (int x = 0; x < border && x < w; x++) out[x] = compute_hessian(ii, kxx, kyy, kxy, norm, x * skip, yy);
		for (int x = std::max(w - border, 0); x < w; x++) out[x] = compute_hessian(ii, kxx, kyy, kxy, norm, x * skip, yy);
		long indexTop = (long)(yy - radiusSkinny - 1) * ii.stride + lostPixel;
		long indexBottom = indexTop + (long)blockLarge * ii.stride;
		long indexL = (long)(yy - radiusFeature - 1) * ii.stride + (radiusFeature - radiusSkinny) + lostPixel;
		long indexR = indexL + blockLarge;
		long indexY1 = (long)(yy - blockSmall - 1) * ii.stride + (radiusFeature - blockSmall) + lostPixel;
		long indexY2 = indexY1 + (long)blockSmall * ii.stride;
		long indexY3 = indexY2 + ii.stride;
		long indexY4 = indexY3 + (long)blockSmall * ii.stride;
		const int x3 = blockSmall + 1, x4 = x3 + blockSmall;
		for (int x = border; x < endX; x++) {
			float Dxx = (float)sum4(d[indexBottom + blockW3], d[indexTop + blockW3], d[indexBottom], d[indexTop]);
			Dxx -= times3(sum4(d[indexBottom + blockW2], d[indexTop + blockW2], d[indexBottom + blockSmall], d[indexTop + blockSmall]));
			float Dyy = (float)sum4(d[indexR + rowOff3], d[indexL + rowOff3], d[indexR], d[indexL]);
			Dyy -= times3(sum4(d[indexR + rowOff2], d[indexL + rowOff2], d[indexR + rowOff1], d[indexL + rowOff1]));
			float Dxy = (float)sum4(d[indexY2 + blockSmall], d[indexY1 + blockSmall], d[indexY2], d[indexY1]);
			Dxy -= (float)sum4(d[indexY2 + x4], d[indexY1 + x4], d[indexY2 + x3], d[indexY1 + x3]);
			Dxy += (float)sum4(d[indexY4 + x4], d[indexY3 + x4], d[indexY4 + x3], d[indexY3 + x3]);
			Dxy -= (float)sum4(d[indexY4 + blockSmall], d[indexY3 + blockSmall], d[indexY4], d[indexY3]);
			Dxx *= norm; Dxy *= norm; Dyy *= norm;
			float a = Dxx * Dyy;
			float b = 0.81f * Dxy;
			b = b * Dxy;
			out[x] = a - b;
			indexTop += skip; indexBottom += skip; indexL += skip; indexR += skip;
			indexY1 += skip; indexY2 += skip; indexY3 += skip; indexY4 += skip;
		}
	};
	parallel_for(0, h, nthreads, row);
}
template void hessian_naive<float>(Img<float>, int, int, ImgMut<float>);
template void hessian_naive<int32_t>(Img<int32_t>, int, int, ImgMut<float>);
template void hessian<float>(Img<float>, int, int, ImgMut<float>, int);
template void hessian<int32_t>(Img<int32_t>, int, int, ImgMut<float>, int);

// ===================================================================== non-max suppression
// NonMaxBlock.process :69-94 with NonMaxBlockSearchStrict.Max.searchBlock :56-79 and checkLocalMax :196-221.
// Single-thread emission order is block-row-major; nthreads>1 threads over block rows (NonMaxBlock_MT.java:77)
// but concatenates in block-row order so the result equals the single-thread list.
void nonmax_block_strict_max(Img<float> img, int radius, float thresholdMax, int border, std::vector<PointI16>& out,
                             int nthreads) {
	out.clear();
	const int endX = img.width - border, endY = img.height - border;
	const int step = radius + 1;
	if (endY <= border || endX <= border) return;
	const int nrows = (endY - border + step - 1) / step;
	std::vector<std::vector<PointI16>> per_row(nrows);
	auto block_row = [&](int br) {
		int y = border + br * step;
		int y1 = std::min(y + step, endY);
		for (int x = border; x < endX; x += step) {
			int x1 = std::min(x + step, endX);
			int peakX = 0, peakY = 0; float peakVal = -FLT_MAX;
			for (int yy = y; yy < y1; yy++)
				for (int xx = x; xx < x1; xx++) {
					float v = img.at(xx, yy);
					if (v > peakVal) { peakVal = v; peakX = xx; peakY = yy; }
				}
			if (peakVal >= thresholdMax && peakVal != FLT_MAX) {
				int x0 = std::max(peakX - radius, 0), xe = std::min(peakX + radius, img.width - 1);
				int y0 = std::max(peakY - radius, 0), ye = std::min(peakY + radius, img.height - 1);
				bool ok = true;
				for (int yy = y0; yy <= ye && ok; yy++)
					for (int xx = x0; xx <= xe; xx++) {
						float v = img.at(xx, yy);
						if (v >= peakVal && !(xx == peakX && yy == peakY)) { ok = false; break; }
					}
				if (ok) per_row[br].push_back(PointI16{(int16_t)peakX, (int16_t)peakY});
			}
		}
	};
	parallel_for(0, nrows, nthreads, block_row);
	for (auto& v : per_row) out.insert(out.end(), v.begin(), v.end());
}

// NonMaxExtractorNaive.strictRule (F/alg/feature/detect/extract/NonMaxExtractorNaive.java:82-125)
void nonmax_naive_strict_max(Img<float> img, int radius, float thresh, int border, std::vector<PointI16>& out) {
	out.clear();
	for (int y = border; y < img.height - border; y++)
		for (int x = border; x < img.width - border; x++) {
			float val = img.at(x, y);
			if (val < thresh) continue;
			int x0 = std::max(x - radius, 0), x1 = std::min(x + radius, img.width - 1);
			int y0 = std::max(y - radius, 0), y1 = std::min(y + radius, img.height - 1);
			bool max = true;
			for (int i = y0; i <= y1 && max; i++)
				for (int j = x0; j <= x1; j++) {
					if (i == y && j == x) continue;
					if (val <= img.at(j, i)) { max = false; break; }
				}
			if (max && val != FLT_MAX) out.push_back(PointI16{(int16_t)x, (int16_t)y});
		}
}


// ===================================================================== Fast-Hessian detector
// polyPeak(float,float,float) FastHessianFeatureDetector.java:336-350
float poly_peak(float lower, float middle, float upper) {
	float a = 0.5f * lower - middle + 0.5f * upper;
	float b = 0.5f * upper - 0.5f * lower;
	if (a == 0.0f) return 0.0f;
	return -b / (2.0f * a);
}

// checkMax :304-313 through ImageBorderValue(0): out-of-image reads give 0
static bool check_max(Img<float> inten, float best, int c_x, int c_y) {
	for (int y = c_y - 1; y <= c_y + 1; y++)
		for (int x = c_x - 1; x <= c_x + 1; x++) {
			float v = (x >= 0 && y >= 0 && x < inten.width && y < inten.height) ? inten.at(x, y) : 0.0f;
			if (v >= best) return false;
		}
	return true;
}

// SelectNBestFeatures.process :51-93.  ddogleg QuickSelect's order among the kept points is not pinned by the
// reference (third-party, "parity unpinned"); this restatement keeps the N largest responses, ties broken towards the
// lower candidate index, and emits them in candidate (block-raster) order.
static void select_n_best(Img<float> inten, const std::vector<PointI16>& in, int target, std::vector<PointI16>& out) {
	out.clear();
	if ((int)in.size() <= target) { out = in; return; }
	std::vector<int> idx(in.size());
	std::iota(idx.begin(), idx.end(), 0);
	std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) {
		return inten.at(in[a].x, in[a].y) > inten.at(in[b].x, in[b].y);
	});
	idx.resize(target);
	std::sort(idx.begin(), idx.end());
	for (int i : idx) out.push_back(in[i]);
}

// detect :156-188, detectOctave :198-221, findLocalScaleSpaceMax :230-298
template <typename T>
void fast_hessian_detect(Img<T> ii, const FastHessianConfig& cfg, std::vector<ScalePoint>& found, int nthreads) {
	found.clear();
	const int ns = cfg.scales_per_octave;
	std::vector<int> sizes(ns);
	std::vector<float> buf[3];
	int skip = cfg.initial_sample_size, sizeStep = cfg.scale_step_size, octaveSize = cfg.initial_size;
	std::vector<PointI16> cand, best;
	for (int octave = 0; octave < cfg.octaves; octave++) {
		for (int i = 0; i < ns; i++) sizes[i] = octaveSize + i * sizeStep;
		int maxSize = sizes[ns - 1];
		if (maxSize > ii.width || maxSize > ii.height) break;
		const int w = ii.width / skip, h = ii.height / skip;
		for (auto& b : buf) b.assign((size_t)w * h, 0.0f);
		int spaceIndex = 0;   // the ring position is irrelevant to the result (:231-233 always pick oldest/middle/newest)
		for (int i = 0; i < ns; i++) {
			ImgMut<float> cur{buf[spaceIndex].data(), w, w, h};
			hessian(ii, skip, sizes[i], cur, nthreads);
			spaceIndex = (spaceIndex + 1) % 3;
			if (i < 2) continue;
			const int level = i - 1;
			Img<float> inten0{buf[spaceIndex].data(), w, w, h};
			Img<float> inten1{buf[(spaceIndex + 1) % 3].data(), w, w, h};
			Img<float> inten2{buf[(spaceIndex + 2) % 3].data(), w, w, h};
			const int ignoreBorder = sizes[level] / (2 * skip);
			nonmax_block_strict_max(inten1, cfg.extract_radius, cfg.detect_threshold, ignoreBorder, cand, nthreads);
			const int ignoreRadius = ignoreBorder + cfg.extract_radius;
			const int ignoreWidth = w - ignoreRadius, ignoreHeight = h - ignoreRadius;
			const std::vector<PointI16>* features = &cand;
			long numberRemaining = 0x7fffffffL;
			if (cfg.max_features_per_scale > 0) {
				select_n_best(inten1, cand, cfg.max_features_per_scale, best);
				features = &best; numberRemaining = cfg.max_features_per_scale;
			}
			const int levelSize = sizes[level];
			const int levelStep = levelSize - sizes[level - 1];
			for (size_t k = 0; k < features->size() && numberRemaining > 0; k++) {
				PointI16 f = (*features)[k];
				if (f.x < ignoreRadius || f.x >= ignoreWidth || f.y < ignoreRadius || f.y >= ignoreHeight) continue;
				float val = inten1.at(f.x, f.y);
				if (check_max(inten0, val, f.x, f.y) && check_max(inten2, val, f.x, f.y)) {
					float peakX = poly_peak(inten1.at(f.x - 1, f.y), inten1.at(f.x, f.y), inten1.at(f.x + 1, f.y));
					float peakY = poly_peak(inten1.at(f.x, f.y - 1), inten1.at(f.x, f.y), inten1.at(f.x, f.y + 1));
					float peakS = poly_peak(inten0.at(f.x, f.y), inten1.at(f.x, f.y), inten2.at(f.x, f.y));
					float interpX = ((float)f.x + peakX) * (float)skip;
					float interpY = ((float)f.y + peakY) * (float)skip;
					float interpS = (float)levelSize + peakS * (float)levelStep;
					double scale = 1.2 * (double)interpS / 9.0;
					found.push_back(ScalePoint{(double)interpX, (double)interpY, scale, octave, level, f.x, f.y, val});
					numberRemaining--;
				}
			}
		}
		skip += skip; octaveSize += sizeStep; sizeStep += sizeStep;
	}
}
template void fast_hessian_detect<float>(Img<float>, const FastHessianConfig&, std::vector<ScalePoint>&, int);
template void fast_hessian_detect<int32_t>(Img<int32_t>, const FastHessianConfig&, std::vector<ScalePoint>&, int);

// ===================================================================== Gaussian weight tables
// ddogleg UtilGaussian.computePDF(mean, sigma, sample) = exp(-(d*d)/(2 sigma^2)) / (sigma sqrt(2 pi))
static double gaussian_pdf(double mean, double sigma, double sample) {
	double delta = sample - mean;
	return std::exp(-delta * delta / (2.0 * sigma * sigma)) / (sigma * std::sqrt(2.0 * M_PI));
}
// FactoryKernelGaussian.sigmaForRadius :388-393 (order 0)
static double sigma_for_radius(double radius) { return (radius * 2.0 + 1.0) / 5.0; }

// gaussian(2,true,64,sigma,radius) :120-135 -> gaussian2D_F64(sigma,radius,odd=true,normalize=true) :297-306:
// 1-D un-normalised pdf (gaussian1D_F64 :240-262), outer product (KernelMath.convolve2D :365-383), sum-normalise (:446-452)
std::vector<double> gaussian2d_f64(double sigma, int radius) {
	if (sigma <= 0) sigma = sigma_for_radius(radius);
	const int w = radius * 2 + 1;
	std::vector<double> k1(w);
	int index = 0;
	for (int i = radius; i >= -radius; i--) k1[index++] = gaussian_pdf(0, sigma, i);
	std::vector<double> k2((size_t)w * w);
	index = 0;
	for (int i = 0; i < w; i++) for (int j = 0; j < w; j++) k2[index++] = k1[i] * k1[j];
	double total = 0;
	for (double v : k2) total += v;
	for (double& v : k2) v /= total;
	return k2;
}
// gaussianWidth :418-448
std::vector<double> gaussian_width(double sigma, int width) {
	if (sigma <= 0) sigma = sigma_for_radius(width / 2);
	else if (width <= 0) throw std::invalid_argument("Must specify the width");
	if (width % 2 == 0) {
		int r = width / 2 - 1;
		std::vector<double> ret((size_t)width * width);
		double sum = 0;
		for (int y = 0; y < width; y++) {
			double dy = y <= r ? std::abs(y - r) + 0.5 : std::abs(y - r - 1) + 0.5;
			for (int x = 0; x < width; x++) {
				double dx = x <= r ? std::abs(x - r) + 0.5 : std::abs(x - r - 1) + 0.5;
				double d = std::sqrt(dx * dx + dy * dy);
				double val = gaussian_pdf(0, sigma, d);
				ret[(size_t)y * width + x] = val;
				sum += val;
			}
		}
		for (double& v : ret) v /= sum;
		return ret;
	}
	return gaussian2d_f64(sigma, width / 2);
}

// ===================================================================== sparse gradient
// SparseIntegralGradient_NoBorder.setWidth :41-47
template <typename T> void SparseGradient<T>::set_width(double width) {
	r = ((int)(width + 0.5)) / 2;
	if (r <= 0) r = 1;
	w = r * 2 + 1;
	if (haar) w = 2 * r;                    // SparseIntegralHaar_NoBorder_F32.setWidth :37-43: w = 2r, x0 = y0 = -r, x1 = y1 = r
}
// SparseScaleGradient.isInBounds :47-50 with x0=y0=-r-1, x1=y1=r (SparseIntegralGradient_NoBorder_F32.java:37-43); Haar: x0=y0=-r
template <typename T> bool SparseGradient<T>::in_bounds(int x, int y) const {
	if (haar) return x - r >= 0 && y - r >= 0 && x + r < ii.width && y + r < ii.height;
	return x - r - 1 >= 0 && y - r - 1 >= 0 && x + r < ii.width && y + r < ii.height;
}
// SparseIntegralHaar_NoBorder_F32.compute :46-73 / _I32 :46-73
static void haar_compute(const Img<float>& ii, int r, int w, int x, int y, double& dx, double& dy) {
	const float* d = ii.data;
	long i1 = (long)(y - r) * ii.stride + (x - r), i2 = (long)y * ii.stride + (x - r), i3 = (long)(y + r) * ii.stride + (x - r);
	float p0 = d[i1], p1 = d[i1 + r], p2 = d[i1 + w], p3 = d[i2], p5 = d[i2 + w], p6 = d[i3], p7 = d[i3 + r], p8 = d[i3 + w];
	float left = p7 - p1 - p6 + p0, right = p8 - p2 - p7 + p1, top = p5 - p2 - p3 + p0, bottom = p8 - p5 - p6 + p3;
	dx = (double)(right - left); dy = (double)(bottom - top);
}
static void haar_compute(const Img<int32_t>& ii, int r, int w, int x, int y, double& dx, double& dy) {
	const int32_t* d = ii.data;
	long i1 = (long)(y - r) * ii.stride + (x - r), i2 = (long)y * ii.stride + (x - r), i3 = (long)(y + r) * ii.stride + (x - r);
	int32_t p0 = d[i1], p1 = d[i1 + r], p2 = d[i1 + w], p3 = d[i2], p5 = d[i2 + w], p6 = d[i3], p7 = d[i3 + r], p8 = d[i3 + w];
	int32_t left = sum4(p7, p1, p6, p0), right = sum4(p8, p2, p7, p1), top = sum4(p5, p2, p3, p0), bottom = sum4(p8, p5, p6, p3);
	dx = (double)wrap_sub(right, left); dy = (double)wrap_sub(bottom, top);
}
// SparseIntegralGradient_NoBorder_F32.compute :46-76 (float) / _I32 (int, wrapping)
template <> void SparseGradient<float>::compute(int x, int y, double& dx, double& dy) const {
	if (haar) { haar_compute(ii, r, w, x, y, dx, dy); return; }
	const float* d = ii.data;
	long i1 = (long)(y - r - 1) * ii.stride + (x - r - 1);
	long i2 = i1 + (long)r * ii.stride, i3 = i2 + ii.stride, i4 = i3 + (long)r * ii.stride;
	float p0 = d[i1], p1 = d[i1 + r], p2 = d[i1 + r + 1], p3 = d[i1 + w];
	float p11 = d[i2], p4 = d[i2 + w];
	float p10 = d[i3], p5 = d[i3 + w];
	float p9 = d[i4], p8 = d[i4 + r], p7 = d[i4 + r + 1], p6 = d[i4 + w];
	float left = p8 - p9 - p1 + p0;
	float right = p6 - p7 - p3 + p2;
	float top = p4 - p11 - p3 + p0;
	float bottom = p6 - p9 - p5 + p10;
	dx = (double)(right - left);
	dy = (double)(bottom - top);
}
template <> void SparseGradient<int32_t>::compute(int x, int y, double& dx, double& dy) const {
	if (haar) { haar_compute(ii, r, w, x, y, dx, dy); return; }
	const int32_t* d = ii.data;
	long i1 = (long)(y - r - 1) * ii.stride + (x - r - 1);
	long i2 = i1 + (long)r * ii.stride, i3 = i2 + ii.stride, i4 = i3 + (long)r * ii.stride;
	int32_t p0 = d[i1], p1 = d[i1 + r], p2 = d[i1 + r + 1], p3 = d[i1 + w];
	int32_t p11 = d[i2], p4 = d[i2 + w];
	int32_t p10 = d[i3], p5 = d[i3 + w];
	int32_t p9 = d[i4], p8 = d[i4 + r], p7 = d[i4 + r + 1], p6 = d[i4 + w];
	int32_t left = sum4(p8, p9, p1, p0), right = sum4(p6, p7, p3, p2);
	int32_t top = sum4(p4, p11, p3, p0), bottom = sum4(p6, p9, p5, p10);
	dx = (double)wrap_sub(right, left);
	dy = (double)wrap_sub(bottom, top);
}
template <typename T> void SparseGradient<T>::compute_safe(int x, int y, double& dx, double& dy) const {
	if (in_bounds(x, y)) compute(x, y, dx, dy); else { dx = 0; dy = 0; }
}
template struct SparseGradient<float>;
template struct SparseGradient<int32_t>;

// ===================================================================== containment tests
// SurfDescribeOps.isInside(ii,X,Y,radiusRegions,kernelSize,scale,c,s) :119-159
bool is_inside_rotated(int width, int height, double X, double Y, int radiusRegions, int kernelSize, double scale,
                       double c, double s) {
	int c_x = (int)std::floor(X + 0.5);   // Math.round
	int c_y = (int)std::floor(Y + 0.5);
	kernelSize = (int)std::ceil(kernelSize * scale);
	int kernelRadius = kernelSize / 2 + (kernelSize % 2);
	int radius = (int)std::ceil(radiusRegions * scale);
	int kernelPaddingMinus = radius + kernelRadius + 1;
	int kernelPaddingPlus = radius + kernelRadius;
	if (c != 0 || s != 0) {
		double xx = std::fabs(c * kernelPaddingMinus - s * kernelPaddingMinus);
		double yy = std::fabs(s * kernelPaddingMinus + c * kernelPaddingMinus);
		double delta = xx > yy ? xx - kernelPaddingMinus : yy - kernelPaddingMinus;
		kernelPaddingMinus += (int)std::ceil(delta);
		kernelPaddingPlus += (int)std::ceil(delta);
	}
	int x0 = c_x - kernelPaddingMinus; if (x0 < 0) return false;
	int x1 = c_x + kernelPaddingPlus;  if (x1 >= width) return false;
	int y0 = c_y - kernelPaddingMinus; if (y0 < 0) return false;
	int y1 = c_y + kernelPaddingPlus;  if (y1 >= height) return false;
	return true;
}
// SurfDescribeOps.isInside(width,height,tl_x,tl_y,regionSize,sampleSize) :183-204
bool is_inside_rect(int width, int height, double tl_x, double tl_y, double regionSize, double sampleSize) {
	int w = (int)(sampleSize + 0.5);
	int r = w / 2 + w % 2;
	int x0 = (int)(tl_x + 0.5) - r - 1;
	int y0 = (int)(tl_y + 0.5) - r - 1;
	if (x0 < 0 || y0 < 0) return false;
	int x1 = (int)(tl_x + regionSize + 0.5) + r;
	int y1 = (int)(tl_y + regionSize + 0.5) + r;
	if (x1 >= width || y1 >= height) return false;
	return true;
}


// ===================================================================== orientation
// OrientationIntegralBase ctor :75-92: weights = gaussian(2,true,64,weightSigma,sampleRadius) unless sigma == 0;
// assignDefaultRadius => setObjectRadius(1/objectRadiusToScale)
template <typename T> Orientation<T>::Orientation(const OrientationConfig& c) : cfg(c) {
	if (cfg.weight_sigma != 0) weights = gaussian2d_f64(cfg.weight_sigma, cfg.radius);
	set_object_radius(1.0 / cfg.object_radius_to_scale);
}
// :94-98
template <typename T> void Orientation<T>::set_object_radius(double radius) {
	scale = radius * cfg.object_radius_to_scale;
	g.set_width(scale * cfg.sample_width);
}

// georegression UtilAngle.dist (0.21-SNAPSHOT, source not in tree): shortest angular distance in [0,pi]
static double angle_dist(double a, double b) {
	double diff = std::fabs(a - b);
	if (diff > M_PI) return 2.0 * M_PI - diff;
	return diff;
}

template <typename T> double Orientation<T>::compute(double c_x, double c_y) {
	const int sampleRadius = cfg.radius, sampleWidth = cfg.radius * 2 + 1;
	double period = scale * cfg.sample_period;
	double tl_x = c_x - sampleRadius * period;
	double tl_y = c_y - sampleRadius * period;
	if (cfg.variant == 0) {
		// ImplOrientationAverageGradientIntegral.compute :54-72, computeUnweighted :77-101, computeWeighted :106-127
		bool safe = !is_inside_rect(g.ii.width, g.ii.height, tl_x, tl_y, sampleWidth * period, cfg.sample_width * scale);
		tl_x += 0.5; tl_y += 0.5;
		double Dx = 0, Dy = 0;
		int i = 0;
		for (int y = 0; y < sampleWidth; y++) {
			int pixelsY = (int)(tl_y + y * period);
			for (int x = 0; x < sampleWidth; x++, i++) {
				int pixelsX = (int)(tl_x + x * period);
				double dx, dy;
				if (safe) g.compute_safe(pixelsX, pixelsY, dx, dy); else g.compute(pixelsX, pixelsY, dx, dy);
				if (weights.empty()) { Dx += dx; Dy += dy; }
				else { double w = weights[i]; Dx += w * dx; Dy += w * dy; }
			}
		}
		return fd_atan2(Dy, Dx);
	}
	// ImplOrientationSlidingWindowIntegral.compute :81-107, computeGradient :109-137
	const int total = sampleWidth * sampleWidth;
	std::vector<double> derivX(total), derivY(total), angles(total);
	std::vector<int> order(total);
	{
		double tx = tl_x + 0.5, ty = tl_y + 0.5;
		int k = 0;
		for (int y = 0; y < sampleWidth; y++)
			for (int x = 0; x < sampleWidth; x++, k++) {
				int xx = (int)(tx + x * period);
				int yy = (int)(ty + y * period);
				g.compute_safe(xx, yy, derivX[k], derivY[k]);
			}
	}
	if (!weights.empty())
		for (int i = 0; i < total; i++) { double w = weights[i]; derivX[i] *= w; derivY[i] *= w; }
	for (int i = 0; i < total; i++) angles[i] = fd_atan2(derivY[i], derivX[i]);
	// ddogleg QuickSort_F64.sort(angles,0,n,order): ascending argsort, angles untouched.  Tie order is not pinned
	// by the reference ("parity unpinned"); a stable sort (ties by lower index) is used.
	std::iota(order.begin(), order.end(), 0);
	std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return angles[a] < angles[b]; });
	// estimateAngle :139-188
	int start = 0, end = 1;
	int startIndex = order[start], endIndex = order[end];
	double sumX = derivX[startIndex], sumY = derivY[startIndex];
	double best = sumX * sumX + sumY * sumY, bestX = sumX, bestY = sumY;
	double endAngle = angles[endIndex];
	while (start != total) {
		startIndex = order[start];
		double startAngle = angles[startIndex];
		while (angle_dist(startAngle, endAngle) <= cfg.window_size) {
			sumX += derivX[endIndex];
			sumY += derivY[endIndex];
			double mag = sumX * sumX + sumY * sumY;
			if (mag > best) { best = mag; bestX = sumX; bestY = sumY; }
			end++;
			if (end >= total) end = 0;
			endIndex = order[end];
			endAngle = angles[endIndex];
			if (endIndex == startIndex) break;
		}
		sumX -= derivX[startIndex];
		sumY -= derivY[startIndex];
		start++;
	}
	return fd_atan2(bestY, bestX);
}
template struct Orientation<float>;
template struct Orientation<int32_t>;

// ===================================================================== descriptor
// UtilFeature.normalizeL2 :101-114
void normalize_l2(double* v, int n) {
	double norm = 0;
	for (int i = 0; i < n; i++) { double t = v[i]; norm += t * t; }
	if (norm == 0) return;
	norm = std::sqrt(norm);
	for (int i = 0; i < n; i++) v[i] /= norm;
}

// DescribePointSurf ctor :112-139, DescribePointSurfMod ctor :71-99
template <typename T> Describe<T>::Describe(const DescribeConfig& c) : cfg(c) {
	const int region = cfg.width_large_grid * cfg.width_sub_region;
	dof = cfg.width_large_grid * cfg.width_large_grid * 4;
	radius_descriptor = region / 2;
	if (cfg.variant == 0) {
		int radius = region / 2;
		weight = gaussian_width(cfg.weight_sigma, radius * 2);
		double div = weight[(size_t)radius * (radius * 2) + radius];
		for (double& v : weight) v /= div;
	} else {
		weight_grid = gaussian_width(cfg.sigma_large_grid, cfg.width_large_grid);
		int wsub = cfg.width_sub_region + 2 * cfg.overlap;
		weight_sub = gaussian_width(cfg.sigma_sub_region, wsub);
		int rg = cfg.width_large_grid / 2;      // Kernel2D.getRadius() = width/2
		double div = weight_grid[(size_t)rg * cfg.width_large_grid + rg];
		for (double& v : weight_grid) v /= div;
		int rs = wsub / 2;
		div = weight_sub[(size_t)rs * wsub + rs];
		for (double& v : weight_sub) v /= div;
		radius_descriptor = region / 2 + cfg.overlap;
	}
}

// DescribePointSurf.describe :169-213
template <typename T> void Describe<T>::describe_raw(double x, double y, double angle, double scale, double* desc) {
	double c = fd_cos(angle), s = fd_sin(angle);
	bool isInBounds = is_inside_rotated(ii.width, ii.height, x, y, radius_descriptor, cfg.width_sample, scale, c, s);
	g.ii = ii;
	g.haar = cfg.use_haar != 0;             // DescribePointSurf :135: SurfDescribeOps.createGradient(useHaar, inputType)
	g.set_width(cfg.width_sample * scale);
	if (cfg.variant == 0) features_fast(x, y, c, s, scale, !isInBounds, desc);
	else features_mod(x, y, c, s, scale, !isInBounds, desc);
}
template <typename T> void Describe<T>::describe(double x, double y, double angle, double scale, double* desc, bool* white) {
	describe_raw(x, y, angle, scale, desc);
	normalize_l2(desc, dof);
	if (white) *white = laplace_sign((int)(x + 0.5), (int)(y + 0.5), scale);
}

// DescribePointSurf.features :235-295
template <typename T> void Describe<T>::features_fast(double c_x, double c_y, double c, double s, double scale, bool safe, double* features) {
	const int widthSubRegion = cfg.width_sub_region;
	const int regionSize = cfg.width_large_grid * widthSubRegion;
	const int regionR = regionSize / 2, regionEnd = regionSize - regionR;
	int regionIndex = 0;
	c_x += 0.5; c_y += 0.5;
	for (int rY = -regionR; rY < regionEnd; rY += widthSubRegion)
		for (int rX = -regionR; rX < regionEnd; rX += widthSubRegion) {
			double sum_dx = 0, sum_dy = 0, sum_adx = 0, sum_ady = 0;
			for (int i = 0; i < widthSubRegion; i++) {
				double regionY = (rY + i) * scale;
				for (int j = 0; j < widthSubRegion; j++) {
					double w = weight[(size_t)(regionR + rY + i) * regionSize + (regionR + rX + j)];
					double regionX = (rX + j) * scale;
					int pixelX = (int)(c_x + c * regionX - s * regionY);
					int pixelY = (int)(c_y + s * regionX + c * regionY);
					double gx, gy;
					if (safe) g.compute_safe(pixelX, pixelY, gx, gy); else g.compute(pixelX, pixelY, gx, gy);
					double dx = w * gx, dy = w * gy;
					double pdx = c * dx + s * dy;
					double pdy = -s * dx + c * dy;
					sum_dx += pdx; sum_adx += std::fabs(pdx);
					sum_dy += pdy; sum_ady += std::fabs(pdy);
				}
			}
			features[regionIndex++] = sum_dx; features[regionIndex++] = sum_adx;
			features[regionIndex++] = sum_dy; features[regionIndex++] = sum_ady;
		}
}

// DescribePointSurfMod.features :121-192
template <typename T> void Describe<T>::features_mod(double c_x, double c_y, double c, double s, double scale, bool safe, double* features) {
	const int widthSubRegion = cfg.width_sub_region, overLap = cfg.overlap;
	const int regionSize = cfg.width_large_grid * widthSubRegion;
	const int totalSampleWidth = widthSubRegion + overLap * 2;
	const int regionR = regionSize / 2, regionEnd = regionSize - regionR;
	const int sampleGridWidth = regionSize + 2 * overLap;
	std::vector<double> samplesX((size_t)sampleGridWidth * sampleGridWidth), samplesY(samplesX.size());
	int regionIndex = 0;
	c_x += 0.5; c_y += 0.5;
	int index = 0;
	for (int rY = -regionR - overLap; rY < regionEnd + overLap; rY++) {
		double regionY = rY * scale;
		for (int rX = -regionR - overLap; rX < regionEnd + overLap; rX++, index++) {
			double regionX = rX * scale;
			int pixelX = (int)(c_x + c * regionX - s * regionY);
			int pixelY = (int)(c_y + s * regionX + c * regionY);
			if (safe) g.compute_safe(pixelX, pixelY, samplesX[index], samplesY[index]);
			else g.compute(pixelX, pixelY, samplesX[index], samplesY[index]);
		}
	}
	int indexGridWeight = 0;
	for (int rY = -regionR; rY < regionEnd; rY += widthSubRegion)
		for (int rX = -regionR; rX < regionEnd; rX += widthSubRegion) {
			double sum_dx = 0, sum_dy = 0, sum_adx = 0, sum_ady = 0;
			for (int i = 0; i < totalSampleWidth; i++) {
				index = (rY + regionR + i) * sampleGridWidth + rX + regionR;
				for (int j = 0; j < totalSampleWidth; j++, index++) {
					double w = weight_sub[(size_t)i * totalSampleWidth + j];
					double dx = w * samplesX[index], dy = w * samplesY[index];
					double pdx = c * dx + s * dy;
					double pdy = -s * dx + c * dy;
					sum_dx += pdx; sum_adx += std::fabs(pdx);
					sum_dy += pdy; sum_ady += std::fabs(pdy);
				}
			}
			double w = weight_grid[indexGridWeight++];
			features[regionIndex++] = w * sum_dx; features[regionIndex++] = w * sum_adx;
			features[regionIndex++] = w * sum_dy; features[regionIndex++] = w * sum_ady;
		}
}

// computeLaplaceSign :305-313: double lap = convolveSparse(XX); lap += convolveSparse(YY); lap > 0
template <typename T> bool Describe<T>::laplace_sign(int x, int y, double scale) {
	int s = (int)std::ceil(scale);
	IntegralKernel kxx = kernel_deriv_xx(9 * s), kyy = kernel_deriv_yy(9 * s);
	double lap = (double)convolve_sparse(ii, kxx, x, y);
	lap += (double)convolve_sparse(ii, kyy, x, y);
	return lap > 0;
}
template struct Describe<float>;
template struct Describe<int32_t>;

// ===================================================================== detect + describe wrapper
// WrapDetectDescribeSurf.detect :75-96, computeDescriptors :98-110; _MT: loopBlocks over features with per-thread
// copies of orientation/describe (WrapDetectDescribeSurf_MT.java:44-62) — per-feature results do not depend on the split.
template <typename TIn, typename TII> static void do_integral(Img<TIn> in, ImgMut<TII> out);
template <> void do_integral<float, float>(Img<float> in, ImgMut<float> out) { integral_f32(in, out); }
template <> void do_integral<uint8_t, int32_t>(Img<uint8_t> in, ImgMut<int32_t> out) { integral_u8(in, out); }
template <> void do_integral<int32_t, int32_t>(Img<int32_t> in, ImgMut<int32_t> out) { integral_s32(in, out); }

template <typename TIn, typename TII>
void surf_detect_describe(Img<TIn> input, const FastHessianConfig& fh, const OrientationConfig& oc, const DescribeConfig& dc,
                          SurfResult& out, int nthreads) {
	std::vector<TII> iibuf((size_t)input.width * input.height);
	ImgMut<TII> iim{iibuf.data(), input.width, input.width, input.height};
	do_integral<TIn, TII>(input, iim);      // never threaded in the reference (IntegralImageOps.java:47-53)
	Img<TII> ii = iim.view();
	fast_hessian_detect(ii, fh, out.points, nthreads);
	const int n = (int)out.points.size();
	Describe<TII> proto(dc);
	const int dof = proto.dof;
	out.angle.assign(n, 0.0); out.desc.assign((size_t)n * dof, 0.0); out.white.assign(n, 0);
	const int nt = std::max(1, std::min(nthreads, n));
	auto work = [&](int t) {
		Orientation<TII> ori(oc); Describe<TII> des(dc);
		ori.set_image(ii); des.set_image(ii);
		int b0 = (int)((long)n * t / nt), b1 = (int)((long)n * (t + 1) / nt);
		for (int i = b0; i < b1; i++) {
			const ScalePoint& p = out.points[i];
			double radius = p.scale * 2.0;           // BoofDefaults.SURF_SCALE_TO_RADIUS (IP/BoofDefaults.java:38)
			ori.set_object_radius(radius);
			double angle = ori.compute(p.x, p.y);
			bool white = false;
			des.describe(p.x, p.y, angle, p.scale, &out.desc[(size_t)i * dof], &white);
			out.angle[i] = angle; out.white[i] = white ? 1 : 0;
		}
	};
	parallel_for(0, nt, nt, work);
}
template void surf_detect_describe<float, float>(Img<float>, const FastHessianConfig&, const OrientationConfig&, const DescribeConfig&, SurfResult&, int);
template void surf_detect_describe<uint8_t, int32_t>(Img<uint8_t>, const FastHessianConfig&, const OrientationConfig&, const DescribeConfig&, SurfResult&, int);
template void surf_detect_describe<int32_t, int32_t>(Img<int32_t>, const FastHessianConfig&, const OrientationConfig&, const DescribeConfig&, SurfResult&, int);

// ===================================================================== colour SURF
// ImplConvertPlanarToGray.average: the 3-band and the general loop give the same values (sum left to right, then divide)
template <> void planar_average<float>(const float* bands, int nb, int w, int h, float* gray) {
	const size_t px = (size_t)w * h;
	for (size_t i = 0; i < px; i++) {
		if (nb == 1) { gray[i] = bands[i]; continue; }
		float sum = nb == 3 ? bands[i] : 0.0f;
		for (int b = nb == 3 ? 1 : 0; b < nb; b++) sum += bands[b * px + i];
		gray[i] = sum / nb;
	}
}
template <> void planar_average<uint8_t>(const uint8_t* bands, int nb, int w, int h, uint8_t* gray) {
	const size_t px = (size_t)w * h;
	for (size_t i = 0; i < px; i++) {
		int sum = 0;
		for (int b = 0; b < nb; b++) sum += bands[b * px + i] & 0xFF;
		gray[i] = (uint8_t)(sum / nb);
	}
}
template <> void planar_average<int32_t>(const int32_t* bands, int nb, int w, int h, int32_t* gray) {
	const size_t px = (size_t)w * h;
	for (size_t i = 0; i < px; i++) {
		uint32_t sum = 0;                              // Java int adds wrap
		for (int b = 0; b < nb; b++) sum += (uint32_t)bands[b * px + i];
		gray[i] = (int32_t)sum / nb;
	}
}

template <typename TIn, typename TII>
void surf_planar_detect_describe(const TIn* bands, int nb, int w, int h, const FastHessianConfig& fh, const OrientationConfig& oc,
                                 const DescribeConfig& dc, SurfResult& out, int nthreads) {
	const size_t px = (size_t)w * h;
	// SurfPlanar_to_DetectDescribePoint.detect :58-74
	std::vector<TIn> gray(px);
	planar_average<TIn>(bands, nb, w, h, gray.data());
	std::vector<TII> gbuf(px), bbuf(px * nb);
	do_integral<TIn, TII>(Img<TIn>{gray.data(), w, w, h}, ImgMut<TII>{gbuf.data(), w, w, h});
	for (int b = 0; b < nb; b++) do_integral<TIn, TII>(Img<TIn>{bands + b * px, w, w, h}, ImgMut<TII>{bbuf.data() + b * px, w, w, h});
	Img<TII> grayII{gbuf.data(), w, w, h};
	// DetectDescribeSurfPlanar.detect / describe :87-116
	fast_hessian_detect(grayII, fh, out.points, nthreads);
	const int n = (int)out.points.size();
	Describe<TII> proto(dc);
	const int dof = proto.dof, total = dof * nb;
	out.angle.assign(n, 0.0); out.desc.assign((size_t)n * total, 0.0); out.white.assign(n, 0);
	const int nt = std::max(1, std::min(nthreads, n));
	auto work = [&](int t) {
		Orientation<TII> ori(oc); Describe<TII> des(dc);
		ori.set_image(grayII);
		int b0 = (int)((long)n * t / nt), b1 = (int)((long)n * (t + 1) / nt);
		for (int i = b0; i < b1; i++) {
			const ScalePoint& p = out.points[i];
			ori.set_object_radius(p.scale);          // :108 — the scale itself, not scale*SURF_SCALE_TO_RADIUS as in WrapDetectDescribeSurf
			double angle = ori.compute(p.x, p.y);
			double* d = &out.desc[(size_t)i * total];
			// DescribePointSurfPlanar.describe :97-111
			for (int b = 0; b < nb; b++) {
				des.set_image(Img<TII>{bbuf.data() + b * px, w, w, h});
				des.describe_raw(p.x, p.y, angle, p.scale, d + b * dof);
			}
			normalize_l2(d, total);
			des.set_image(grayII);
			out.white[i] = des.laplace_sign((int)(p.x + 0.5), (int)(p.y + 0.5), p.scale) ? 1 : 0;
			out.angle[i] = angle;
		}
	};
	parallel_for(0, nt, nt, work);
}
template void surf_planar_detect_describe<float, float>(const float*, int, int, int, const FastHessianConfig&, const OrientationConfig&, const DescribeConfig&, SurfResult&, int);
template void surf_planar_detect_describe<uint8_t, int32_t>(const uint8_t*, int, int, int, const FastHessianConfig&, const OrientationConfig&, const DescribeConfig&, SurfResult&, int);

// ===================================================================== association
// DescriptorDistance.euclideanSq :55-64
double euclidean_sq(const double* a, const double* b, int n) {
	double total = 0;
	for (int i = 0; i < n; i++) { double d = a[i] - b[i]; total += d * d; }
	return total;
}
static inline double score_of(const double* a, const double* b, int dof, int score_type) {
	double e = euclidean_sq(a, b, dof);
	return score_type == 1 ? std::sqrt(e) : e;     // ScoreAssociateEuclidean_F64 = Math.sqrt(euclideanSq)
}

// AssociateGreedy.associate :65-118 (and _MT: loopFor over src rows, both passes)
void associate_greedy(const double* src, int n, const double* dst, int m, int dof, double maxFitError, bool backwards,
                      int score_type, int32_t* pairs, double* fit, int nthreads) {
	std::vector<double> work((size_t)n * m);
	parallel_for(0, n, nthreads, [&](int i) {
		const double* a = src + (size_t)i * dof;
		double bestScore = maxFitError; int bestIndex = -1;
		double* wrow = work.data() + (size_t)i * m;
		for (int j = 0; j < m; j++) {
			double f = score_of(a, dst + (size_t)j * dof, dof, score_type);
			wrow[j] = f;
			if (f <= bestScore) { bestIndex = j; bestScore = f; }
		}
		pairs[i] = bestIndex; fit[i] = bestScore;
	});
	if (!backwards) return;
	// each row only writes its own pairs[i]/fit[i] and reads work[] — safe to thread like the _MT twin
	std::vector<int32_t> fwd(pairs, pairs + n);
	parallel_for(0, n, nthreads, [&](int i) {
		int match = fwd[i];
		if (match == -1) return;
		double scoreToBeat = work[(size_t)i * m + match];
		for (int j = 0; j < n; j++) {
			if (work[(size_t)j * m + match] <= scoreToBeat && j != i) { pairs[i] = -1; fit[i] = DBL_MAX; break; }
		}
	});
}

// Same results, O(n+m) memory: recompute the column instead of reading the work buffer.
void associate_greedy_nobuffer(const double* src, int n, const double* dst, int m, int dof, double maxFitError,
                               bool backwards, int32_t* pairs, double* fit, int nthreads) {
	parallel_for(0, n, nthreads, [&](int i) {
		const double* a = src + (size_t)i * dof;
		double bestScore = maxFitError; int bestIndex = -1;
		for (int j = 0; j < m; j++) {
			double f = euclidean_sq(a, dst + (size_t)j * dof, dof);
			if (f <= bestScore) { bestIndex = j; bestScore = f; }
		}
		pairs[i] = bestIndex; fit[i] = bestScore;
	});
	if (!backwards) return;
	std::vector<int32_t> fwd(pairs, pairs + n);
	std::vector<double> fwdfit(fit, fit + n);
	parallel_for(0, n, nthreads, [&](int i) {
		int match = fwd[i];
		if (match == -1) return;
		const double* b = dst + (size_t)match * dof;
		double scoreToBeat = fwdfit[i];
		for (int j = 0; j < n; j++) {
			if (j != i && euclidean_sq(src + (size_t)j * dof, b, dof) <= scoreToBeat) { pairs[i] = -1; fit[i] = DBL_MAX; break; }
		}
	});
}

}  // namespace orc
