// bs_capi.cu — the extern "C" surface declared in include/boofsurf.h: handles, device workspaces, stream orchestration.
// Reference-side counterparts: WrapDetectDescribeSurf (F/abst/feature/detdesc/WrapDetectDescribeSurf.java:75-141) and
// WrapAssociateGreedy (F/abst/feature/associate/WrapAssociateGreedy.java:58-123).
#include "bs_common.cuh"
#include <cstdlib>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <memory>
#include <type_traits>
#include <vector>

using namespace bs;

namespace {

// RAII device buffer that only grows
struct DevBuf {
	void* p = nullptr;
	size_t bytes = 0;
	void reserve(size_t n) {
		if (n <= bytes) return;
		if (p) cudaFree(p);
		p = nullptr; bytes = 0;
		BS_CUDA(cudaMalloc(&p, n));
		bytes = n;
	}
	void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
	template <typename T> T* as() const { return static_cast<T*>(p); }
};

struct DeviceGuard {
	int prev = -1;
	explicit DeviceGuard(int dev) {
		BS_CUDA(cudaGetDevice(&prev));
		if (prev != dev) BS_CUDA(cudaSetDevice(dev));
		else prev = -1;
	}
	~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

template <typename F> int guarded(F&& f) {
	try { f(); return BS_OK; }
	catch (const Error& e) { set_last_error(e.what()); return e.code; }
	catch (const std::exception& e) { set_last_error(e.what()); return BS_ERR_CUDA; }
}

size_t elem_size(int pixel_type) { return pixel_type == BS_PIXEL_U8 ? 1 : 4; }

// Host->device copy for setup paths.  cudaMemcpy from pageable memory may return once the data is staged, before the DMA
// has landed; the library's streams are cudaStreamNonBlocking and would not wait for it, so drain the device here.
void h2d_sync(void* dst, const void* src, size_t bytes) {
	if (bytes == 0) return;
	BS_CUDA(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
	BS_CUDA(cudaDeviceSynchronize());
}

// OrientationIntegralBase ctor (:75-92) + DescribePointSurf / DescribePointSurfMod ctors
// (DescribePointSurf.java:112-139, DescribePointSurfMod.java:71-99): parameters and weight tables
struct DescribeSetup {
	DescribeParams p{};
	std::vector<double> ori_w, w, wgrid;
	std::vector<uint16_t> perm;
	DevBuf d_ori_w, d_w, d_wgrid, d_perm, d_walk;
	DescribeTables tables{};

	// visiting order of the G x G sample grid per orientation bin (see DescribeTables::d_perm)
	void build_perm(int G) {
		const int n = G * G, R = G / 2;
		perm.assign((size_t)DESC_PERM_BINS * n, 0);
		std::vector<std::pair<std::pair<long, double>, int>> key(n);
		for (int b = 0; b < DESC_PERM_BINS; b++) {
			const double th = (b + 0.5) * (3.14159265358979323846 / DESC_PERM_BINS), c = std::cos(th), s = std::sin(th);
			for (int k = 0; k < n; k++) {
				const int gy = k / G - R, gx = k % G - R;
				const double y = s * gx + c * gy, x = c * gx - s * gy;
				key[k] = {{(long)std::floor(y * 4.0 + 1e-9), x}, k};     // bands a quarter of a sample spacing high
			}
			std::sort(key.begin(), key.end());
			for (int k = 0; k < n; k++) perm[(size_t)b * n + k] = (uint16_t)key[k].second;
		}
	}

	void build(const bs_ori_cfg& o, const bs_desc_cfg& d) {
		p.ori_variant = o.variant; p.ori_radius = o.radius; p.ori_sample_width = o.sample_width;
		p.ori_radius_to_scale = o.object_radius_to_scale; p.ori_period = o.sample_period; p.ori_window = o.window_size;
		p.ori_weighted = o.weight_sigma != 0 ? 1 : 0;
		const int sw = 2 * o.radius + 1;
		ori_w.assign((size_t)sw * sw, 0.0);
		if (p.ori_weighted) gaussian2d_odd(o.weight_sigma, o.radius, ori_w.data());

		p.variant = d.variant; p.grid = d.width_large_grid; p.sub = d.width_sub_region; p.sample_width = d.width_sample;
		p.use_haar = d.use_haar ? 1 : 0;
		p.overlap = d.variant == 1 ? d.overlap : 0;
		p.dof = d.width_large_grid * d.width_large_grid * 4;
		const int region = d.width_large_grid * d.width_sub_region;
		p.radius_descriptor = region / 2 + p.overlap;
		p.ori_object_radius_mult = 2.0;              // BoofDefaults.SURF_SCALE_TO_RADIUS (WrapDetectDescribeSurf.java:103)
		p.out_stride = p.dof; p.out_offset = 0; p.finish = 1;
		if (region + 2 * p.overlap > 255) fail(BS_ERR_ARG, "descriptor sample grid too large");
		build_perm(region + 2 * p.overlap);
		if (d.variant == 0) {
			// weight = gaussianWidth(weightSigma, radius*2) / weight.get(radius, radius)
			const int radius = region / 2, width = radius * 2;
			if (width != region) fail(BS_ERR_ARG, "widthLargeGrid*widthSubRegion must be even for the SURF descriptor weights");
			w.assign((size_t)width * width, 0.0);
			gaussian_width_table(d.weight_sigma, width, w.data());
			const double div = w[(size_t)radius * width + radius];
			for (double& v : w) v /= div;
			wgrid.clear();
		} else {
			wgrid.assign((size_t)d.width_large_grid * d.width_large_grid, 0.0);
			gaussian_width_table(d.sigma_large_grid, d.width_large_grid, wgrid.data());
			const int wsub = d.width_sub_region + 2 * d.overlap;
			w.assign((size_t)wsub * wsub, 0.0);
			gaussian_width_table(d.sigma_sub_region, wsub, w.data());
			const int rg = d.width_large_grid / 2;      // Kernel2D.getRadius() = width/2
			double div = wgrid[(size_t)rg * d.width_large_grid + rg];
			for (double& v : wgrid) v /= div;
			const int rs = wsub / 2;
			div = w[(size_t)rs * wsub + rs];
			for (double& v : w) v /= div;
		}
	}
	void upload() {
		auto up = [](DevBuf& b, const std::vector<double>& v) -> const double* {
			if (v.empty()) return nullptr;
			b.reserve(v.size() * sizeof(double));
			h2d_sync(b.p, v.data(), v.size() * sizeof(double));
			return b.as<double>();
		};
		tables.d_ori_w = p.ori_weighted ? up(d_ori_w, ori_w) : nullptr;
		tables.d_w = up(d_w, w);
		tables.d_wgrid = up(d_wgrid, wgrid);
		d_perm.reserve(perm.size() * sizeof(uint16_t));
		h2d_sync(d_perm.p, perm.data(), perm.size() * sizeof(uint16_t));
		tables.d_perm = d_perm.as<uint16_t>();
		tables.d_walk = nullptr; tables.walk_cap = 0;
	}
	// scratch of the sliding-window walks (DescribeTables::d_walk) for up to `features` features per launch, at most 6 GB
	// limit > 0 (bs_surf_set_walk_scratch): never more than that many features, also when the scratch is already larger
	void reserve_walk(long features, long limit = 0) {
		if (p.ori_variant != 1) return;
		const size_t per = ori_w.size() * 32;
		long cap = std::max<long>(1, std::min<long>(features, (long)((6UL << 30) / per)));
		if (limit > 0) cap = std::min(cap, limit);
		if (cap > tables.walk_cap) {
			d_walk.reserve((size_t)cap * per);
			tables.d_walk = d_walk.as<double>(); tables.walk_cap = cap;
		} else if (limit > 0 && tables.d_walk) {
			tables.walk_cap = cap;
		}
	}
	void release() { d_ori_w.release(); d_w.release(); d_wgrid.release(); d_perm.release(); d_walk.release(); }
};

struct PoolBufs {
	DevBuf xy, scale, meta, response, frame, angle, desc, white;
	Pool view{};
	void reserve(long cap, int dof) {
		if (cap < 1) cap = 1;
		xy.reserve(cap * 16); scale.reserve(cap * 8); meta.reserve(cap * 16); response.reserve(cap * 4);
		frame.reserve(cap * 4); angle.reserve(cap * 8); desc.reserve((size_t)cap * dof * 8); white.reserve(cap);
		view.xy = xy.as<double>(); view.scale = scale.as<double>(); view.meta = meta.as<int32_t>();
		view.response = response.as<float>(); view.frame = frame.as<int32_t>(); view.angle = angle.as<double>();
		view.desc = desc.as<double>(); view.white = white.as<uint8_t>(); view.cap = cap;
	}
	void release() {
		xy.release(); scale.release(); meta.release(); response.release(); frame.release(); angle.release(); desc.release();
		white.release();
	}
};

}  // namespace

// ====================================================================================================== bs_surf
// bs_surf_retain: one frame's features, copied out of the pool so that they survive the next detect (include/boofsurf.h)
struct bs_features {
	int device = 0, n = 0, dof = 64;
	DevBuf xy, scale, angle, desc, white, offsets;
};

struct bs_surf {
	int device = 0;
	bs_fh_cfg fh{}; bs_ori_cfg ori{}; bs_desc_cfg desc{};
	int pixel_type = BS_PIXEL_F32;
	int bands = 0;                // 0: grey SURF; >= 1: colour SURF over a Planar image of that many bands
	int out_dof() const { return ds.p.dof * std::max(bands, 1); }
	cudaStream_t stream = nullptr;
	cudaEvent_t done = nullptr;
	// A device-resident batch is worked on by up to MAX_WAYS streams ("ways"), each with its own integral / intensity /
	// staging buffers, group after group in turn: the latency-bound kernels of one group (NMS, compaction, the tails of the
	// persistent kernels) overlap the throughput-bound ones of another.  All ways write one pool: the compaction of a chunk
	// continues the frame offsets of the chunk before it, so it waits for that chunk's compaction event.  Way 0 runs on
	// the caller's stream.  (Measured at 1080p, 1024 frames: see DESIGN.md section 4.)
	static constexpr int MAX_WAYS = 4;
	struct Way {
		DevBuf d_ii, d_inten, d_stage, d_counts, d_offsets, d_keys, d_keep, d_scan /* band partials of the parallel integral scan */;
		cudaStream_t st = nullptr;                 // internal stream (ways >= 1)
		cudaEvent_t ev_compact = nullptr, ev_done = nullptr;
		void release() { d_ii.release(); d_inten.release(); d_stage.release(); d_counts.release(); d_offsets.release(); d_keys.release(); d_keep.release(); d_scan.release(); }
	};
	Way ways[MAX_WAYS];
	int n_ways = 3, ways_used = 1;
	cudaEvent_t ev_in = nullptr;
	DescribeSetup ds;
	// workspace
	int chunk_req = 0; long pool_req = 0; long walk_req = 0;
	int W = 0, H = 0, chunk = 0;
	Plan plan{};
	DevBuf d_plan, d_in, d_frame_offset;
	DevBuf d_gray, d_ii_bands;    // colour SURF: band average (pixel type) and the band integrals [band][group][H][W]
	PoolBufs pool;
	void* h_pinned = nullptr; size_t h_pinned_bytes = 0;
	// last result
	int batch = 0;
	bool pending = false;
	std::vector<int> h_frame_offset;
	// optional per-stage timing: a pair of CUDA events on the launching stream around every stage launch
	bool timing = false;
	std::vector<cudaEvent_t> ev;          // event pool
	std::vector<int> ev_stage;            // stage of pair k (events 2k, 2k+1)
	double stage_ms[BS_N_STAGES] = {0, 0, 0, 0, 0};
	long stage_launches[BS_N_STAGES] = {0, 0, 0, 0, 0};
	void stage_begin(int stage, cudaStream_t st) {
		if (!timing) return;
		while (ev.size() < 2 * (ev_stage.size() + 1)) { cudaEvent_t e; BS_CUDA(cudaEventCreate(&e)); ev.push_back(e); }
		BS_CUDA(cudaEventRecord(ev[2 * ev_stage.size()], st));
		ev_stage.push_back(stage);
	}
	void stage_end(cudaStream_t st) {
		if (!timing) return;
		BS_CUDA(cudaEventRecord(ev[2 * ev_stage.size() - 1], st));
	}
	void collect_timing() {
		for (size_t k = 0; k < ev_stage.size(); k++) {
			float ms = 0;
			BS_CUDA(cudaEventElapsedTime(&ms, ev[2 * k], ev[2 * k + 1]));
			stage_ms[ev_stage[k]] += ms;
			stage_launches[ev_stage[k]] += 1;
		}
		ev_stage.clear();
	}

	// bs_surf_detect_host_batch: workers, each with its own stream and workspace.  A lane's turn is upload -> kernels ->
	// download in stream order; with three or more lanes the GPU always has the kernels of another turn queued while one lane
	// uploads and one downloads (two lanes leave it idle for the difference between copy and kernel time).
	static constexpr int MAX_LANES = 6;
	bs_surf* lane[MAX_LANES] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
	int n_lanes = 6;

	int group = 0;
	// integral images of `n` frames: integer pixel types take the parallel scan (exact under any association); float keeps the
	// reference's serial order — measured: a parallel-association float integral is 6.25 vs 7.05 us per 1080p frame and moves
	// 85 % of the descriptors by more than north_star's 1e-4 (float integrals of this size carry rounding noise of several units,
	// another association is another noise realisation), so it is not offered
	template <typename TI, typename TII>
	void integral(const TI* in, long frame_stride, int stride, TII* out, int n, DevBuf& scan_scratch, cudaStream_t st) {
		if (!std::is_same<TII, float>::value) {
			scan_scratch.reserve(integral_scan_scratch_bytes(W, H, n));
			if (launch_integral_scan<TI, TII>(in, frame_stride, stride, out, scan_scratch.p, W, H, n, st)) return;
		}
		launch_integral<TI, TII>(in, frame_stride, stride, out, W, H, n, st);
	}
	template <typename TI, typename TII>
	void run_typed(const void* d_pixels, long frame_stride, int stride, int nbatch, cudaStream_t st);
	template <typename TI, typename TII>
	void run_planar_typed(const void* d_pixels, long frame_stride, long band_stride, int stride, int nbatch, cudaStream_t st);
	void ensure(int w, int h, int nbatch);
	void run(const void* d_pixels, long frame_stride, int stride, int w, int h, int nbatch, cudaStream_t st);
	void run_planar(const void* d_pixels, long frame_stride, long band_stride, int stride, int w, int h, int nbatch, cudaStream_t st);
	void finish();
};

void bs_surf::ensure(int w, int h, int nbatch) {
	if (w <= 0 || h <= 0) fail(BS_ERR_ARG, "image width and height must be positive");
	if (w > 32767 || h > 32767) fail(BS_ERR_ARG, "image larger than 32767 (QueueCorner stores int16 coordinates)");
	if (w != W || h != H) {
		plan = make_plan(fh, w, h);
		d_plan.reserve(sizeof(Plan));
		h2d_sync(d_plan.p, &plan, sizeof(Plan));
		W = w; H = h;
	}
	// frames per wave: measured 5001 / 5209 / 5317 frames/s at 8 / 16 / 32 (1080p): launches must be big enough to fill 148 SMs
	// and amortise the tails of the persistent kernels; capped so that the intensity layers stay around 1.5 GB
	int c = chunk_req > 0 ? chunk_req : (int)std::max<long>(1, std::min<long>(32, (3L << 29) / std::max<long>(plan.inten_per_frame * 4, 1)));
	c = std::min(c, nbatch);
	chunk = std::max(c, 1);
	const size_t px = (size_t)W * H;
	// integral group: a multiple of the chunk, about 32 frames (or ~1 GB of integral images, whichever is smaller)
	static const int group_cap = getenv("BS_GROUP") ? std::max(1, atoi(getenv("BS_GROUP"))) : 32;
	int gframes = (int)std::max<size_t>(1, std::min<size_t>(group_cap, ((size_t)1 << 30) / (px * 4)));
	gframes = std::max(chunk, (gframes / chunk) * chunk);
	group = std::min(std::max(gframes, chunk), std::max(nbatch, 1));
	if (group < chunk) group = chunk;
	// the walk scratch of the sliding-window orientation and the colour buffers exist once per handle: one way for those
	const int groups = (nbatch + group - 1) / group;
	ways_used = (ds.p.ori_variant == 1 || bands > 0) ? 1 : std::max(1, std::min(std::min(n_ways, (int)MAX_WAYS), groups));
	const size_t nctas = (size_t)std::max(plan.ctas_per_frame, 1) * chunk;
	for (int k = 0; k < ways_used; k++) {
		Way& w = ways[k];
		w.d_ii.reserve(px * 4 * group);
		w.d_inten.reserve((size_t)std::max<long>(plan.inten_per_frame, 1) * 4 * chunk);
		w.d_stage.reserve(nctas * NMS_CTA_BLOCKS * sizeof(RawPoint));
		w.d_counts.reserve(nctas * sizeof(int));
		w.d_offsets.reserve((nctas + chunk) * sizeof(int));      // + one frame total per frame of the chunk
		if (plan.max_per_level > 0) {          // SelectNBestFeatures scratch: one key and one flag per NMS block
			w.d_keys.reserve(nctas * NMS_CTA_BLOCKS * sizeof(unsigned));
			w.d_keep.reserve(nctas * NMS_CTA_BLOCKS);
		}
	}
	d_frame_offset.reserve((size_t)(nbatch + 1) * sizeof(int));
	long cap = pool_req > 0 ? pool_req : (long)nbatch * std::max<long>((long)px / 64, 256);
	pool.reserve(cap, out_dof());
	ds.reserve_walk(std::min<long>(cap, (long)chunk * std::max<long>((long)px / 96, 256)), walk_req);   // a chunk's features (~3x the synthetic density)
	if (bands > 0) {
		d_gray.reserve(px * elem_size(pixel_type) * group);
		d_ii_bands.reserve(px * 4 * group * bands);
	}
}

// the whole pipeline for a device-resident batch, all on `st`.  Two granularities:
//   group  (<= group frames): integral images.  The ordered float scan only has rows x frames / columns x frames worth of
//          parallelism, so it wants many frames per launch;
//   chunk  (<= chunk frames): Fast-Hessian -> NMS -> compaction -> describe.  Small, so that the intensity layers a chunk
//          writes are still in L2 when the NMS kernel reads them.
template <typename TI, typename TII>
void bs_surf::run_typed(const void* d_pixels, long frame_stride, int stride, int nbatch, cudaStream_t st) {
	int* fo = d_frame_offset.as<int>();
	const long px = (long)W * H;
	const int nw = ways_used;
	if (nw > 1) {                                           // the other ways start once the caller's stream has the input (and fo[0])
		BS_CUDA(cudaEventRecord(ev_in, st));
		for (int k = 1; k < nw; k++) BS_CUDA(cudaStreamWaitEvent(ways[k].st, ev_in, 0));
	}
	int gi = 0, prev_way = -1;
	for (int g0 = 0; g0 < nbatch; g0 += group, gi++) {
		const int g = std::min(group, nbatch - g0);
		const int wi = gi % nw;
		Way& w = ways[wi];
		cudaStream_t ws = wi == 0 ? st : w.st;
		TII* ii = w.d_ii.as<TII>();
		const TI* in = static_cast<const TI*>(d_pixels) + (long)g0 * frame_stride;
		stage_begin(0, ws);
		integral<TI, TII>(in, frame_stride, stride, ii, g, w.d_scan, ws);
		stage_end(ws);
		for (int f0 = g0; f0 < g0 + g; f0 += chunk) {
			const int c = std::min(chunk, g0 + g - f0);
			const TII* cii = ii + (long)(f0 - g0) * px;
			stage_begin(1, ws);
			launch_hessian_plan<TII>(cii, plan, d_plan.as<Plan>(), w.d_inten.as<float>(), c, ws);
			stage_end(ws);
			stage_begin(2, ws);
			launch_nms(plan, d_plan.as<Plan>(), w.d_inten.as<float>(), c, w.d_stage.as<RawPoint>(), w.d_counts.as<int>(), w.d_keys.as<unsigned>(),
			           w.d_keep.as<unsigned char>(), ws);
			stage_end(ws);
			if (prev_way >= 0 && prev_way != wi) BS_CUDA(cudaStreamWaitEvent(ws, ways[prev_way].ev_compact, 0));   // fo[f0] is final
			stage_begin(3, ws);
			launch_compact(plan, c, f0, w.d_stage.as<RawPoint>(), w.d_counts.as<int>(), w.d_offsets.as<int>(), fo, pool.view, ws);
			stage_end(ws);
			if (nw > 1) BS_CUDA(cudaEventRecord(w.ev_compact, ws));
			prev_way = wi;
			stage_begin(4, ws);
			launch_describe<TII>(cii, W, H, ds.p, ds.tables, pool.view, fo + f0, fo + f0 + c, f0, true, true, ws);
			stage_end(ws);
		}
	}
	for (int k = 1; k < nw; k++) {                          // everything later on `st` (and the `done` event) follows all ways
		BS_CUDA(cudaEventRecord(ways[k].ev_done, ways[k].st));
		BS_CUDA(cudaStreamWaitEvent(st, ways[k].ev_done, 0));
	}
}

// Colour SURF for a device-resident batch of Planar frames (band b of frame f at d_pixels + f*frame_stride + b*band_stride):
// SurfPlanar_to_DetectDescribePoint.detect (F/abst/feature/detdesc/SurfPlanar_to_DetectDescribePoint.java:58-74: band average,
// grey integral, one integral per band) -> DetectDescribeSurfPlanar.detect/describe (F/alg/feature/detdesc/
// DetectDescribeSurfPlanar.java:87-116: detection and orientation on the grey integral, object radius = scale) ->
// DescribePointSurfPlanar.describe (F/alg/feature/describe/DescribePointSurfPlanar.java:97-111: raw sums per band,
// one L2 normalisation over all bands, Laplacian sign from the grey integral).
template <typename TI, typename TII>
void bs_surf::run_planar_typed(const void* d_pixels, long frame_stride, long band_stride, int stride, int nbatch, cudaStream_t st) {
	int* fo = d_frame_offset.as<int>();
	const long px = (long)W * H;
	const int dof = ds.p.dof, total = dof * bands;
	for (int g0 = 0; g0 < nbatch; g0 += group) {
		const int g = std::min(group, nbatch - g0);
		TII* ii = ways[0].d_ii.as<TII>();
		TII* bii = d_ii_bands.as<TII>();
		const TI* in = static_cast<const TI*>(d_pixels) + (long)g0 * frame_stride;
		stage_begin(0, st);
		launch_average_bands<TI>(in, frame_stride, band_stride, stride, bands, d_gray.as<TI>(), W, H, g, st);
		integral<TI, TII>(d_gray.as<TI>(), px, W, ii, g, ways[0].d_scan, st);
		for (int b = 0; b < bands; b++) integral<TI, TII>(in + b * band_stride, frame_stride, stride, bii + (long)b * group * px, g, ways[0].d_scan, st);
		stage_end(st);
		for (int f0 = g0; f0 < g0 + g; f0 += chunk) {
			const int c = std::min(chunk, g0 + g - f0);
			const TII* cii = ii + (long)(f0 - g0) * px;
			stage_begin(1, st);
			launch_hessian_plan<TII>(cii, plan, d_plan.as<Plan>(), ways[0].d_inten.as<float>(), c, st);
			stage_end(st);
			stage_begin(2, st);
			launch_nms(plan, d_plan.as<Plan>(), ways[0].d_inten.as<float>(), c, ways[0].d_stage.as<RawPoint>(), ways[0].d_counts.as<int>(), ways[0].d_keys.as<unsigned>(),
			           ways[0].d_keep.as<unsigned char>(), st);
			stage_end(st);
			stage_begin(3, st);
			launch_compact(plan, c, f0, ways[0].d_stage.as<RawPoint>(), ways[0].d_counts.as<int>(), ways[0].d_offsets.as<int>(), fo, pool.view, st);
			stage_end(st);
			stage_begin(4, st);
			DescribeParams po = ds.p;
			po.ori_object_radius_mult = 1.0;
			launch_describe<TII>(cii, W, H, po, ds.tables, pool.view, fo + f0, fo + f0 + c, f0, true, false, st);
			for (int b = 0; b < bands; b++) {
				DescribeParams pb = ds.p;
				pb.out_stride = total; pb.out_offset = b * dof; pb.finish = 0;
				launch_describe<TII>(bii + (long)b * group * px + (long)(f0 - g0) * px, W, H, pb, ds.tables, pool.view, fo + f0, fo + f0 + c, f0,
				                     false, true, st);
			}
			launch_planar_finish<TII>(cii, W, H, total, pool.view, fo + f0, fo + f0 + c, f0, st);
			stage_end(st);
		}
	}
}

void bs_surf::run_planar(const void* d_pixels, long frame_stride, long band_stride, int stride, int w, int h, int nbatch, cudaStream_t st) {
	if (bands <= 0) fail(BS_ERR_STATE, "this handle was created for single-band images (bs_surf_create)");
	if (nbatch <= 0) fail(BS_ERR_ARG, "batch must be >= 1");
	if (stride < w) fail(BS_ERR_ARG, "stride must be >= width");
	ensure(w, h, nbatch);
	batch = nbatch;
	BS_CUDA(cudaMemsetAsync(d_frame_offset.p, 0, sizeof(int), st));
	if (pixel_type == BS_PIXEL_F32) run_planar_typed<float, float>(d_pixels, frame_stride, band_stride, stride, nbatch, st);
	else if (pixel_type == BS_PIXEL_U8) run_planar_typed<uint8_t, int32_t>(d_pixels, frame_stride, band_stride, stride, nbatch, st);
	else run_planar_typed<int32_t, int32_t>(d_pixels, frame_stride, band_stride, stride, nbatch, st);
	BS_CUDA(cudaEventRecord(done, st));
	pending = true;
}

void bs_surf::run(const void* d_pixels, long frame_stride, int stride, int w, int h, int nbatch, cudaStream_t st) {
	if (bands > 0) fail(BS_ERR_STATE, "this handle was created for Planar images (bs_surf_detect_planar)");
	if (nbatch <= 0) fail(BS_ERR_ARG, "batch must be >= 1");
	if (stride < w) fail(BS_ERR_ARG, "stride must be >= width");
	ensure(w, h, nbatch);
	batch = nbatch;
	BS_CUDA(cudaMemsetAsync(d_frame_offset.p, 0, sizeof(int), st));
	if (pixel_type == BS_PIXEL_F32) run_typed<float, float>(d_pixels, frame_stride, stride, nbatch, st);
	else if (pixel_type == BS_PIXEL_U8) run_typed<uint8_t, int32_t>(d_pixels, frame_stride, stride, nbatch, st);
	else run_typed<int32_t, int32_t>(d_pixels, frame_stride, stride, nbatch, st);
	BS_CUDA(cudaEventRecord(done, st));
	pending = true;
}

void bs_surf::finish() {
	if (!pending) return;
	BS_CUDA(cudaEventSynchronize(done));
	h_frame_offset.assign(batch + 1, 0);
	BS_CUDA(cudaMemcpy(h_frame_offset.data(), d_frame_offset.p, (size_t)(batch + 1) * sizeof(int), cudaMemcpyDeviceToHost));
	pending = false;
	collect_timing();
}

// ====================================================================================================== extern "C": misc
extern "C" {

void bs_fh_cfg_default(bs_fh_cfg* c) {
	c->detect_threshold = 1; c->extract_radius = 2; c->max_features_per_scale = -1; c->initial_sample_size = 1;
	c->initial_size = 9; c->scales_per_octave = 4; c->octaves = 4; c->scale_step_size = 6;
}
void bs_ori_cfg_default(bs_ori_cfg* c, int variant) {
	c->variant = variant; c->object_radius_to_scale = 1.0 / 2.0; c->sample_width = 6; c->weight_sigma = -1;
	c->window_size = M_PI / 3.0;
	if (variant == 1) { c->radius = 8; c->sample_period = 0.65; }
	else { c->radius = 6; c->sample_period = 1; }
}
void bs_desc_cfg_default(bs_desc_cfg* c, int variant) {
	c->variant = variant; c->width_large_grid = 4; c->width_sub_region = 5; c->width_sample = 3; c->use_haar = 0;
	c->weight_sigma = 4.5; c->overlap = 2; c->sigma_large_grid = 2.5; c->sigma_sub_region = 2.5;
}
const char* bs_last_error(void) { return last_error_cstr(); }
const char* bs_version(void) { return "boofsurf-b200 0.1 (sm_100a)"; }
int bs_device_count(int* n) {
	return guarded([&] { BS_CUDA(cudaGetDeviceCount(n)); });
}
long bs_launch_count(void) { return launch_count(); }

// ====================================================================================================== extern "C": bs_surf
int bs_surf_create(int device, const bs_fh_cfg* fh, const bs_ori_cfg* ori, const bs_desc_cfg* desc, int pixel_type, bs_surf** out) {
	return guarded([&] {
		if (!out) fail(BS_ERR_ARG, "out is null");
		*out = nullptr;
		if (pixel_type != BS_PIXEL_F32 && pixel_type != BS_PIXEL_U8 && pixel_type != BS_PIXEL_S32)
			fail(BS_ERR_ARG, "Unknown input type");   // GIntegralImageOps.getIntegralType :35-46
		bs_fh_cfg f; bs_ori_cfg o; bs_desc_cfg d;
		if (fh) f = *fh; else bs_fh_cfg_default(&f);
		if (desc) d = *desc; else bs_desc_cfg_default(&d, ori ? ori->variant : 0);
		if (ori) o = *ori; else bs_ori_cfg_default(&o, d.variant);
		check_fh_cfg(f); check_ori_cfg(o); check_desc_cfg(d);
		int ndev = 0;
		BS_CUDA(cudaGetDeviceCount(&ndev));
		if (device < 0 || device >= ndev) fail(BS_ERR_ARG, "no such CUDA device");
		DeviceGuard g(device);
		std::unique_ptr<bs_surf> s(new bs_surf());
		s->device = device; s->fh = f; s->ori = o; s->desc = d; s->pixel_type = pixel_type;
		BS_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
		BS_CUDA(cudaEventCreateWithFlags(&s->done, cudaEventDisableTiming));
		BS_CUDA(cudaEventCreateWithFlags(&s->ev_in, cudaEventDisableTiming));
		for (int k = 0; k < bs_surf::MAX_WAYS; k++) {
			if (k > 0) BS_CUDA(cudaStreamCreateWithFlags(&s->ways[k].st, cudaStreamNonBlocking));
			BS_CUDA(cudaEventCreateWithFlags(&s->ways[k].ev_compact, cudaEventDisableTiming));
			BS_CUDA(cudaEventCreateWithFlags(&s->ways[k].ev_done, cudaEventDisableTiming));
		}
		if (const char* e = getenv("BS_WAYS")) s->n_ways = std::min(std::max(atoi(e), 1), (int)bs_surf::MAX_WAYS);
		if (const char* e = getenv("BS_LANES")) s->n_lanes = std::min(std::max(atoi(e), 2), (int)bs_surf::MAX_LANES);
		s->ds.build(o, d);
		s->ds.upload();
		*out = s.release();
	});
}

int bs_surf_create_planar(int device, const bs_fh_cfg* fh, const bs_ori_cfg* ori, const bs_desc_cfg* desc, int pixel_type, int num_bands,
                          bs_surf** out) {
	if (num_bands < 1 || num_bands > 16) {
		set_last_error("number of bands must be in [1,16]");
		if (out) *out = nullptr;
		return BS_ERR_ARG;
	}
	const int rc = bs_surf_create(device, fh, ori, desc, pixel_type, out);
	if (rc == BS_OK) (*out)->bands = num_bands;
	return rc;
}

int bs_surf_detect_planar_device(bs_surf* s, const void* d_pixels, long frame_stride, long band_stride, int stride, int width, int height,
                                 int batch, void* stream) {
	return guarded([&] {
		if (!s || !d_pixels) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(s->device);
		s->run_planar(d_pixels, frame_stride, band_stride, stride, width, height, batch, stream ? static_cast<cudaStream_t>(stream) : s->stream);
	});
}

int bs_surf_detect_planar(bs_surf* s, const void* const* bands, int num_bands, int start_index, int stride, int width, int height) {
	return guarded([&] {
		if (!s || !bands) fail(BS_ERR_ARG, "null argument");
		if (s->bands <= 0) fail(BS_ERR_STATE, "this handle was created for single-band images (bs_surf_create)");
		// SurfPlanar_to_DetectDescribePoint.detect :59-61
		if (num_bands != s->bands)
			fail(BS_ERR_ARG, "Unexpected number of bands. Expected " + std::to_string(s->bands) + " found " + std::to_string(num_bands));
		if (width <= 0 || height <= 0) fail(BS_ERR_ARG, "image width and height must be positive");
		if (stride < width || start_index < 0) fail(BS_ERR_ARG, "bad stride/startIndex");
		DeviceGuard g(s->device);
		const size_t es = elem_size(s->pixel_type);
		const size_t dense = (size_t)width * height * es;
		s->d_in.reserve(dense * num_bands);
		for (int b = 0; b < num_bands; b++) {
			if (!bands[b]) fail(BS_ERR_ARG, "null band");
			const char* src = static_cast<const char*>(bands[b]) + (size_t)start_index * es;
			BS_CUDA(cudaMemcpy2DAsync(static_cast<char*>(s->d_in.p) + dense * b, (size_t)width * es, src, (size_t)stride * es, (size_t)width * es,
			                          height, cudaMemcpyHostToDevice, s->stream));
		}
		const long px = (long)width * height;
		s->run_planar(s->d_in.p, px * num_bands, px, width, width, height, 1, s->stream);
		s->finish();
	});
}

int bs_surf_destroy(bs_surf* s) {
	if (!s) return BS_OK;
	for (int k = 0; k < bs_surf::MAX_LANES; k++) {
		if (s->lane[k]) bs_surf_destroy(s->lane[k]);
		s->lane[k] = nullptr;
	}
	return guarded([&] {
		DeviceGuard g(s->device);
		cudaDeviceSynchronize();
		s->ds.release(); s->pool.release();
		s->d_plan.release(); s->d_in.release(); s->d_frame_offset.release();
		s->d_gray.release(); s->d_ii_bands.release();
		if (s->h_pinned) cudaFreeHost(s->h_pinned);
		if (s->done) cudaEventDestroy(s->done);
		if (s->ev_in) cudaEventDestroy(s->ev_in);
		for (int k = 0; k < bs_surf::MAX_WAYS; k++) {
			s->ways[k].release();
			if (s->ways[k].ev_compact) cudaEventDestroy(s->ways[k].ev_compact);
			if (s->ways[k].ev_done) cudaEventDestroy(s->ways[k].ev_done);
			if (s->ways[k].st) cudaStreamDestroy(s->ways[k].st);
		}
		for (cudaEvent_t e : s->ev) cudaEventDestroy(e);
		if (s->stream) cudaStreamDestroy(s->stream);
		delete s;
	});
}

int bs_surf_set_workspace(bs_surf* s, int chunk_frames, long pool_features) {
	return guarded([&] {
		if (!s) fail(BS_ERR_ARG, "handle is null");
		if (chunk_frames < 0 || pool_features < 0) fail(BS_ERR_ARG, "workspace sizes must be >= 0");
		s->chunk_req = chunk_frames; s->pool_req = pool_features;
	});
}

int bs_surf_set_walk_scratch(bs_surf* s, long max_features) {
	return guarded([&] {
		if (!s) fail(BS_ERR_ARG, "handle is null");
		if (max_features < 0) fail(BS_ERR_ARG, "max_features must be >= 0");
		s->finish();
		s->walk_req = max_features;
	});
}

int bs_surf_set_ways(bs_surf* s, int ways) {
	return guarded([&] {
		if (!s) fail(BS_ERR_ARG, "handle is null");
		if (ways < 1 || ways > bs_surf::MAX_WAYS) fail(BS_ERR_ARG, "ways must be 1..4");
		s->finish();
		s->n_ways = ways;
	});
}

int bs_surf_dof(const bs_surf* s) { return s ? s->out_dof() : 0; }

int bs_surf_set_timing(bs_surf* s, int on) {
	return guarded([&] {
		if (!s) fail(BS_ERR_ARG, "handle is null");
		s->finish();
		s->timing = on != 0;
		for (int k = 0; k < BS_N_STAGES; k++) { s->stage_ms[k] = 0; s->stage_launches[k] = 0; }
	});
}

int bs_surf_get_timing(bs_surf* s, double* stage_ms, long* stage_calls) {
	return guarded([&] {
		if (!s) fail(BS_ERR_ARG, "handle is null");
		DeviceGuard g(s->device);
		s->finish();
		for (int k = 0; k < BS_N_STAGES; k++) {
			if (stage_ms) stage_ms[k] = s->stage_ms[k];
			if (stage_calls) stage_calls[k] = s->stage_launches[k];
		}
	});
}

int bs_surf_detect_device(bs_surf* s, const void* d_pixels, long frame_stride, int stride, int width, int height, int batch,
                          void* stream) {
	return guarded([&] {
		if (!s || !d_pixels) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(s->device);
		s->run(d_pixels, frame_stride, stride, width, height, batch, stream ? static_cast<cudaStream_t>(stream) : s->stream);
	});
}

int bs_host_register(void* p, size_t bytes) {
	return guarded([&] {
		if (!p || bytes == 0) fail(BS_ERR_ARG, "null or empty host range");
		BS_CUDA(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
	});
}
int bs_host_unregister(void* p) {
	return guarded([&] {
		if (!p) fail(BS_ERR_ARG, "null pointer");
		BS_CUDA(cudaHostUnregister(p));
	});
}

int bs_surf_detect_host_batch(bs_surf* s, const void* pixels, long frame_stride, int stride, int width, int height, int batch,
                              long cap, int32_t* frame_offset, double* xy, double* scale, double* angle, double* desc,
                              uint8_t* white, long* total) {
	return guarded([&] {
		if (!s || !pixels || !frame_offset) fail(BS_ERR_ARG, "null argument");
		if (width <= 0 || height <= 0) fail(BS_ERR_ARG, "image width and height must be positive");
		if (batch < 0 || cap < 0 || stride < width) fail(BS_ERR_ARG, "bad batch/cap/stride");
		if (frame_stride < (long)stride * (height - 1) + width) fail(BS_ERR_ARG, "frames overlap");
		// the worker lanes are grey handles: a planar handle's bands and 64*bands descriptor rows do not fit them
		if (s->bands > 0) fail(BS_ERR_STATE, "bs_surf_detect_host_batch is not available on a planar (colour) handle");
		DeviceGuard g(s->device);
		s->finish();
		const int NL = s->n_lanes;
		for (int k = 0; k < NL; k++)
			if (!s->lane[k]) {
				bs_surf* h = nullptr;
				if (bs_surf_create(s->device, &s->fh, &s->ori, &s->desc, s->pixel_type, &h) != BS_OK) fail(BS_ERR_CUDA, last_error_cstr());
				s->lane[k] = h;
			}
		const size_t es = elem_size(s->pixel_type);
		const int dof = s->out_dof();
		static const int sub_env = getenv("BS_HOST_SUB") ? std::max(1, atoi(getenv("BS_HOST_SUB"))) : 16;
		// frames per lane turn.  Measured at 1080p (e2e frames/s, lanes x sub): 4x32 5354, 3x32 5352, 6x32 5184, 4x16 5269, 6x16 5544,
		// 4x64 5157 — against a PCIe ceiling of 55.6 GB/s one way, 45.7 GB/s each way at once (tools/probe/pcie_bw.py): 8.3 MB in and
		// 5.5 MB out per frame make the host path PCIe-bound at about 5.5 k frames/s
		const int sub = sub_env;
		frame_offset[0] = 0;
		long base = 0;                             // features of the frames drained so far
		struct Turn { int first = 0, n = 0; bool busy = false; } turn[bs_surf::MAX_LANES];
		auto drain = [&](int k) {
			Turn& t = turn[k];
			if (!t.busy) return;
			bs_surf* h = s->lane[k];
			h->finish();                           // waits for this lane only; the other lane keeps the GPU busy
			const int* fo = h->h_frame_offset.data();
			const long found = fo[t.n];
			if (found > h->pool.view.cap) fail(BS_ERR_CAPACITY, "feature pool overflow inside a sub-batch (bs_surf_set_workspace pool_features)");
			for (int i = 0; i < t.n; i++) frame_offset[t.first + i + 1] = (int32_t)(base + fo[i + 1]);
			const long room = std::max<long>(0, std::min<long>(found, cap - base));
			if (room > 0) {
				const Pool& p = h->pool.view;
				cudaStream_t st = h->stream;   // ordered before this lane's next turn, concurrent with the other lane
				if (xy) BS_CUDA(cudaMemcpyAsync(xy + 2 * base, p.xy, (size_t)room * 16, cudaMemcpyDeviceToHost, st));
				if (scale) BS_CUDA(cudaMemcpyAsync(scale + base, p.scale, (size_t)room * 8, cudaMemcpyDeviceToHost, st));
				if (angle) BS_CUDA(cudaMemcpyAsync(angle + base, p.angle, (size_t)room * 8, cudaMemcpyDeviceToHost, st));
				if (desc) BS_CUDA(cudaMemcpyAsync(desc + (size_t)dof * base, p.desc, (size_t)room * dof * 8, cudaMemcpyDeviceToHost, st));
				if (white) BS_CUDA(cudaMemcpyAsync(white + base, p.white, (size_t)room, cudaMemcpyDeviceToHost, st));
			}
			base += found;
			t.busy = false;
		};
		int k = 0, last = 0;
		for (int f0 = 0; f0 < batch; f0 += sub, k = (k + 1) % NL) {
			drain(k);                              // normally a no-op: this lane was drained right after the previous one started
			bs_surf* h = s->lane[k];
			// the caller sized pool_features for the whole batch; a lane holds one turn of `sub` frames (twice its share: frames differ)
			h->chunk_req = s->chunk_req; h->walk_req = s->walk_req;
			h->pool_req = s->pool_req > 0 ? std::min<long>(s->pool_req, std::max<long>(1024, 2 * ((s->pool_req * sub + batch - 1) / batch))) : 0;
			const int n = std::min(sub, batch - f0);
			const size_t dense = (size_t)width * height * es;
			h->d_in.reserve(dense * sub);
			const char* src = static_cast<const char*>(pixels) + (size_t)f0 * frame_stride * es;
			if (stride == width && frame_stride == (long)width * height) {
				BS_CUDA(cudaMemcpyAsync(h->d_in.p, src, dense * n, cudaMemcpyHostToDevice, h->stream));
			} else {
				for (int i = 0; i < n; i++)
					BS_CUDA(cudaMemcpy2DAsync(static_cast<char*>(h->d_in.p) + dense * i, (size_t)width * es, src + (size_t)i * frame_stride * es,
					                          (size_t)stride * es, (size_t)width * es, height, cudaMemcpyHostToDevice, h->stream));
			}
			h->run(h->d_in.p, (long)width * height, width, width, height, n, h->stream);
			turn[k].first = f0; turn[k].n = n; turn[k].busy = true;
			last = k;
			drain((k + 1) % NL);                   // results of the oldest turn go home while the younger ones upload and compute
		}
		for (int j = 1; j <= NL; j++) drain((last + j) % NL);      // the remaining turns, oldest first
		for (int j = 0; j < NL; j++) BS_CUDA(cudaStreamSynchronize(s->lane[j]->stream));
		if (total) *total = base;
		s->batch = 0; s->pending = false;
		if (base > cap) fail(BS_ERR_CAPACITY, "more features than the host buffers hold");
	});
}

int bs_surf_detect(bs_surf* s, const void* pixels, int start_index, int stride, int width, int height) {
	return guarded([&] {
		if (!s || !pixels) fail(BS_ERR_ARG, "null argument");
		if (width <= 0 || height <= 0) fail(BS_ERR_ARG, "image width and height must be positive");
		if (stride < width || start_index < 0) fail(BS_ERR_ARG, "bad stride/startIndex");
		DeviceGuard g(s->device);
		const size_t es = elem_size(s->pixel_type);
		const size_t dense = (size_t)width * height * es;
		s->d_in.reserve(dense);
		const char* src = static_cast<const char*>(pixels) + (size_t)start_index * es;
		// sub-images arrive as a pitched copy; the device copy is dense
		BS_CUDA(cudaMemcpy2DAsync(s->d_in.p, (size_t)width * es, src, (size_t)stride * es, (size_t)width * es, height,
		                          cudaMemcpyHostToDevice, s->stream));
		s->run(s->d_in.p, (long)width * height, width, width, height, 1, s->stream);
		s->finish();
	});
}

static void check_ready(bs_surf* s) {
	if (!s) fail(BS_ERR_ARG, "handle is null");
	if (s->batch <= 0) fail(BS_ERR_STATE, "detect has not been called");
	s->finish();
}

int bs_surf_total(bs_surf* s, long* n) {
	return guarded([&] {
		DeviceGuard g(s ? s->device : 0);
		check_ready(s);
		if (n) *n = s->h_frame_offset[s->batch];
		if (s->h_frame_offset[s->batch] > s->pool.view.cap) fail(BS_ERR_CAPACITY, "feature pool too small for this batch");
	});
}

int bs_surf_count(bs_surf* s, int frame, int* n) {
	return guarded([&] {
		DeviceGuard g(s ? s->device : 0);
		check_ready(s);
		if (frame < 0 || frame >= s->batch) fail(BS_ERR_ARG, "frame out of range");
		if (n) *n = s->h_frame_offset[frame + 1] - s->h_frame_offset[frame];
		if (s->h_frame_offset[frame + 1] > s->pool.view.cap) fail(BS_ERR_CAPACITY, "feature pool too small for this batch");
	});
}

int bs_surf_read(bs_surf* s, int frame, double* xy, double* scale, double* angle, double* desc, uint8_t* white, int32_t* meta,
                 float* response) {
	return guarded([&] {
		DeviceGuard g(s ? s->device : 0);
		check_ready(s);
		if (frame < 0 || frame >= s->batch) fail(BS_ERR_ARG, "frame out of range");
		const long b = s->h_frame_offset[frame], e = s->h_frame_offset[frame + 1];
		if (e > s->pool.view.cap) fail(BS_ERR_CAPACITY, "feature pool too small for this batch");
		const size_t n = (size_t)(e - b);
		if (n == 0) return;
		const Pool& p = s->pool.view;
		const int dof = s->out_dof();
		if (xy) BS_CUDA(cudaMemcpy(xy, p.xy + 2 * b, n * 16, cudaMemcpyDeviceToHost));
		if (scale) BS_CUDA(cudaMemcpy(scale, p.scale + b, n * 8, cudaMemcpyDeviceToHost));
		if (angle) BS_CUDA(cudaMemcpy(angle, p.angle + b, n * 8, cudaMemcpyDeviceToHost));
		if (desc) BS_CUDA(cudaMemcpy(desc, p.desc + b * dof, n * dof * 8, cudaMemcpyDeviceToHost));
		if (white) BS_CUDA(cudaMemcpy(white, p.white + b, n, cudaMemcpyDeviceToHost));
		if (meta) BS_CUDA(cudaMemcpy(meta, p.meta + 4 * b, n * 16, cudaMemcpyDeviceToHost));
		if (response) BS_CUDA(cudaMemcpy(response, p.response + b, n * 4, cudaMemcpyDeviceToHost));
	});
}

int bs_surf_retain(bs_surf* s, int frame, void* stream, bs_features** out) {
	return guarded([&] {
		if (!out) fail(BS_ERR_ARG, "out is null");
		*out = nullptr;
		DeviceGuard g(s ? s->device : 0);
		check_ready(s);
		if (frame < 0 || frame >= s->batch) fail(BS_ERR_ARG, "frame out of range");
		const std::vector<int>& fo = s->h_frame_offset;
		if (fo[s->batch] > s->pool.view.cap) fail(BS_ERR_CAPACITY, "feature pool too small for this batch");
		std::unique_ptr<bs_features> f(new bs_features());
		f->device = s->device; f->dof = s->out_dof();
		const long b = fo[frame];
		const int n = fo[frame + 1] - fo[frame];
		f->n = n;
		const size_t m = (size_t)std::max(n, 1);
		f->xy.reserve(m * 16); f->scale.reserve(m * 8); f->angle.reserve(m * 8); f->desc.reserve(m * f->dof * 8); f->white.reserve(m);
		f->offsets.reserve(8);
		cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : s->stream;
		const Pool& p = s->pool.view;
		if (n > 0) {
			BS_CUDA(cudaMemcpyAsync(f->xy.p, p.xy + 2 * b, (size_t)n * 16, cudaMemcpyDeviceToDevice, st));
			BS_CUDA(cudaMemcpyAsync(f->scale.p, p.scale + b, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
			BS_CUDA(cudaMemcpyAsync(f->angle.p, p.angle + b, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
			BS_CUDA(cudaMemcpyAsync(f->desc.p, p.desc + b * f->dof, (size_t)n * f->dof * 8, cudaMemcpyDeviceToDevice, st));
			BS_CUDA(cudaMemcpyAsync(f->white.p, p.white + b, (size_t)n, cudaMemcpyDeviceToDevice, st));
		}
		const int32_t h_off[2] = {0, n};
		BS_CUDA(cudaMemcpyAsync(f->offsets.p, h_off, 8, cudaMemcpyHostToDevice, st));
		BS_CUDA(cudaStreamSynchronize(st));        // the next detect may overwrite the pool on another stream
		*out = f.release();
	});
}
int bs_features_count(const bs_features* f, int* n, int* dof) {
	return guarded([&] {
		if (!f) fail(BS_ERR_ARG, "handle is null");
		if (n) *n = f->n;
		if (dof) *dof = f->dof;
	});
}
int bs_features_view(const bs_features* f, bs_surf_device_view* v) {
	return guarded([&] {
		if (!f || !v) fail(BS_ERR_ARG, "null argument");
		v->d_xy = f->xy.as<double>(); v->d_scale = f->scale.as<double>(); v->d_angle = f->angle.as<double>();
		v->d_desc = f->desc.as<double>(); v->d_white = f->white.as<uint8_t>(); v->d_frame_offset = f->offsets.as<int32_t>();
		v->total = f->n; v->batch = 1; v->dof = f->dof;
	});
}
int bs_features_destroy(bs_features* f) {
	return guarded([&] {
		if (!f) return;
		DeviceGuard g(f->device);
		f->xy.release(); f->scale.release(); f->angle.release(); f->desc.release(); f->white.release(); f->offsets.release();
		delete f;
	});
}

int bs_surf_view(bs_surf* s, bs_surf_device_view* v) {
	return guarded([&] {
		if (!v) fail(BS_ERR_ARG, "view is null");
		DeviceGuard g(s ? s->device : 0);
		check_ready(s);
		const Pool& p = s->pool.view;
		v->d_xy = p.xy; v->d_scale = p.scale; v->d_angle = p.angle; v->d_desc = p.desc; v->d_white = p.white;
		v->d_frame_offset = s->d_frame_offset.as<int32_t>();
		v->total = s->h_frame_offset[s->batch]; v->batch = s->batch; v->dof = s->out_dof();
		if (v->total > p.cap) fail(BS_ERR_CAPACITY, "feature pool too small for this batch");
	});
}

}  // extern "C"

// ====================================================================================================== stage-level entry points
namespace {

struct TempStream {
	cudaStream_t st = nullptr;
	TempStream() { BS_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)); }
	~TempStream() { if (st) cudaStreamDestroy(st); }
};

void check_device(int device) {
	int ndev = 0;
	BS_CUDA(cudaGetDeviceCount(&ndev));
	if (device < 0 || device >= ndev) fail(BS_ERR_ARG, "no such CUDA device");
}

// upload a dense integral image of either type
void upload_ii(DevBuf& d, const void* ii, int width, int height) {
	if (!ii || width <= 0 || height <= 0) fail(BS_ERR_ARG, "bad integral image");
	if ((long)width * height >= (1L << 31)) fail(BS_ERR_ARG, "integral image of 2^31 pixels or more (the sample kernels index it with 32-bit offsets)");
	d.reserve((size_t)width * height * 4);
	h2d_sync(d.p, ii, (size_t)width * height * 4);
}

// pool with n caller-given points (frame 0) and the range [0,n] on the device
struct PointSet {
	PoolBufs pool; DevBuf range;
	void load(const double* xys, const double* angle, int n, int dof) {
		pool.reserve(n, dof);
		std::vector<double> xy((size_t)2 * n), sc(n);
		for (int i = 0; i < n; i++) { xy[2 * i] = xys[3 * i]; xy[2 * i + 1] = xys[3 * i + 1]; sc[i] = xys[3 * i + 2]; }
		h2d_sync(pool.view.xy, xy.data(), xy.size() * 8);
		h2d_sync(pool.view.scale, sc.data(), sc.size() * 8);
		BS_CUDA(cudaMemset(pool.view.frame, 0, (size_t)n * 4));
		BS_CUDA(cudaDeviceSynchronize());
		if (angle) h2d_sync(pool.view.angle, angle, (size_t)n * 8);
		const int r[2] = {0, n};
		range.reserve(8);
		h2d_sync(range.p, r, 8);
	}
	void release() { pool.release(); range.release(); }
};

}  // namespace

extern "C" {

int bs_integral(int device, int pixel_type, const void* in, int in_stride, int width, int height, void* out) {
	return guarded([&] {
		if (!in || !out) fail(BS_ERR_ARG, "null argument");
		if (width <= 0 || height <= 0) return;      // empty image: nothing to do, like the reference loops
		if (in_stride < width) fail(BS_ERR_ARG, "stride must be >= width");
		check_device(device);
		DeviceGuard g(device);
		TempStream ts;
		DevBuf d_in, d_out;
		const size_t es = elem_size(pixel_type);
		d_in.reserve((size_t)width * height * es);
		d_out.reserve((size_t)width * height * 4);
		BS_CUDA(cudaMemcpy2D(d_in.p, width * es, in, in_stride * es, width * es, height, cudaMemcpyHostToDevice));
		BS_CUDA(cudaDeviceSynchronize());
		if (pixel_type == BS_PIXEL_F32) launch_integral<float, float>(d_in.as<float>(), 0, width, d_out.as<float>(), width, height, 1, ts.st);
		else if (pixel_type == BS_PIXEL_U8 || pixel_type == BS_PIXEL_S32) {
			// integer images: the parallel scan (what the pipeline runs for them), serial passes only for shapes it does not cover
			DevBuf scratch;
			scratch.reserve(integral_scan_scratch_bytes(width, height, 1));
			bool done = pixel_type == BS_PIXEL_U8
			    ? launch_integral_scan<uint8_t, int32_t>(d_in.as<uint8_t>(), 0, width, d_out.as<int32_t>(), scratch.p, width, height, 1, ts.st)
			    : launch_integral_scan<int32_t, int32_t>(d_in.as<int32_t>(), 0, width, d_out.as<int32_t>(), scratch.p, width, height, 1, ts.st);
			if (!done) {
				if (pixel_type == BS_PIXEL_U8) launch_integral<uint8_t, int32_t>(d_in.as<uint8_t>(), 0, width, d_out.as<int32_t>(), width, height, 1, ts.st);
				else launch_integral<int32_t, int32_t>(d_in.as<int32_t>(), 0, width, d_out.as<int32_t>(), width, height, 1, ts.st);
			}
			BS_CUDA(cudaStreamSynchronize(ts.st));
			scratch.release();
		}
		else fail(BS_ERR_ARG, "Unknown input type");
		BS_CUDA(cudaStreamSynchronize(ts.st));
		BS_CUDA(cudaMemcpy(out, d_out.p, (size_t)width * height * 4, cudaMemcpyDeviceToHost));
		d_in.release(); d_out.release();
	});
}

int bs_nonmax(int device, const float* intensity, int stride, int width, int height, int radius, float threshold_max, int ignore_border,
              int max_out, int16_t* xy, int* n) {
	return guarded([&] {
		if (!n) fail(BS_ERR_ARG, "null argument");
		*n = 0;
		if (width <= 0 || height <= 0) return;
		if (!intensity || (max_out > 0 && !xy)) fail(BS_ERR_ARG, "null argument");
		if (stride < width) fail(BS_ERR_ARG, "stride must be >= width");
		if (radius < 1) fail(BS_ERR_ARG, "radius must be >= 1");          // ConfigExtract.checkValidity
		if (ignore_border < 0 || max_out < 0) fail(BS_ERR_ARG, "ignore_border and max_out must be >= 0");
		if (width > 32767 || height > 32767) fail(BS_ERR_ARG, "image larger than 32767 (QueueCorner stores int16 coordinates)");
		check_device(device);
		DeviceGuard g(device);
		TempStream ts;
		DevBuf d_img, d_cand, d_out, d_total;
		const int step = radius + 1;
		const long blocks = (long)div_up(width, step) * div_up(height, step);
		d_img.reserve((size_t)width * height * 4); d_cand.reserve((size_t)std::max<long>(blocks, 1) * 4);
		d_out.reserve((size_t)std::max(max_out, 1) * 4); d_total.reserve(4);
		BS_CUDA(cudaMemcpy2DAsync(d_img.p, (size_t)width * 4, intensity, (size_t)stride * 4, (size_t)width * 4, height, cudaMemcpyHostToDevice, ts.st));
		launch_nonmax_single(d_img.as<float>(), width, width, height, radius, threshold_max, ignore_border, d_cand.as<int>(), max_out, d_out.as<int16_t>(),
		                     d_total.as<int>(), ts.st);
		int total = 0;
		BS_CUDA(cudaMemcpyAsync(&total, d_total.p, 4, cudaMemcpyDeviceToHost, ts.st));
		BS_CUDA(cudaStreamSynchronize(ts.st));
		*n = total;
		if (std::min(total, max_out) > 0) BS_CUDA(cudaMemcpy(xy, d_out.p, (size_t)std::min(total, max_out) * 4, cudaMemcpyDeviceToHost));
		d_img.release(); d_cand.release(); d_out.release(); d_total.release();
		if (total > max_out) fail(BS_ERR_CAPACITY, "more maxima than max_out");
	});
}

int bs_hessian(int device, int ii_type, const void* ii, int width, int height, int skip, int size, float* out) {
	return guarded([&] {
		if (!out) fail(BS_ERR_ARG, "null argument");
		if (skip < 1 || size < 3) fail(BS_ERR_ARG, "bad skip/size");
		check_device(device);
		DeviceGuard g(device);
		TempStream ts;
		DevBuf d_ii, d_out;
		upload_ii(d_ii, ii, width, height);
		const size_t n = (size_t)(width / skip) * (height / skip);
		if (n == 0) { d_ii.release(); return; }
		d_out.reserve(n * 4);
		if (ii_type == BS_PIXEL_F32) launch_hessian_single<float>(d_ii.as<float>(), width, height, skip, size, d_out.as<float>(), ts.st);
		else launch_hessian_single<int32_t>(d_ii.as<int32_t>(), width, height, skip, size, d_out.as<float>(), ts.st);
		BS_CUDA(cudaStreamSynchronize(ts.st));
		BS_CUDA(cudaMemcpy(out, d_out.p, n * 4, cudaMemcpyDeviceToHost));
		d_ii.release(); d_out.release();
	});
}

int bs_fh_intensity(int device, int ii_type, const void* ii, int width, int height, const bs_fh_cfg* cfg, int* n_layers, long* layers,
                    long* n_floats, float* out, long max_floats) {
	return guarded([&] {
		bs_fh_cfg f;
		if (cfg) f = *cfg; else bs_fh_cfg_default(&f);
		check_fh_cfg(f);
		if (!n_layers || !n_floats) fail(BS_ERR_ARG, "null argument");
		if (width <= 0 || height <= 0) fail(BS_ERR_ARG, "image width and height must be positive");
		Plan plan = make_plan(f, width, height);
		*n_layers = plan.n_layers; *n_floats = plan.inten_per_frame;
		if (layers)
			for (int k = 0; k < plan.n_layers; k++) {
				const Layer& L = plan.layers[k];
				layers[5 * k] = L.size; layers[5 * k + 1] = L.skip; layers[5 * k + 2] = L.w; layers[5 * k + 3] = L.h; layers[5 * k + 4] = L.offset;
			}
		if (!out || plan.inten_per_frame == 0) return;
		if (max_floats < plan.inten_per_frame) fail(BS_ERR_CAPACITY, "output buffer smaller than the scale space");
		check_device(device);
		DeviceGuard g(device);
		TempStream ts;
		DevBuf d_ii, d_plan, d_inten;
		upload_ii(d_ii, ii, width, height);
		d_plan.reserve(sizeof(Plan));
		h2d_sync(d_plan.p, &plan, sizeof(Plan));
		d_inten.reserve((size_t)plan.inten_per_frame * 4);
		BS_CUDA(cudaDeviceSynchronize());
		if (ii_type == BS_PIXEL_F32) launch_hessian_plan<float>(d_ii.as<float>(), plan, d_plan.as<Plan>(), d_inten.as<float>(), 1, ts.st);
		else launch_hessian_plan<int32_t>(d_ii.as<int32_t>(), plan, d_plan.as<Plan>(), d_inten.as<float>(), 1, ts.st);
		BS_CUDA(cudaStreamSynchronize(ts.st));
		BS_CUDA(cudaMemcpy(out, d_inten.p, (size_t)plan.inten_per_frame * 4, cudaMemcpyDeviceToHost));
		d_ii.release(); d_plan.release(); d_inten.release();
	});
}

int bs_fh_detect(int device, int ii_type, const void* ii, int width, int height, const bs_fh_cfg* cfg, int max_out, int* n_out,
                 double* xys, int32_t* meta, float* response) {
	return guarded([&] {
		bs_fh_cfg f;
		if (cfg) f = *cfg; else bs_fh_cfg_default(&f);
		check_fh_cfg(f);
		if (max_out < 0) fail(BS_ERR_ARG, "max_out must be >= 0");
		check_device(device);
		DeviceGuard g(device);
		TempStream ts;
		DevBuf d_ii, d_plan, d_inten, d_stage, d_counts, d_offsets, d_fo, d_keys, d_keep;
		PoolBufs pool;
		upload_ii(d_ii, ii, width, height);
		Plan plan = make_plan(f, width, height);
		d_plan.reserve(sizeof(Plan));
		h2d_sync(d_plan.p, &plan, sizeof(Plan));
		d_inten.reserve((size_t)std::max<long>(plan.inten_per_frame, 1) * 4);
		const size_t nctas = std::max(plan.ctas_per_frame, 1);
		d_stage.reserve(nctas * NMS_CTA_BLOCKS * sizeof(RawPoint));
		d_counts.reserve(nctas * 4); d_offsets.reserve((nctas + 1) * 4); d_fo.reserve(8);
		if (plan.max_per_level > 0) { d_keys.reserve(nctas * NMS_CTA_BLOCKS * 4); d_keep.reserve(nctas * NMS_CTA_BLOCKS); }
		BS_CUDA(cudaMemset(d_fo.p, 0, 8));
		BS_CUDA(cudaDeviceSynchronize());
		pool.reserve(std::max(max_out, 1), 1);
		if (ii_type == BS_PIXEL_F32) launch_hessian_plan<float>(d_ii.as<float>(), plan, d_plan.as<Plan>(), d_inten.as<float>(), 1, ts.st);
		else launch_hessian_plan<int32_t>(d_ii.as<int32_t>(), plan, d_plan.as<Plan>(), d_inten.as<float>(), 1, ts.st);
		launch_nms(plan, d_plan.as<Plan>(), d_inten.as<float>(), 1, d_stage.as<RawPoint>(), d_counts.as<int>(), d_keys.as<unsigned>(),
		           d_keep.as<unsigned char>(), ts.st);
		launch_compact(plan, 1, 0, d_stage.as<RawPoint>(), d_counts.as<int>(), d_offsets.as<int>(), d_fo.as<int>(), pool.view, ts.st);
		BS_CUDA(cudaStreamSynchronize(ts.st));
		int fo[2];
		BS_CUDA(cudaMemcpy(fo, d_fo.p, 8, cudaMemcpyDeviceToHost));
		if (n_out) *n_out = fo[1];
		const int k = std::min(fo[1], max_out);
		if (k > 0) {
			if (xys) {
				std::vector<double> xy((size_t)2 * k), sc(k);
				BS_CUDA(cudaMemcpy(xy.data(), pool.view.xy, xy.size() * 8, cudaMemcpyDeviceToHost));
				BS_CUDA(cudaMemcpy(sc.data(), pool.view.scale, sc.size() * 8, cudaMemcpyDeviceToHost));
				for (int i = 0; i < k; i++) { xys[3 * i] = xy[2 * i]; xys[3 * i + 1] = xy[2 * i + 1]; xys[3 * i + 2] = sc[i]; }
			}
			if (meta) BS_CUDA(cudaMemcpy(meta, pool.view.meta, (size_t)k * 16, cudaMemcpyDeviceToHost));
			if (response) BS_CUDA(cudaMemcpy(response, pool.view.response, (size_t)k * 4, cudaMemcpyDeviceToHost));
		}
		d_ii.release(); d_plan.release(); d_inten.release(); d_stage.release(); d_counts.release(); d_offsets.release(); d_fo.release();
		d_keys.release(); d_keep.release();
		pool.release();
	});
}

static void run_describe_stage(int device, int ii_type, const void* ii, int width, int height, const bs_ori_cfg& o, const bs_desc_cfg& d,
                               const double* xys, const double* angle_in, int n, bool do_ori, bool do_desc, double* angle_out,
                               double* desc_out, uint8_t* white_out) {
	if (n < 0) fail(BS_ERR_ARG, "n must be >= 0");
	if (n == 0) return;
	if (!xys) fail(BS_ERR_ARG, "null argument");
	check_device(device);
	DeviceGuard g(device);
	TempStream ts;
	DevBuf d_ii;
	upload_ii(d_ii, ii, width, height);
	DescribeSetup ds;
	ds.build(o, d);
	ds.upload();
	PointSet ps;
	ps.load(xys, angle_in, n, ds.p.dof);
	const int* r = ps.range.as<int>();
	if (ii_type == BS_PIXEL_F32) launch_describe<float>(d_ii.as<float>(), width, height, ds.p, ds.tables, ps.pool.view, r, r + 1, 0, do_ori, do_desc, ts.st);
	else launch_describe<int32_t>(d_ii.as<int32_t>(), width, height, ds.p, ds.tables, ps.pool.view, r, r + 1, 0, do_ori, do_desc, ts.st);
	BS_CUDA(cudaStreamSynchronize(ts.st));
	if (angle_out) BS_CUDA(cudaMemcpy(angle_out, ps.pool.view.angle, (size_t)n * 8, cudaMemcpyDeviceToHost));
	if (desc_out) BS_CUDA(cudaMemcpy(desc_out, ps.pool.view.desc, (size_t)n * ds.p.dof * 8, cudaMemcpyDeviceToHost));
	if (white_out) BS_CUDA(cudaMemcpy(white_out, ps.pool.view.white, (size_t)n, cudaMemcpyDeviceToHost));
	ps.release(); ds.release(); d_ii.release();
}

int bs_orientation(int device, int ii_type, const void* ii, int width, int height, const bs_ori_cfg* cfg, const double* xys, int n,
                   double* angle) {
	return guarded([&] {
		bs_ori_cfg o; bs_desc_cfg d;
		if (cfg) o = *cfg; else bs_ori_cfg_default(&o, 0);
		bs_desc_cfg_default(&d, 0);
		check_ori_cfg(o);
		run_describe_stage(device, ii_type, ii, width, height, o, d, xys, nullptr, n, true, false, angle, nullptr, nullptr);
	});
}

int bs_describe(int device, int ii_type, const void* ii, int width, int height, const bs_desc_cfg* cfg, const double* xys,
                const double* angle, int n, double* desc, uint8_t* white) {
	return guarded([&] {
		bs_ori_cfg o; bs_desc_cfg d;
		if (cfg) d = *cfg; else bs_desc_cfg_default(&d, 0);
		bs_ori_cfg_default(&o, 0);
		check_desc_cfg(d);
		if (!angle) fail(BS_ERR_ARG, "null argument");
		run_describe_stage(device, ii_type, ii, width, height, o, d, xys, angle, n, false, true, nullptr, desc, white);
	});
}

}  // extern "C"

// ====================================================================================================== bs_assoc
struct bs_assoc {
	int device = 0;
	double max_fit = DBL_MAX;
	int backwards = 0;
	int mode = -1;            // -1 auto, 0 exact, 1 tensor-core filter + exact re-score
	int sm_count = 148;
	cudaStream_t stream = nullptr;
	DevBuf p_min, p_first, p_last;            // split partials
	DevBuf col_min, col_first, col_last;      // column statistics
	DevBuf d_src, d_dst, d_pairs, d_fit;      // staging for the host entry point
	// tensor-core filter workspace: split operands, norms, candidate lists
	DevBuf op_src, op_dst, norm_src, norm_dst, tc_scalars /* maxnorm src, maxnorm dst, overflow */, cand_count, cand;
	long tc_runs = 0, tc_fallbacks = 0;
	// AssociateSurfBasic workspace: class index lists, class matrices, class results, staged Laplacian signs
	DevBuf sb_idx_src, sb_idx_dst, sb_counts, sb_src, sb_dst, sb_pairs, sb_fit, sb_white_src, sb_white_dst;

	// AssociateSurfBasic.associate (F/alg/feature/associate/AssociateSurfBasic.java:83-122) on device arrays: pairs[n] = index
	// into dst (global) or -1, fit[n].  Synchronises `st` (the class sizes decide the launches).
	void run_surf_basic(const double* src, const uint8_t* white_src, int n, const double* dst, const uint8_t* white_dst, int m, int dof,
	                    int32_t* pairs, double* fit, cudaStream_t st) {
		if (n < 0 || m < 0 || dof < 1) fail(BS_ERR_ARG, "bad sizes");
		if (n == 0) return;
		sb_idx_src.reserve((size_t)n * 4); sb_idx_dst.reserve((size_t)std::max(m, 1) * 4); sb_counts.reserve(8);
		sb_src.reserve((size_t)n * dof * 8); sb_dst.reserve((size_t)std::max(m, 1) * dof * 8);
		sb_pairs.reserve((size_t)n * 4); sb_fit.reserve((size_t)n * 8);
		int32_t* cnt = sb_counts.as<int32_t>();
		launch_partition_white(white_src, n, sb_idx_src.as<int32_t>(), cnt + 0, st);
		launch_partition_white(white_dst, m, sb_idx_dst.as<int32_t>(), cnt + 1, st);
		launch_gather_rows(src, sb_idx_src.as<int32_t>(), n, dof, sb_src.as<double>(), st);
		launch_gather_rows(dst, sb_idx_dst.as<int32_t>(), m, dof, sb_dst.as<double>(), st);
		int32_t h_cnt[2] = {0, 0};
		BS_CUDA(cudaMemcpyAsync(h_cnt, cnt, 8, cudaMemcpyDeviceToHost, st));
		BS_CUDA(cudaStreamSynchronize(st));
		const int ns[2] = {h_cnt[0], n - h_cnt[0]}, ms[2] = {h_cnt[1], m - h_cnt[1]};
		const int so[2] = {0, h_cnt[0]}, mo[2] = {0, h_cnt[1]};
		for (int c = 0; c < 2; c++) {            // positive (white) lists first, then negative (:94-121)
			if (ns[c] == 0) continue;
			run(sb_src.as<double>() + (size_t)so[c] * dof, ns[c], sb_dst.as<double>() + (size_t)mo[c] * dof, ms[c], dof,
			    sb_pairs.as<int32_t>() + so[c], sb_fit.as<double>() + so[c], st);
			launch_scatter_pairs(sb_pairs.as<int32_t>() + so[c], sb_fit.as<double>() + so[c], ns[c], sb_idx_src.as<int32_t>() + so[c],
			                     sb_idx_dst.as<int32_t>() + mo[c], pairs, fit, st);
		}
	}

	RowminScratch scratch(long entries) {
		p_min.reserve(entries * 8); p_first.reserve(entries * 4); p_last.reserve(entries * 4);
		return RowminScratch{p_min.as<double>(), p_first.as<int>(), p_last.as<int>(), entries};
	}
	// forward pass of the local rows: pairs/fit
	void forward(const double* src, int n, const double* dst, int m, int dof, int32_t* pairs, double* fit, cudaStream_t st) {
		const int splits = rowmin_splits(n, m, sm_count);
		RowminScratch s = scratch((long)splits * n);
		launch_rowmin(src, n, dst, m, dof, splits, s, st);
		launch_rowmin_finalize(s, n, splits, 0, max_fit, 0, pairs, fit, nullptr, nullptr, nullptr, st);
	}
	// column statistics over the local rows (global numbering = row_offset + local)
	void columns(const double* src, int n, int row_offset, const double* dst, int m, int dof, double* cmin, int32_t* cfirst,
	             int32_t* clast, cudaStream_t st) {
		const int splits = rowmin_splits(m, n, sm_count);
		RowminScratch s = scratch((long)splits * m);
		launch_rowmin(dst, m, src, n, dof, splits, s, st);
		launch_rowmin_finalize(s, m, splits, 1, 0.0, row_offset, nullptr, nullptr, cmin, cfirst, clast, st);
	}
	void run(const double* src, int n, const double* dst, int m, int dof, int32_t* pairs, double* fit, cudaStream_t st) {
		if (n < 0 || m < 0 || dof < 1) fail(BS_ERR_ARG, "bad sizes");
		if (n == 0) return;
		if (m == 0) {   // no destination: every source is unmatched with the untouched running best
			std::vector<int32_t> hp(n, -1); std::vector<double> hf(n, max_fit);
			BS_CUDA(cudaMemcpyAsync(pairs, hp.data(), (size_t)n * 4, cudaMemcpyHostToDevice, st));
			BS_CUDA(cudaMemcpyAsync(fit, hf.data(), (size_t)n * 8, cudaMemcpyHostToDevice, st));
			BS_CUDA(cudaStreamSynchronize(st));
			return;
		}
		const bool want_tc = mode == 1 || (mode == -1 && tc_supported(dof) && (long)n * m >= (1L << 22));
		if (want_tc) {
			if (!tc_supported(dof)) fail(BS_ERR_ARG, "the tensor-core filter is built for 64-element descriptors (SURF-64)");
			if (run_tc(src, n, dst, m, pairs, fit, st)) return;
			tc_fallbacks++;                    // a row had more candidates than the list holds: redo everything exactly
		}
		run_exact(src, n, dst, m, dof, pairs, fit, st);
	}
	void run_exact(const double* src, int n, const double* dst, int m, int dof, int32_t* pairs, double* fit, cudaStream_t st) {
		forward(src, n, dst, m, dof, pairs, fit, st);
		if (backwards) {
			col_min.reserve((size_t)m * 8); col_first.reserve((size_t)m * 4); col_last.reserve((size_t)m * 4);
			columns(src, n, 0, dst, m, dof, col_min.as<double>(), col_first.as<int32_t>(), col_last.as<int32_t>(), st);
			launch_validate(n, 0, pairs, fit, col_min.as<double>(), col_first.as<int32_t>(), col_last.as<int32_t>(), st);
		}
	}
	void tc_prepare(const double* src, int n, const double* dst, int m, cudaStream_t st) {
		op_src.reserve(tc_operand_bytes(n)); op_dst.reserve(tc_operand_bytes(m));
		norm_src.reserve(tc_norm_bytes(n)); norm_dst.reserve(tc_norm_bytes(m));
		tc_scalars.reserve(16);
		const size_t rows = (size_t)std::max(n, m);
		cand_count.reserve(rows * 4); cand.reserve(tc_cand_bytes(rows));
		unsigned* sc = tc_scalars.as<unsigned>();
		launch_tc_split(src, n, op_src.p, norm_src.as<float>(), sc + 0, st);
		launch_tc_split(dst, m, op_dst.p, norm_dst.as<float>(), sc + 1, st);
		BS_CUDA(cudaMemsetAsync(sc + 2, 0, 4, st));
	}
	// sharded step 1 on the tensor cores: forward pass of the local rows and the column statistics over them (global row
	// numbering); false if a candidate list overflowed.  Synchronises `st`.
	bool forward_tc(const double* src, int n, int row_offset, const double* dst, int m, int32_t* pairs, double* fit, double* cmin,
	                int32_t* cfirst, int32_t* clast, cudaStream_t st) {
		tc_prepare(src, n, dst, m, st);
		unsigned* sc = tc_scalars.as<unsigned>();
		int* overflow = reinterpret_cast<int*>(sc + 2);
		launch_tc_filter(op_src.p, norm_src.as<float>(), n, op_dst.p, norm_dst.as<float>(), sc + 1, m, cand_count.as<int>(), cand.as<int>(),
		                 nullptr, sm_count, st);
		launch_tc_rescore(src, n, dst, cand_count.as<int>(), cand.as<int>(), norm_src.as<float>(), sc + 1, 0, max_fit, 0, pairs, fit, nullptr, nullptr, nullptr, overflow, st);
		tc_bound.reserve((size_t)m * 8);
		launch_tc_row_bounds(pairs, fit, n, m, tc_bound.as<double>(), st);
		launch_tc_filter(op_dst.p, norm_dst.as<float>(), m, op_src.p, norm_src.as<float>(), sc + 0, n, cand_count.as<int>(), cand.as<int>(),
		                 nullptr, sm_count, st, tc_bound.as<double>());
		launch_tc_rescore(dst, m, src, cand_count.as<int>(), cand.as<int>(), norm_dst.as<float>(), sc + 0, 1, 0.0, row_offset, nullptr, nullptr, cmin, cfirst, clast, overflow, st);
		int h_overflow = 0;
		BS_CUDA(cudaMemcpyAsync(&h_overflow, overflow, 4, cudaMemcpyDeviceToHost, st));
		BS_CUDA(cudaStreamSynchronize(st));
		tc_runs++;
		return h_overflow == 0;
	}
	// workspace of the tensor-core path for problems up to n x m (so that queued runs never reallocate under a running kernel)
	void tc_reserve(int n, int m) {
		op_src.reserve(tc_operand_bytes(n)); op_dst.reserve(tc_operand_bytes(m));
		norm_src.reserve(tc_norm_bytes(n)); norm_dst.reserve(tc_norm_bytes(m));
		tc_scalars.reserve(16);
		const size_t rows = (size_t)std::max(n, m);
		cand_count.reserve(rows * 4); cand.reserve(tc_cand_bytes(rows));
		col_min.reserve((size_t)m * 8); col_first.reserve((size_t)m * 4); col_last.reserve((size_t)m * 4);
		tc_bound.reserve((size_t)m * 8);
	}
	// tensor-core filter + exact re-score, queued on `st` without a host round trip; *overflow (device, zeroed by the caller)
	// becomes non-zero if a candidate list overflowed (results then invalid)
	void queue_tc(const double* src, int n, const double* dst, int m, int32_t* pairs, double* fit, int* overflow, cudaStream_t st) {
		tc_prepare(src, n, dst, m, st);
		unsigned* sc = tc_scalars.as<unsigned>();
		launch_tc_filter(op_src.p, norm_src.as<float>(), n, op_dst.p, norm_dst.as<float>(), sc + 1, m, cand_count.as<int>(), cand.as<int>(),
		                 nullptr, sm_count, st);
		launch_tc_rescore(src, n, dst, cand_count.as<int>(), cand.as<int>(), norm_src.as<float>(), sc + 1, 0, max_fit, 0, pairs, fit, nullptr, nullptr, nullptr, overflow, st);
		if (backwards) {
			col_min.reserve((size_t)m * 8); col_first.reserve((size_t)m * 4); col_last.reserve((size_t)m * 4);
			// the column minimum of j is at most the fit of any row that chose j: the second direction starts from it
			tc_bound.reserve((size_t)m * 8);
			launch_tc_row_bounds(pairs, fit, n, m, tc_bound.as<double>(), st);
			launch_tc_filter(op_dst.p, norm_dst.as<float>(), m, op_src.p, norm_src.as<float>(), sc + 0, n, cand_count.as<int>(), cand.as<int>(),
			                 nullptr, sm_count, st, tc_bound.as<double>());
			launch_tc_rescore(dst, m, src, cand_count.as<int>(), cand.as<int>(), norm_dst.as<float>(), sc + 0, 1, 0.0, 0, nullptr, nullptr, col_min.as<double>(),
			                  col_first.as<int32_t>(), col_last.as<int32_t>(), overflow, st);
			launch_validate(n, 0, pairs, fit, col_min.as<double>(), col_first.as<int32_t>(), col_last.as<int32_t>(), st);
		}
		tc_runs++;
	}
	// the same, synchronising `st`; false if a candidate list overflowed
	bool run_tc(const double* src, int n, const double* dst, int m, int32_t* pairs, double* fit, cudaStream_t st) {
		tc_scalars.reserve(16);
		int* overflow = reinterpret_cast<int*>(tc_scalars.as<unsigned>() + 2);    // zeroed by tc_prepare
		queue_tc(src, n, dst, m, pairs, fit, overflow, st);
		int h_overflow = 0;
		BS_CUDA(cudaMemcpyAsync(&h_overflow, overflow, 4, cudaMemcpyDeviceToHost, st));
		BS_CUDA(cudaStreamSynchronize(st));
		return h_overflow == 0;
	}
	DevBuf batch_flags, tc_bound;
};

extern "C" {

int bs_assoc_create(int device, double max_fit_error, int backwards_validation, bs_assoc** out) {
	return guarded([&] {
		if (!out) fail(BS_ERR_ARG, "out is null");
		*out = nullptr;
		check_device(device);
		DeviceGuard g(device);
		std::unique_ptr<bs_assoc> a(new bs_assoc());
		a->device = device; a->max_fit = max_fit_error; a->backwards = backwards_validation ? 1 : 0;
		BS_CUDA(cudaDeviceGetAttribute(&a->sm_count, cudaDevAttrMultiProcessorCount, device));
		BS_CUDA(cudaStreamCreateWithFlags(&a->stream, cudaStreamNonBlocking));
		*out = a.release();
	});
}

int bs_assoc_destroy(bs_assoc* a) {
	if (!a) return BS_OK;
	return guarded([&] {
		DeviceGuard g(a->device);
		cudaDeviceSynchronize();
		a->p_min.release(); a->p_first.release(); a->p_last.release();
		a->col_min.release(); a->col_first.release(); a->col_last.release();
		a->d_src.release(); a->d_dst.release(); a->d_pairs.release(); a->d_fit.release();
		a->op_src.release(); a->op_dst.release(); a->norm_src.release(); a->norm_dst.release(); a->tc_scalars.release();
		a->cand_count.release(); a->cand.release(); a->batch_flags.release(); a->tc_bound.release();
		a->sb_idx_src.release(); a->sb_idx_dst.release(); a->sb_counts.release(); a->sb_src.release(); a->sb_dst.release();
		a->sb_pairs.release(); a->sb_fit.release(); a->sb_white_src.release(); a->sb_white_dst.release();
		if (a->stream) cudaStreamDestroy(a->stream);
		delete a;
	});
}

int bs_assoc_set_max_fit_error(bs_assoc* a, double max_fit_error) {
	return guarded([&] { if (!a) fail(BS_ERR_ARG, "handle is null"); a->max_fit = max_fit_error; });
}

int bs_assoc_set_mode(bs_assoc* a, int mode) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		if (mode < -1 || mode > 1) fail(BS_ERR_ARG, "mode must be -1 (auto), 0 (exact) or 1 (tensor-core filter)");
		a->mode = mode;
	});
}

int bs_assoc_filter_scores(bs_assoc* a, const double* src, int n, const double* dst, int m, int dof, float* scores,
                           int32_t* cand_count, long* stats) {
	return guarded([&] {
		if (!a || !src || !dst) fail(BS_ERR_ARG, "null argument");
		if (n <= 0 || m <= 0) fail(BS_ERR_ARG, "bad sizes");
		if (!tc_supported(dof)) fail(BS_ERR_ARG, "the tensor-core filter is built for 64-element descriptors (SURF-64)");
		DeviceGuard g(a->device);
		cudaStream_t st = a->stream;
		DevBuf d_scores;
		a->d_src.reserve((size_t)n * dof * 8); a->d_dst.reserve((size_t)m * dof * 8);
		BS_CUDA(cudaMemcpyAsync(a->d_src.p, src, (size_t)n * dof * 8, cudaMemcpyHostToDevice, st));
		BS_CUDA(cudaMemcpyAsync(a->d_dst.p, dst, (size_t)m * dof * 8, cudaMemcpyHostToDevice, st));
		if (scores) { d_scores.reserve((size_t)n * m * 4); BS_CUDA(cudaMemsetAsync(d_scores.p, 0xff, (size_t)n * m * 4, st)); }
		a->tc_prepare(a->d_src.as<double>(), n, a->d_dst.as<double>(), m, st);
		unsigned* sc = a->tc_scalars.as<unsigned>();
		launch_tc_filter(a->op_src.p, a->norm_src.as<float>(), n, a->op_dst.p, a->norm_dst.as<float>(), sc + 1, m, a->cand_count.as<int>(),
		                 a->cand.as<int>(), scores ? d_scores.as<float>() : nullptr, a->sm_count, st);
		if (scores) BS_CUDA(cudaMemcpyAsync(scores, d_scores.p, (size_t)n * m * 4, cudaMemcpyDeviceToHost, st));
		if (cand_count) BS_CUDA(cudaMemcpyAsync(cand_count, a->cand_count.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
		BS_CUDA(cudaStreamSynchronize(st));
		d_scores.release();
		if (stats) { stats[0] = a->tc_runs; stats[1] = a->tc_fallbacks; }
	});
}

int bs_assoc_run_device(bs_assoc* a, const double* d_src, int n, const double* d_dst, int m, int dof, int32_t* d_pairs, double* d_fit,
                        void* stream) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		if ((n > 0 && (!d_src || !d_pairs || !d_fit)) || (m > 0 && !d_dst)) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(a->device);
		a->run(d_src, n, d_dst, m, dof, d_pairs, d_fit, stream ? static_cast<cudaStream_t>(stream) : a->stream);
	});
}

int bs_assoc_run(bs_assoc* a, const double* src, int n, const double* dst, int m, int dof, int32_t* pairs, double* fit) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		if (!src && n > 0) fail(BS_ERR_STATE, "source features not specified");
		if (!dst && m > 0) fail(BS_ERR_STATE, "destination features not specified");
		if (n < 0 || m < 0 || dof < 1) fail(BS_ERR_ARG, "bad sizes");
		if (n == 0) return;
		if (!pairs || !fit) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(a->device);
		a->d_src.reserve((size_t)n * dof * 8); a->d_dst.reserve((size_t)std::max(m, 1) * dof * 8);
		a->d_pairs.reserve((size_t)n * 4); a->d_fit.reserve((size_t)n * 8);
		BS_CUDA(cudaMemcpyAsync(a->d_src.p, src, (size_t)n * dof * 8, cudaMemcpyHostToDevice, a->stream));
		if (m > 0) BS_CUDA(cudaMemcpyAsync(a->d_dst.p, dst, (size_t)m * dof * 8, cudaMemcpyHostToDevice, a->stream));
		a->run(a->d_src.as<double>(), n, a->d_dst.as<double>(), m, dof, a->d_pairs.as<int32_t>(), a->d_fit.as<double>(), a->stream);
		BS_CUDA(cudaMemcpyAsync(pairs, a->d_pairs.p, (size_t)n * 4, cudaMemcpyDeviceToHost, a->stream));
		BS_CUDA(cudaMemcpyAsync(fit, a->d_fit.p, (size_t)n * 8, cudaMemcpyDeviceToHost, a->stream));
		BS_CUDA(cudaStreamSynchronize(a->stream));
	});
}

int bs_assoc_surf_run_device(bs_assoc* a, const double* d_src, const uint8_t* d_src_white, int n, const double* d_dst,
                             const uint8_t* d_dst_white, int m, int dof, int32_t* d_pairs, double* d_fit, void* stream) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		if ((n > 0 && (!d_src || !d_src_white || !d_pairs || !d_fit)) || (m > 0 && (!d_dst || !d_dst_white))) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(a->device);
		a->run_surf_basic(d_src, d_src_white, n, d_dst, d_dst_white, m, dof, d_pairs, d_fit,
		                  stream ? static_cast<cudaStream_t>(stream) : a->stream);
	});
}

int bs_assoc_surf_run(bs_assoc* a, const double* src, const uint8_t* src_white, int n, const double* dst, const uint8_t* dst_white, int m,
                      int dof, int32_t* pairs, double* fit) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		if ((!src || !src_white) && n > 0) fail(BS_ERR_STATE, "source features not specified");
		if ((!dst || !dst_white) && m > 0) fail(BS_ERR_STATE, "destination features not specified");
		if (n < 0 || m < 0 || dof < 1) fail(BS_ERR_ARG, "bad sizes");
		if (n == 0) return;
		if (!pairs || !fit) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(a->device);
		cudaStream_t st = a->stream;
		a->d_src.reserve((size_t)n * dof * 8); a->d_dst.reserve((size_t)std::max(m, 1) * dof * 8);
		a->d_pairs.reserve((size_t)n * 4); a->d_fit.reserve((size_t)n * 8);
		a->sb_white_src.reserve((size_t)n); a->sb_white_dst.reserve((size_t)std::max(m, 1));
		BS_CUDA(cudaMemcpyAsync(a->d_src.p, src, (size_t)n * dof * 8, cudaMemcpyHostToDevice, st));
		BS_CUDA(cudaMemcpyAsync(a->sb_white_src.p, src_white, (size_t)n, cudaMemcpyHostToDevice, st));
		if (m > 0) {
			BS_CUDA(cudaMemcpyAsync(a->d_dst.p, dst, (size_t)m * dof * 8, cudaMemcpyHostToDevice, st));
			BS_CUDA(cudaMemcpyAsync(a->sb_white_dst.p, dst_white, (size_t)m, cudaMemcpyHostToDevice, st));
		}
		a->run_surf_basic(a->d_src.as<double>(), a->sb_white_src.as<uint8_t>(), n, a->d_dst.as<double>(), a->sb_white_dst.as<uint8_t>(), m,
		                  dof, a->d_pairs.as<int32_t>(), a->d_fit.as<double>(), st);
		BS_CUDA(cudaMemcpyAsync(pairs, a->d_pairs.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
		BS_CUDA(cudaMemcpyAsync(fit, a->d_fit.p, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
		BS_CUDA(cudaStreamSynchronize(st));
	});
}

// AssociateSurfBasic.associate :83-122 — match list and unassociated sources in the reference's order: the white class in
// ascending source index, then the other class; nothing at all when either list is empty (:89-92)
int bs_assoc_surf_matches(const int32_t* pairs, const double* fit, const uint8_t* src_white, int n, int m, int32_t* match_src,
                          int32_t* match_dst, double* match_score, int* n_matches, int32_t* unassoc_src, int* n_unassoc_src,
                          int32_t* unassoc_dst, int* n_unassoc_dst) {
	return guarded([&] {
		if (n < 0 || m < 0 || (n > 0 && (!pairs || !fit || !src_white))) fail(BS_ERR_ARG, "bad arguments");
		std::vector<char> hit((size_t)std::max(m, 1), 0);
		int nm = 0, nus = 0, nud = 0;
		if (n > 0 && m > 0) {
			for (int cls = 1; cls >= 0; cls--)
				for (int i = 0; i < n; i++) {
					if ((src_white[i] != 0) != (cls == 1)) continue;
					const int d = pairs[i];
					if (d >= 0) {
						if (d >= m) fail(BS_ERR_ARG, "pair index out of range");
						if (match_src) match_src[nm] = i;
						if (match_dst) match_dst[nm] = d;
						if (match_score) match_score[nm] = fit[i];
						nm++; hit[d] = 1;
					}
				}
			// the unassociated sources of a class are listed after that class's matches, class by class (:104-107,118-121)
			for (int cls = 1; cls >= 0; cls--)
				for (int i = 0; i < n; i++)
					if ((src_white[i] != 0) == (cls == 1) && pairs[i] < 0) { if (unassoc_src) unassoc_src[nus] = i; nus++; }
		}
		// WrapAssociateSurfBasic.getUnassociatedDestination: FindUnassociated.checkDestination over totalDestination()
		for (int j = 0; j < m; j++)
			if (!hit[j]) { if (unassoc_dst) unassoc_dst[nud] = j; nud++; }
		if (n_matches) *n_matches = nm;
		if (n_unassoc_src) *n_unassoc_src = nus;
		if (n_unassoc_dst) *n_unassoc_dst = nud;
	});
}

// detect -> associate without leaving the GPU: the features of two frames of the detector's last batch are the source and
// destination lists (what DetectDescribeAssociate / ExampleAssociatePoints do through host lists in the reference)
int bs_assoc_frames_device(bs_assoc* a, bs_surf* s, int frame_src, int frame_dst, int surf_basic, int32_t* d_pairs, double* d_fit,
                           void* stream) {
	return guarded([&] {
		if (!a || !s) fail(BS_ERR_ARG, "handle is null");
		if (a->device != s->device) fail(BS_ERR_ARG, "detector and associator live on different devices");
		DeviceGuard g(a->device);
		check_ready(s);
		if (frame_src < 0 || frame_src >= s->batch || frame_dst < 0 || frame_dst >= s->batch) fail(BS_ERR_ARG, "frame out of range");
		const std::vector<int>& fo = s->h_frame_offset;
		if (fo[s->batch] > s->pool.view.cap) fail(BS_ERR_CAPACITY, "feature pool too small for this batch");
		const int n = fo[frame_src + 1] - fo[frame_src], m = fo[frame_dst + 1] - fo[frame_dst];
		if (n > 0 && (!d_pairs || !d_fit)) fail(BS_ERR_ARG, "null argument");
		const Pool& p = s->pool.view;
		const int dof = s->out_dof();
		cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : a->stream;
		const double* src = p.desc + (long)fo[frame_src] * dof;
		const double* dst = p.desc + (long)fo[frame_dst] * dof;
		if (surf_basic) a->run_surf_basic(src, p.white + fo[frame_src], n, dst, p.white + fo[frame_dst], m, dof, d_pairs, d_fit, st);
		else a->run(src, n, dst, m, dof, d_pairs, d_fit, st);
	});
}

int bs_assoc_features_device(bs_assoc* a, const bs_features* src, const bs_features* dst, int surf_basic, int32_t* d_pairs, double* d_fit,
                             void* stream) {
	return guarded([&] {
		if (!a || !src || !dst) fail(BS_ERR_ARG, "handle is null");
		if (a->device != src->device || a->device != dst->device) fail(BS_ERR_ARG, "feature sets and associator live on different devices");
		if (src->dof != dst->dof) fail(BS_ERR_ARG, "source and destination descriptors differ in length");
		if (src->n > 0 && (!d_pairs || !d_fit)) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(a->device);
		cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : a->stream;
		if (surf_basic) a->run_surf_basic(src->desc.as<double>(), src->white.as<uint8_t>(), src->n, dst->desc.as<double>(), dst->white.as<uint8_t>(),
		                                  dst->n, src->dof, d_pairs, d_fit, st);
		else a->run(src->desc.as<double>(), src->n, dst->desc.as<double>(), dst->n, src->dof, d_pairs, d_fit, st);
	});
}

int bs_assoc_frames_batch_device(bs_assoc* a, bs_surf* s, int n_pairs, const int32_t* frame_src, const int32_t* frame_dst,
                                 int32_t* d_pairs, double* d_fit, long cap, long* offsets, void* stream) {
	return guarded([&] {
		if (!a || !s) fail(BS_ERR_ARG, "handle is null");
		if (a->device != s->device) fail(BS_ERR_ARG, "detector and associator live on different devices");
		if (n_pairs < 0 || (n_pairs > 0 && (!frame_src || !frame_dst)) || !offsets) fail(BS_ERR_ARG, "bad pair list");
		DeviceGuard g(a->device);
		check_ready(s);
		const std::vector<int>& fo = s->h_frame_offset;
		if (fo[s->batch] > s->pool.view.cap) fail(BS_ERR_CAPACITY, "feature pool too small for this batch");
		int max_n = 0, max_m = 0;
		offsets[0] = 0;
		for (int k = 0; k < n_pairs; k++) {
			if (frame_src[k] < 0 || frame_src[k] >= s->batch || frame_dst[k] < 0 || frame_dst[k] >= s->batch) fail(BS_ERR_ARG, "frame out of range");
			const int n = fo[frame_src[k] + 1] - fo[frame_src[k]], m = fo[frame_dst[k] + 1] - fo[frame_dst[k]];
			offsets[k + 1] = offsets[k] + n;
			max_n = std::max(max_n, n); max_m = std::max(max_m, m);
		}
		if (offsets[n_pairs] > cap) fail(BS_ERR_CAPACITY, "pairs / fit arrays smaller than the source features of the batch");
		if (offsets[n_pairs] == 0) return;
		if (!d_pairs || !d_fit) fail(BS_ERR_ARG, "null argument");
		const Pool& p = s->pool.view;
		const int dof = s->out_dof();
		cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : a->stream;
		a->batch_flags.reserve((size_t)n_pairs * 4);
		BS_CUDA(cudaMemsetAsync(a->batch_flags.p, 0, (size_t)n_pairs * 4, st));
		if (tc_supported(dof)) a->tc_reserve(max_n, std::max(max_m, 1));
		std::vector<char> queued(n_pairs, 0);
		for (int k = 0; k < n_pairs; k++) {
			const int n = fo[frame_src[k] + 1] - fo[frame_src[k]], m = fo[frame_dst[k] + 1] - fo[frame_dst[k]];
			if (n == 0) continue;
			const double* src = p.desc + (long)fo[frame_src[k]] * dof;
			const double* dst = p.desc + (long)fo[frame_dst[k]] * dof;
			const bool want_tc = m > 0 && tc_supported(dof) && (a->mode == 1 || (a->mode == -1 && (long)n * m >= (1L << 22)));
			if (want_tc) {
				a->queue_tc(src, n, dst, m, d_pairs + offsets[k], d_fit + offsets[k], a->batch_flags.as<int>() + k, st);
				queued[k] = 1;
			} else {
				a->run(src, n, dst, m, dof, d_pairs + offsets[k], d_fit + offsets[k], st);
			}
		}
		std::vector<int> flags(n_pairs, 0);
		BS_CUDA(cudaMemcpyAsync(flags.data(), a->batch_flags.p, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, st));
		BS_CUDA(cudaStreamSynchronize(st));
		for (int k = 0; k < n_pairs; k++)
			if (queued[k] && flags[k] != 0) {        // a row had more candidates than the list holds: redo this pair exactly
				const int n = fo[frame_src[k] + 1] - fo[frame_src[k]], m = fo[frame_dst[k] + 1] - fo[frame_dst[k]];
				a->tc_fallbacks++;
				a->run_exact(p.desc + (long)fo[frame_src[k]] * dof, n, p.desc + (long)fo[frame_dst[k]] * dof, m, dof, d_pairs + offsets[k],
				             d_fit + offsets[k], st);
			}
	});
}

int bs_assoc_forward_device(bs_assoc* a, const double* d_src, int n, int row_offset, const double* d_dst, int m, int dof,
                            int32_t* d_pairs, double* d_fit, double* d_col_min, int32_t* d_col_first, int32_t* d_col_last,
                            void* stream) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		if (n <= 0 || m <= 0 || dof < 1) fail(BS_ERR_ARG, "bad sizes");
		if (!d_src || !d_dst || !d_pairs || !d_fit || !d_col_min || !d_col_first || !d_col_last) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(a->device);
		cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : a->stream;
		const bool want_tc = a->mode == 1 || (a->mode == -1 && tc_supported(dof) && (long)n * m >= (1L << 22));
		if (want_tc) {
			if (!tc_supported(dof)) fail(BS_ERR_ARG, "the tensor-core filter is built for 64-element descriptors (SURF-64)");
			if (a->forward_tc(d_src, n, row_offset, d_dst, m, d_pairs, d_fit, d_col_min, d_col_first, d_col_last, st)) return;
			a->tc_fallbacks++;
		}
		a->forward(d_src, n, d_dst, m, dof, d_pairs, d_fit, st);
		a->columns(d_src, n, row_offset, d_dst, m, dof, d_col_min, d_col_first, d_col_last, st);
	});
}

int bs_assoc_forward_only_device(bs_assoc* a, const double* d_src, int n, const double* d_dst, int m, int dof, int32_t* d_pairs,
                                 double* d_fit, void* stream) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		if ((n > 0 && (!d_src || !d_pairs || !d_fit)) || (m > 0 && !d_dst)) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(a->device);
		struct Restore { bs_assoc* a; int b; ~Restore() { a->backwards = b; } } restore{a, a->backwards};
		a->backwards = 0;
		a->run(d_src, n, d_dst, m, dof, d_pairs, d_fit, stream ? static_cast<cudaStream_t>(stream) : a->stream);
	});
}

int bs_assoc_merge_columns_device(bs_assoc* a, int ranks, int m, const double* d_all_min, const int32_t* d_all_first,
                                  const int32_t* d_all_last, double* d_col_min, int32_t* d_col_first, int32_t* d_col_last,
                                  void* stream) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		if (ranks < 1 || m < 0) fail(BS_ERR_ARG, "bad sizes");
		DeviceGuard g(a->device);
		launch_merge_columns(ranks, m, d_all_min, d_all_first, d_all_last, d_col_min, d_col_first, d_col_last,
		                     stream ? static_cast<cudaStream_t>(stream) : a->stream);
	});
}

int bs_assoc_merge_packed_device(bs_assoc* a, int ranks, int m, const void* d_packed, double* d_col_min, int32_t* d_col_first,
                                 int32_t* d_col_last, void* stream) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		if (ranks < 1 || m < 0) fail(BS_ERR_ARG, "bad sizes");
		if (m > 0 && (!d_packed || !d_col_min || !d_col_first || !d_col_last)) fail(BS_ERR_ARG, "null argument");
		DeviceGuard g(a->device);
		const char* base = static_cast<const char*>(d_packed);
		launch_merge_columns(ranks, m, reinterpret_cast<const double*>(base), reinterpret_cast<const int32_t*>(base + (size_t)m * 8),
		                     reinterpret_cast<const int32_t*>(base + (size_t)m * 12), d_col_min, d_col_first, d_col_last,
		                     stream ? static_cast<cudaStream_t>(stream) : a->stream, 2L * m, 4L * m);
	});
}

int bs_assoc_validate_device(bs_assoc* a, int n, int row_offset, int32_t* d_pairs, double* d_fit, const double* d_col_min,
                             const int32_t* d_col_first, const int32_t* d_col_last, void* stream) {
	return guarded([&] {
		if (!a) fail(BS_ERR_ARG, "handle is null");
		DeviceGuard g(a->device);
		launch_validate(n, row_offset, d_pairs, d_fit, d_col_min, d_col_first, d_col_last,
		                stream ? static_cast<cudaStream_t>(stream) : a->stream);
	});
}

// WrapAssociateGreedy.associate :85-92, FindUnassociated.checkDestination :56-72 — index bookkeeping on the host
int bs_assoc_matches(const int32_t* pairs, const double* fit, int n, int m, int32_t* match_src, int32_t* match_dst,
                     double* match_score, int* n_matches, int32_t* unassoc_src, int* n_unassoc_src, int32_t* unassoc_dst,
                     int* n_unassoc_dst) {
	return guarded([&] {
		if (n < 0 || m < 0 || (n > 0 && (!pairs || !fit))) fail(BS_ERR_ARG, "bad arguments");
		std::vector<char> hit((size_t)std::max(m, 1), 0);
		int nm = 0, nus = 0, nud = 0;
		for (int i = 0; i < n; i++) {
			const int d = pairs[i];
			if (d >= 0) {
				if (d >= m) fail(BS_ERR_ARG, "pair index out of range");
				if (match_src) match_src[nm] = i;
				if (match_dst) match_dst[nm] = d;
				if (match_score) match_score[nm] = fit[i];
				nm++; hit[d] = 1;
			} else {
				if (unassoc_src) unassoc_src[nus] = i;
				nus++;
			}
		}
		for (int j = 0; j < m; j++)
			if (!hit[j]) { if (unassoc_dst) unassoc_dst[nud] = j; nud++; }
		if (n_matches) *n_matches = nm;
		if (n_unassoc_src) *n_unassoc_src = nus;
		if (n_unassoc_dst) *n_unassoc_dst = nud;
	});
}

}  // extern "C"
