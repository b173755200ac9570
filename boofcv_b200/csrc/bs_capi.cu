// bs_capi.cu — the extern "C" surface declared in include/boofsurf.h: handles, device workspaces, stream orchestration.
// Reference-side counterparts: WrapDetectDescribeSurf (F/abst/feature/detdesc/WrapDetectDescribeSurf.java:75-141) and
// WrapAssociateGreedy (F/abst/feature/associate/WrapAssociateGreedy.java:58-123).
#include "bs_common.cuh"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
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

// OrientationIntegralBase ctor (:75-92) + DescribePointSurf / DescribePointSurfMod ctors
// (DescribePointSurf.java:112-139, DescribePointSurfMod.java:71-99): parameters and weight tables
struct DescribeSetup {
	DescribeParams p{};
	std::vector<double> ori_w, w, wgrid;
	DevBuf d_ori_w, d_w, d_wgrid;
	DescribeTables tables{};

	void build(const bs_ori_cfg& o, const bs_desc_cfg& d) {
		p.ori_variant = o.variant; p.ori_radius = o.radius; p.ori_sample_width = o.sample_width;
		p.ori_radius_to_scale = o.object_radius_to_scale; p.ori_period = o.sample_period; p.ori_window = o.window_size;
		p.ori_weighted = o.weight_sigma != 0 ? 1 : 0;
		const int sw = 2 * o.radius + 1;
		ori_w.assign((size_t)sw * sw, 0.0);
		if (p.ori_weighted) gaussian2d_odd(o.weight_sigma, o.radius, ori_w.data());

		p.variant = d.variant; p.grid = d.width_large_grid; p.sub = d.width_sub_region; p.sample_width = d.width_sample;
		p.overlap = d.variant == 1 ? d.overlap : 0;
		p.dof = d.width_large_grid * d.width_large_grid * 4;
		const int region = d.width_large_grid * d.width_sub_region;
		p.radius_descriptor = region / 2 + p.overlap;
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
			BS_CUDA(cudaMemcpy(b.p, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice));
			return b.as<double>();
		};
		tables.d_ori_w = p.ori_weighted ? up(d_ori_w, ori_w) : nullptr;
		tables.d_w = up(d_w, w);
		tables.d_wgrid = up(d_wgrid, wgrid);
	}
	void release() { d_ori_w.release(); d_w.release(); d_wgrid.release(); }
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
struct bs_surf {
	int device = 0;
	bs_fh_cfg fh{}; bs_ori_cfg ori{}; bs_desc_cfg desc{};
	int pixel_type = BS_PIXEL_F32;
	cudaStream_t stream = nullptr;
	cudaEvent_t done = nullptr;
	DescribeSetup ds;
	// workspace
	int chunk_req = 0; long pool_req = 0;
	int W = 0, H = 0, chunk = 0;
	Plan plan{};
	DevBuf d_plan, d_in, d_ii, d_inten, d_stage, d_counts, d_offsets, d_frame_offset;
	PoolBufs pool;
	void* h_pinned = nullptr; size_t h_pinned_bytes = 0;
	// last result
	int batch = 0;
	bool pending = false;
	std::vector<int> h_frame_offset;

	void ensure(int w, int h, int nbatch);
	void run(const void* d_pixels, long frame_stride, int stride, int w, int h, int nbatch, cudaStream_t st);
	void finish();
};

void bs_surf::ensure(int w, int h, int nbatch) {
	if (w <= 0 || h <= 0) fail(BS_ERR_ARG, "image width and height must be positive");
	if (w > 32767 || h > 32767) fail(BS_ERR_ARG, "image larger than 32767 (QueueCorner stores int16 coordinates)");
	if (w != W || h != H) {
		plan = make_plan(fh, w, h);
		d_plan.reserve(sizeof(Plan));
		BS_CUDA(cudaMemcpy(d_plan.p, &plan, sizeof(Plan), cudaMemcpyHostToDevice));
		W = w; H = h;
	}
	// frames per wave: enough to fill the GPU, small enough to keep the intensity workspace L2-friendly
	int c = chunk_req > 0 ? chunk_req : 4;
	c = std::min(c, nbatch);
	chunk = std::max(c, 1);
	const size_t px = (size_t)W * H;
	d_ii.reserve(px * 4 * chunk);
	d_inten.reserve((size_t)std::max<long>(plan.inten_per_frame, 1) * 4 * chunk);
	const size_t nctas = (size_t)std::max(plan.ctas_per_frame, 1) * chunk;
	d_stage.reserve(nctas * NMS_CTA_BLOCKS * sizeof(RawPoint));
	d_counts.reserve(nctas * sizeof(int));
	d_offsets.reserve(nctas * sizeof(int));
	d_frame_offset.reserve((size_t)(nbatch + 1) * sizeof(int));
	long cap = pool_req > 0 ? pool_req : (long)nbatch * std::max<long>((long)px / 64, 256);
	pool.reserve(cap, ds.p.dof);
}

// the whole pipeline for a device-resident batch, chunk by chunk, all on `st`
void bs_surf::run(const void* d_pixels, long frame_stride, int stride, int w, int h, int nbatch, cudaStream_t st) {
	if (nbatch <= 0) fail(BS_ERR_ARG, "batch must be >= 1");
	if (stride < w) fail(BS_ERR_ARG, "stride must be >= width");
	if (plan.max_per_level > 0 && false) {}
	ensure(w, h, nbatch);
	batch = nbatch;
	int* fo = d_frame_offset.as<int>();
	BS_CUDA(cudaMemsetAsync(fo, 0, sizeof(int), st));
	for (int f0 = 0; f0 < nbatch; f0 += chunk) {
		const int c = std::min(chunk, nbatch - f0);
		if (pixel_type == BS_PIXEL_F32) {
			const float* in = static_cast<const float*>(d_pixels) + (long)f0 * frame_stride;
			launch_integral<float, float>(in, frame_stride, stride, d_ii.as<float>(), W, H, c, st);
			launch_hessian_plan<float>(d_ii.as<float>(), plan, d_plan.as<Plan>(), d_inten.as<float>(), c, st);
		} else if (pixel_type == BS_PIXEL_U8) {
			const uint8_t* in = static_cast<const uint8_t*>(d_pixels) + (long)f0 * frame_stride;
			launch_integral<uint8_t, int32_t>(in, frame_stride, stride, d_ii.as<int32_t>(), W, H, c, st);
			launch_hessian_plan<int32_t>(d_ii.as<int32_t>(), plan, d_plan.as<Plan>(), d_inten.as<float>(), c, st);
		} else {
			const int32_t* in = static_cast<const int32_t*>(d_pixels) + (long)f0 * frame_stride;
			launch_integral<int32_t, int32_t>(in, frame_stride, stride, d_ii.as<int32_t>(), W, H, c, st);
			launch_hessian_plan<int32_t>(d_ii.as<int32_t>(), plan, d_plan.as<Plan>(), d_inten.as<float>(), c, st);
		}
		launch_nms(plan, d_plan.as<Plan>(), d_inten.as<float>(), c, d_stage.as<RawPoint>(), d_counts.as<int>(), st);
		launch_compact(plan, c, f0, d_stage.as<RawPoint>(), d_counts.as<int>(), d_offsets.as<int>(), fo, pool.view, st);
		if (pixel_type == BS_PIXEL_F32)
			launch_describe<float>(d_ii.as<float>(), W, H, ds.p, ds.tables, pool.view, fo + f0, fo + f0 + c, f0, true, true, st);
		else
			launch_describe<int32_t>(d_ii.as<int32_t>(), W, H, ds.p, ds.tables, pool.view, fo + f0, fo + f0 + c, f0, true, true, st);
	}
	BS_CUDA(cudaEventRecord(done, st));
	pending = true;
}

void bs_surf::finish() {
	if (!pending) return;
	BS_CUDA(cudaEventSynchronize(done));
	h_frame_offset.assign(batch + 1, 0);
	BS_CUDA(cudaMemcpy(h_frame_offset.data(), d_frame_offset.p, (size_t)(batch + 1) * sizeof(int), cudaMemcpyDeviceToHost));
	pending = false;
}
