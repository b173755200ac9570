// bs_plan.cu — host-side planning: config validation, the scale-space schedule, Gaussian weight tables.
// (host code only; kept in a .cu so the whole library builds with one nvcc command line)
#include "bs_common.cuh"

#include <atomic>
#include <cmath>
#include <vector>

namespace bs {

static thread_local std::string g_last_error;
void set_last_error(const std::string& m) { g_last_error = m; }
const char* last_error_cstr() { return g_last_error.c_str(); }

static std::atomic<long> g_launches{0};
void count_launch(int n) { g_launches += n; }
long launch_count() { return g_launches.load(); }

// ---------------------------------------------------------------- config validation
// ConfigFastHessian.checkValidity (F/abst/feature/detect/interest/ConfigFastHessian.java:88-90) is empty and
// ConfigExtract.checkValidity (F/abst/feature/detect/extract/ConfigExtract.java:96-101) rejects radius<=0 and border<0;
// the rest are the limits of this implementation.
void check_fh_cfg(const bs_fh_cfg& c) {
	if (c.extract_radius <= 0) fail(BS_ERR_ARG, "Search radius must be >= 1");
	if (c.extract_radius > 7) fail(BS_ERR_ARG, "extract_radius > 7 is not supported");
	if (c.initial_sample_size < 1) fail(BS_ERR_ARG, "initial_sample_size must be >= 1");
	if (c.initial_size < 3) fail(BS_ERR_ARG, "initial_size must be >= 3");
	if (c.scales_per_octave < 3 || c.scales_per_octave > MAX_SCALES)
		fail(BS_ERR_ARG, "scales_per_octave must be in [3, " + std::to_string(MAX_SCALES) + "]");
	if (c.octaves < 1 || c.octaves > MAX_OCTAVES) fail(BS_ERR_ARG, "octaves must be in [1, " + std::to_string(MAX_OCTAVES) + "]");
	if (c.scale_step_size < 1) fail(BS_ERR_ARG, "scale_step_size must be >= 1");
}
void check_ori_cfg(const bs_ori_cfg& c) {
	if (c.variant != 0 && c.variant != 1) fail(BS_ERR_ARG, "orientation variant must be 0 (average) or 1 (sliding)");
	if (c.radius < 1 || 2 * c.radius + 1 > MAX_ORI_W) fail(BS_ERR_ARG, "orientation radius must be in [1,16]");
	if (!(c.object_radius_to_scale > 0)) fail(BS_ERR_ARG, "object_radius_to_scale must be > 0");
	if (!(c.sample_period > 0)) fail(BS_ERR_ARG, "sample_period must be > 0");
	if (c.sample_width < 1) fail(BS_ERR_ARG, "sample_width must be >= 1");
}
void check_desc_cfg(const bs_desc_cfg& c) {
	if (c.variant != 0 && c.variant != 1) fail(BS_ERR_ARG, "descriptor variant must be 0 (Speed) or 1 (Stability)");
	if (c.width_large_grid < 1 || c.width_sub_region < 1 || c.width_sample < 1) fail(BS_ERR_ARG, "descriptor widths must be >= 1");
	const int grid = c.width_large_grid * c.width_sub_region + (c.variant == 1 ? 2 * c.overlap : 0);
	if (grid > MAX_DESC_GRID) fail(BS_ERR_ARG, "descriptor sample grid larger than 32x32 is not supported");
	if (c.width_large_grid * c.width_large_grid * 4 > 256) fail(BS_ERR_ARG, "descriptor longer than 256 is not supported");
	if (c.variant == 1 && c.overlap < 0) fail(BS_ERR_ARG, "overlap must be >= 0");
}

// ---------------------------------------------------------------- layers / plan
Layer make_layer(int size, int skip, int W, int H) {
	Layer L{};
	L.size = size; L.skip = skip;
	L.w = W / skip; L.h = H / skip;
	const int blockW = size / 3, blockH = size - blockW - 1;
	L.r1 = blockW / 2; L.r2 = blockW + L.r1; L.r3 = blockH / 2;
	L.b = blockW;
	L.norm = 1.0f / (float)(size * size);
	return L;
}

// FastHessianFeatureDetector.detect :156-188 — octave/scale schedule; detectOctave :198-221 — which layers detect
Plan make_plan(const bs_fh_cfg& cfg, int W, int H) {
	Plan p{};
	p.W = W; p.H = H;
	p.radius = cfg.extract_radius;
	p.threshold = cfg.detect_threshold;
	p.max_per_level = cfg.max_features_per_scale;
	int skip = cfg.initial_sample_size, sizeStep = cfg.scale_step_size, octaveSize = cfg.initial_size;
	const int ns = cfg.scales_per_octave;
	long offset = 0;
	int ctas = 0;
	for (int o = 0; o < cfg.octaves; o++) {
		const int maxSize = octaveSize + (ns - 1) * sizeStep;
		if (maxSize > W || maxSize > H) break;
		const int first = p.n_layers;
		for (int i = 0; i < ns; i++) {
			Layer L = make_layer(octaveSize + i * sizeStep, skip, W, H);
			L.octave = o; L.index = i; L.offset = offset;
			offset += (long)L.w * L.h;
			p.layers[p.n_layers++] = L;
		}
		for (int i = 2; i < ns; i++) {
			Level v{};
			v.octave = o; v.level = i - 1;
			v.lower = first + i - 2; v.mid = first + i - 1; v.upper = first + i;
			const Layer& M = p.layers[v.mid];
			v.border = M.size / (2 * skip);
			v.level_size = M.size;
			v.level_step = M.size - p.layers[v.lower].size;
			const int step = cfg.extract_radius + 1;
			const int endX = M.w - v.border, endY = M.h - v.border;
			if (endX <= v.border || endY <= v.border) { v.nbx = 0; v.nby = 0; }
			else { v.nbx = div_up(endX - v.border, step); v.nby = div_up(endY - v.border, step); }
			v.cta_begin = ctas;
			ctas += div_up(v.nbx * v.nby, NMS_CTA_BLOCKS);
			p.levels[p.n_levels++] = v;
		}
		skip += skip; octaveSize += sizeStep; sizeStep += sizeStep;
	}
	p.inten_per_frame = offset;
	p.ctas_per_frame = ctas;
	return p;
}

// ---------------------------------------------------------------- Gaussian tables
// ddogleg UtilGaussian.computePDF(mean, sigma, x) = exp(-(x-mean)^2 / (2 sigma^2)) / (sigma sqrt(2 pi))
static double normal_pdf(double sigma, double x) { return std::exp(-x * x / (2.0 * sigma * sigma)) / (sigma * std::sqrt(2.0 * M_PI)); }
// FactoryKernelGaussian.sigmaForRadius(radius, order 0), IP/factory/filter/kernel/FactoryKernelGaussian.java:388-393
static double default_sigma(double radius) { return (radius * 2.0 + 1.0) / 5.0; }

// FactoryKernelGaussian.gaussian2D_F64(sigma, radius, odd, normalize) :297-306: outer product of the 1-D samples
// (gaussian1D_F64 :240-262, KernelMath.convolve2D IP/alg/filter/kernel/KernelMath.java:365-383), then divided by the sum
// (KernelMath.normalizeSumToOne :446-452)
void gaussian2d_odd(double sigma, int radius, double* out) {
	if (sigma <= 0) sigma = default_sigma(radius);
	const int w = 2 * radius + 1;
	std::vector<double> line(w);
	for (int i = 0; i < w; i++) line[i] = normal_pdf(sigma, (double)(radius - i));
	double sum = 0;
	for (int i = 0; i < w; i++)
		for (int j = 0; j < w; j++) { out[i * w + j] = line[i] * line[j]; }
	for (int k = 0; k < w * w; k++) sum += out[k];
	for (int k = 0; k < w * w; k++) out[k] /= sum;
}

// FactoryKernelGaussian.gaussianWidth(sigma, width) :418-448: even widths sample the pdf at half-pixel radial distances
void gaussian_width_table(double sigma, int width, double* out) {
	if (sigma <= 0) sigma = default_sigma(width / 2);
	else if (width <= 0) fail(BS_ERR_ARG, "Must specify the width since it doesn't know if it should be even or odd");
	if (width % 2 != 0) { gaussian2d_odd(sigma, width / 2, out); return; }
	const int r = width / 2 - 1;
	double sum = 0;
	for (int y = 0; y < width; y++) {
		const double dy = (y <= r ? std::abs(y - r) : std::abs(y - r - 1)) + 0.5;
		for (int x = 0; x < width; x++) {
			const double dx = (x <= r ? std::abs(x - r) : std::abs(x - r - 1)) + 0.5;
			const double v = normal_pdf(sigma, std::sqrt(dx * dx + dy * dy));
			out[y * width + x] = v;
			sum += v;
		}
	}
	for (int k = 0; k < width * width; k++) out[k] /= sum;
}

}  // namespace bs
