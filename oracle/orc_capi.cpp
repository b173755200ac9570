// ORACLE — TEST INFRASTRUCTURE ONLY.  extern "C" surface of the CPU restatement, loaded with ctypes by
// oracle/bsoracle.py (used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs).
#include "orc_surf.h"
#include "orc_math.h"
#include <algorithm>
#include <cstring>

using namespace orc;

namespace {
FastHessianConfig fh_from(const int* iv, float thr) {
	// iv = {extract_radius, max_features_per_scale, initial_sample_size, initial_size, scales_per_octave, octaves, scale_step_size}
	FastHessianConfig c; c.detect_threshold = thr; c.extract_radius = iv[0]; c.max_features_per_scale = iv[1];
	c.initial_sample_size = iv[2]; c.initial_size = iv[3]; c.scales_per_octave = iv[4]; c.octaves = iv[5];
	c.scale_step_size = iv[6];
	return c;
}
OrientationConfig oc_from(const double* v) {
	// v = {variant, object_radius_to_scale, radius, sample_period, sample_width, weight_sigma, window_size}
	OrientationConfig c; c.variant = (int)v[0]; c.object_radius_to_scale = v[1]; c.radius = (int)v[2];
	c.sample_period = v[3]; c.sample_width = (int)v[4]; c.weight_sigma = v[5]; c.window_size = v[6];
	return c;
}
DescribeConfig dc_from(const double* v) {
	// v = {variant, width_large_grid, width_sub_region, width_sample, use_haar, weight_sigma, overlap, sigma_large_grid, sigma_sub_region}
	DescribeConfig c; c.variant = (int)v[0]; c.width_large_grid = (int)v[1]; c.width_sub_region = (int)v[2];
	c.width_sample = (int)v[3]; c.use_haar = (int)v[4]; c.weight_sigma = v[5]; c.overlap = (int)v[6];
	c.sigma_large_grid = v[7]; c.sigma_sub_region = v[8];
	return c;
}
template <typename T> int detect_t(const T* ii, int stride, int w, int h, const int* iv, float thr, int max_out,
                                   double* xys, int* meta, float* resp, int nthreads) {
	std::vector<ScalePoint> found;
	fast_hessian_detect(Img<T>{ii, stride, w, h}, fh_from(iv, thr), found, nthreads);
	int n = (int)found.size();
	for (int i = 0; i < n && i < max_out; i++) {
		xys[3 * i] = found[i].x; xys[3 * i + 1] = found[i].y; xys[3 * i + 2] = found[i].scale;
		if (meta) { meta[4 * i] = found[i].octave; meta[4 * i + 1] = found[i].level; meta[4 * i + 2] = found[i].px; meta[4 * i + 3] = found[i].py; }
		if (resp) resp[i] = found[i].response;
	}
	return n;
}
template <typename TIn, typename TII> int surf_t(const TIn* px, int stride, int w, int h, const int* iv, float thr, const double* ov,
                                                 const double* dv, int max_out, double* xys, double* angle, double* desc,
                                                 uint8_t* white, int nthreads) {
	SurfResult r;
	surf_detect_describe<TIn, TII>(Img<TIn>{px, stride, w, h}, fh_from(iv, thr), oc_from(ov), dc_from(dv), r, nthreads);
	int n = (int)r.points.size();
	int dof = n ? (int)(r.desc.size() / n) : 0;
	for (int i = 0; i < n && i < max_out; i++) {
		xys[3 * i] = r.points[i].x; xys[3 * i + 1] = r.points[i].y; xys[3 * i + 2] = r.points[i].scale;
		angle[i] = r.angle[i]; white[i] = r.white[i];
		std::memcpy(desc + (size_t)i * dof, r.desc.data() + (size_t)i * dof, sizeof(double) * dof);
	}
	return n;
}
// colour SURF (surfColorFast / surfColorStable): bands = nb dense w x h planes; desc holds nb*dof doubles per feature
template <typename TIn, typename TII> static int surf_planar_t(const TIn* bands, int nb, int w, int h, const int* iv, float thr, const double* ov,
                                                               const double* dv, int max_out, double* xys, double* angle, double* desc,
                                                               uint8_t* white, int nthreads) {
	SurfResult r;
	surf_planar_detect_describe<TIn, TII>(bands, nb, w, h, fh_from(iv, thr), oc_from(ov), dc_from(dv), r, nthreads);
	int n = (int)r.points.size();
	int dof = n ? (int)(r.desc.size() / n) : 0;
	for (int i = 0; i < n && i < max_out; i++) {
		xys[3 * i] = r.points[i].x; xys[3 * i + 1] = r.points[i].y; xys[3 * i + 2] = r.points[i].scale;
		angle[i] = r.angle[i]; white[i] = r.white[i];
		std::memcpy(desc + (size_t)i * dof, r.desc.data() + (size_t)i * dof, sizeof(double) * dof);
	}
	return n;
}
}  // namespace

extern "C" {

double orc_atan2(double y, double x) { return fd_atan2(y, x); }
double orc_sin(double x) { return fd_sin(x); }
double orc_cos(double x) { return fd_cos(x); }
void orc_atan2_v(const double* y, const double* x, double* out, long n) { for (long i = 0; i < n; i++) out[i] = fd_atan2(y[i], x[i]); }
void orc_sincos_v(const double* a, double* s, double* c, long n) { for (long i = 0; i < n; i++) { s[i] = fd_sin(a[i]); c[i] = fd_cos(a[i]); } }

void orc_integral_f32(const float* in, int in_stride, float* out, int out_stride, int w, int h) {
	integral_f32(Img<float>{in, in_stride, w, h}, ImgMut<float>{out, out_stride, w, h});
}
void orc_integral_u8(const uint8_t* in, int in_stride, int32_t* out, int out_stride, int w, int h) {
	integral_u8(Img<uint8_t>{in, in_stride, w, h}, ImgMut<int32_t>{out, out_stride, w, h});
}
void orc_integral_s32(const int32_t* in, int in_stride, int32_t* out, int out_stride, int w, int h) {
	integral_s32(Img<int32_t>{in, in_stride, w, h}, ImgMut<int32_t>{out, out_stride, w, h});
}
float orc_block_unsafe_f32(const float* ii, int stride, int w, int h, int x0, int y0, int x1, int y1) { return block_unsafe(Img<float>{ii, stride, w, h}, x0, y0, x1, y1); }
float orc_block_zero_f32(const float* ii, int stride, int w, int h, int x0, int y0, int x1, int y1) { return block_zero(Img<float>{ii, stride, w, h}, x0, y0, x1, y1); }
int32_t orc_block_unsafe_s32(const int32_t* ii, int stride, int w, int h, int x0, int y0, int x1, int y1) { return block_unsafe(Img<int32_t>{ii, stride, w, h}, x0, y0, x1, y1); }
int32_t orc_block_zero_s32(const int32_t* ii, int stride, int w, int h, int x0, int y0, int x1, int y1) { return block_zero(Img<int32_t>{ii, stride, w, h}, x0, y0, x1, y1); }

// which: 0 = XX, 1 = YY, 2 = XY.  out = {n, then per block x0,y0,x1,y1,scale}
void orc_kernel_deriv(int which, int size, int* out) {
	IntegralKernel k = which == 0 ? kernel_deriv_xx(size) : which == 1 ? kernel_deriv_yy(size) : kernel_deriv_xy(size);
	out[0] = k.n;
	for (int i = 0; i < k.n; i++) { out[1 + 5 * i] = k.blocks[i].x0; out[2 + 5 * i] = k.blocks[i].y0; out[3 + 5 * i] = k.blocks[i].x1; out[4 + 5 * i] = k.blocks[i].y1; out[5 + 5 * i] = k.scales[i]; }
}

// intensity must hold (w/skip)*(h/skip) floats, dense
void orc_hessian_f32(const float* ii, int stride, int w, int h, int skip, int size, float* inten, int naive, int nthreads) {
	ImgMut<float> o{inten, w / skip, w / skip, h / skip};
	if (naive) hessian_naive(Img<float>{ii, stride, w, h}, skip, size, o); else hessian(Img<float>{ii, stride, w, h}, skip, size, o, nthreads);
}
void orc_hessian_s32(const int32_t* ii, int stride, int w, int h, int skip, int size, float* inten, int naive, int nthreads) {
	ImgMut<float> o{inten, w / skip, w / skip, h / skip};
	if (naive) hessian_naive(Img<int32_t>{ii, stride, w, h}, skip, size, o); else hessian(Img<int32_t>{ii, stride, w, h}, skip, size, o, nthreads);
}

// returns the number found; xy receives min(found,max_out) int16 pairs
int orc_nonmax(const float* inten, int stride, int w, int h, int radius, float thresh, int border, int naive, int16_t* xy, int max_out) {
	std::vector<PointI16> out;
	if (naive) nonmax_naive_strict_max(Img<float>{inten, stride, w, h}, radius, thresh, border, out);
	else nonmax_block_strict_max(Img<float>{inten, stride, w, h}, radius, thresh, border, out, 1);
	for (size_t i = 0; i < out.size() && (int)i < max_out; i++) { xy[2 * i] = out[i].x; xy[2 * i + 1] = out[i].y; }
	return (int)out.size();
}

float orc_poly_peak(float lo, float mid, float hi) { return poly_peak(lo, mid, hi); }

int orc_fh_detect_f32(const float* ii, int stride, int w, int h, const int* iv, float thr, int max_out, double* xys, int* meta, float* resp, int nthreads) {
	return detect_t<float>(ii, stride, w, h, iv, thr, max_out, xys, meta, resp, nthreads);
}
int orc_fh_detect_s32(const int32_t* ii, int stride, int w, int h, const int* iv, float thr, int max_out, double* xys, int* meta, float* resp, int nthreads) {
	return detect_t<int32_t>(ii, stride, w, h, iv, thr, max_out, xys, meta, resp, nthreads);
}

// weight tables: kind 0 = gaussian2d_f64(sigma, radius=n) -> (2n+1)^2 ; kind 1 = gaussian_width(sigma, width=n) -> n^2
int orc_gaussian(int kind, double sigma, int n, double* out) {
	std::vector<double> k = kind == 0 ? gaussian2d_f64(sigma, n) : gaussian_width(sigma, n);
	std::memcpy(out, k.data(), k.size() * sizeof(double));
	return (int)k.size();
}

void orc_gradient_f32(const float* ii, int stride, int w, int h, double width, const int* xy, int n, int safe, double* dxdy, int haar) {
	SparseGradient<float> g; g.ii = Img<float>{ii, stride, w, h}; g.haar = haar != 0; g.set_width(width);
	for (int i = 0; i < n; i++) { if (safe) g.compute_safe(xy[2 * i], xy[2 * i + 1], dxdy[2 * i], dxdy[2 * i + 1]); else g.compute(xy[2 * i], xy[2 * i + 1], dxdy[2 * i], dxdy[2 * i + 1]); }
}

int orc_is_inside_rotated(int w, int h, double X, double Y, int radius_regions, int kernel_size, double scale, double c, double s) { return is_inside_rotated(w, h, X, Y, radius_regions, kernel_size, scale, c, s); }
int orc_is_inside_rect(int w, int h, double tl_x, double tl_y, double region_size, double sample_size) { return is_inside_rect(w, h, tl_x, tl_y, region_size, sample_size); }

// xys = n x (x, y, scale); object radius = scale*2 as in WrapDetectDescribeSurf.computeDescriptors
void orc_orientation_f32(const float* ii, int stride, int w, int h, const double* ov, const double* xys, int n, double* angle) {
	Orientation<float> o(oc_from(ov)); o.set_image(Img<float>{ii, stride, w, h});
	for (int i = 0; i < n; i++) { o.set_object_radius(xys[3 * i + 2] * 2.0); angle[i] = o.compute(xys[3 * i], xys[3 * i + 1]); }
}
void orc_orientation_s32(const int32_t* ii, int stride, int w, int h, const double* ov, const double* xys, int n, double* angle) {
	Orientation<int32_t> o(oc_from(ov)); o.set_image(Img<int32_t>{ii, stride, w, h});
	for (int i = 0; i < n; i++) { o.set_object_radius(xys[3 * i + 2] * 2.0); angle[i] = o.compute(xys[3 * i], xys[3 * i + 1]); }
}
void orc_describe_f32(const float* ii, int stride, int w, int h, const double* dv, const double* xys, const double* angle, int n, double* desc, uint8_t* white) {
	Describe<float> d(dc_from(dv)); d.set_image(Img<float>{ii, stride, w, h});
	for (int i = 0; i < n; i++) { bool wh; d.describe(xys[3 * i], xys[3 * i + 1], angle[i], xys[3 * i + 2], desc + (size_t)i * d.dof, &wh); white[i] = wh; }
}
void orc_describe_s32(const int32_t* ii, int stride, int w, int h, const double* dv, const double* xys, const double* angle, int n, double* desc, uint8_t* white) {
	Describe<int32_t> d(dc_from(dv)); d.set_image(Img<int32_t>{ii, stride, w, h});
	for (int i = 0; i < n; i++) { bool wh; d.describe(xys[3 * i], xys[3 * i + 1], angle[i], xys[3 * i + 2], desc + (size_t)i * d.dof, &wh); white[i] = wh; }
}
void orc_normalize_l2(double* v, int n) { normalize_l2(v, n); }

// full pipeline; returns number of features found (outputs hold min(found, max_out))
int orc_surf_f32(const float* px, int stride, int w, int h, const int* iv, float thr, const double* ov, const double* dv, int max_out,
                 double* xys, double* angle, double* desc, uint8_t* white, int nthreads) {
	return surf_t<float, float>(px, stride, w, h, iv, thr, ov, dv, max_out, xys, angle, desc, white, nthreads);
}
int orc_surf_u8(const uint8_t* px, int stride, int w, int h, const int* iv, float thr, const double* ov, const double* dv, int max_out,
                double* xys, double* angle, double* desc, uint8_t* white, int nthreads) {
	return surf_t<uint8_t, int32_t>(px, stride, w, h, iv, thr, ov, dv, max_out, xys, angle, desc, white, nthreads);
}

int orc_surf_planar_f32(const float* bands, int nb, int w, int h, const int* iv, float thr, const double* ov, const double* dv, int max_out,
                        double* xys, double* angle, double* desc, uint8_t* white, int nthreads) {
	return surf_planar_t<float, float>(bands, nb, w, h, iv, thr, ov, dv, max_out, xys, angle, desc, white, nthreads);
}
int orc_surf_planar_u8(const uint8_t* bands, int nb, int w, int h, const int* iv, float thr, const double* ov, const double* dv, int max_out,
                       double* xys, double* angle, double* desc, uint8_t* white, int nthreads) {
	return surf_planar_t<uint8_t, int32_t>(bands, nb, w, h, iv, thr, ov, dv, max_out, xys, angle, desc, white, nthreads);
}
void orc_planar_average_f32(const float* bands, int nb, int w, int h, float* gray) { planar_average<float>(bands, nb, w, h, gray); }
void orc_planar_average_u8(const uint8_t* bands, int nb, int w, int h, uint8_t* gray) { planar_average<uint8_t>(bands, nb, w, h, gray); }
void orc_describe_raw_f32(const float* ii, int stride, int w, int h, const double* dv, const double* xys, const double* angle, int n, double* desc) {
	Describe<float> d(dc_from(dv)); d.set_image(Img<float>{ii, stride, w, h});
	for (int i = 0; i < n; i++) d.describe_raw(xys[3 * i], xys[3 * i + 1], angle[i], xys[3 * i + 2], desc + (size_t)i * d.dof);
}
int orc_laplace_sign_f32(const float* ii, int stride, int w, int h, const double* dv, int x, int y, double scale) {
	Describe<float> d(dc_from(dv)); d.set_image(Img<float>{ii, stride, w, h});
	return d.laplace_sign(x, y, scale) ? 1 : 0;
}

double orc_euclidean_sq(const double* a, const double* b, int n) { return euclidean_sq(a, b, n); }
void orc_associate(const double* src, int n, const double* dst, int m, int dof, double max_fit, int backwards, int score_type,
                   int32_t* pairs, double* fit, int nthreads) {
	associate_greedy(src, n, dst, m, dof, max_fit, backwards != 0, score_type, pairs, fit, nthreads);
}
void orc_associate_nobuffer(const double* src, int n, const double* dst, int m, int dof, double max_fit, int backwards,
                            int32_t* pairs, double* fit, int nthreads) {
	associate_greedy_nobuffer(src, n, dst, m, dof, max_fit, backwards != 0, pairs, fit, nthreads);
}

}  // extern "C"
