// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_math.h header note).  PARITY STATUS: the reference is 100 % Java and
// cannot be compiled or run in this image (no JDK, dependency jars not vendored), so this oracle is
// a restatement pinned ONLY by the reference's own known-answer / differential tests re-expressed in
// tests/test_oracle_*.py.  End-to-end feature lists have no golden vectors in the reference:
// "parity unpinned" for those.
//
// C++17 scalar restatement of BoofCV's SURF path.  Paths below are relative to /root/reference:
//   IP = main/boofcv-ip/src/main/java/boofcv      F = main/boofcv-feature/src/main/java/boofcv
// float where Java uses float, double where Java uses double, left-to-right evaluation, no FMA
// (-ffp-contract=off), (int) = truncation.
#pragma once
#include <cstdint>
#include <vector>

namespace orc {

// Image view = {data, startIndex folded into the pointer, stride, width, height} (TY/struct/image/ImageBase.java:34-52)
template <typename T> struct Img {
	const T* data; int stride; int width; int height;
	T at(int x, int y) const { return data[(long)y * stride + x]; }
};
template <typename T> struct ImgMut {
	T* data; int stride; int width; int height;
	Img<T> view() const { return Img<T>{data, stride, width, height}; }
};

// ---- IP/alg/transform/ii/impl/ImplIntegralImageOps.java:42-144 ----
void integral_f32(Img<float> in, ImgMut<float> out);
void integral_u8(Img<uint8_t> in, ImgMut<int32_t> out);
void integral_s32(Img<int32_t> in, ImgMut<int32_t> out);

// ---- ImplIntegralImageOps.java:185-258 ----
float block_unsafe(Img<float> ii, int x0, int y0, int x1, int y1);
float block_zero(Img<float> ii, int x0, int y0, int x1, int y1);
int32_t block_unsafe(Img<int32_t> ii, int x0, int y0, int x1, int y1);
int32_t block_zero(Img<int32_t> ii, int x0, int y0, int x1, int y1);

// ---- IP/alg/transform/ii/IntegralKernel.java:30-48, DerivativeIntegralImage.java:102-158 ----
struct Rect { int x0, y0, x1, y1; };
struct IntegralKernel { int n; Rect blocks[4]; int scales[4]; };
IntegralKernel kernel_deriv_xx(int size);
IntegralKernel kernel_deriv_yy(int size);
IntegralKernel kernel_deriv_xy(int size);
float convolve_sparse(Img<float> ii, const IntegralKernel& k, int x, int y);      // ImplIntegralImageOps.java:172-183
int32_t convolve_sparse(Img<int32_t> ii, const IntegralKernel& k, int x, int y);  // :216-227

// ---- F/alg/feature/detect/intensity/impl/ImplIntegralImageFeatureIntensity.java ----
// intensity is (ii.width/skip) x (ii.height/skip), dense.  hessian = hessianBorder + hessianInner
// (IntegralImageFeatureIntensity.java:43-79); nthreads>1 threads over rows like the _MT twin.
template <typename T> void hessian_naive(Img<T> ii, int skip, int size, ImgMut<float> inten);
template <typename T> void hessian(Img<T> ii, int skip, int size, ImgMut<float> inten, int nthreads = 1);

// ---- F/alg/feature/detect/extract/NonMaxBlock.java:69-94 + NonMaxBlockSearchStrict.java:56-79,196-221 ----
struct PointI16 { int16_t x, y; };
void nonmax_block_strict_max(Img<float> inten, int radius, float threshold_max, int border,
                             std::vector<PointI16>& out, int nthreads = 1);
// GenericNonMaxTests' reference-of-reference (NonMaxExtractorNaive) used by the differential self test
void nonmax_naive_strict_max(Img<float> inten, int radius, float threshold_max, int border, std::vector<PointI16>& out);

// ---- F/abst/feature/detect/interest/ConfigFastHessian.java:33-70 ----
struct FastHessianConfig {
	float detect_threshold = 1; int extract_radius = 2; int max_features_per_scale = -1;
	int initial_sample_size = 1; int initial_size = 9; int scales_per_octave = 4; int octaves = 4;
	int scale_step_size = 6;
};
// ScalePoint (F/struct/feature/ScalePoint.java:28-34) + where it came from (for ordering checks)
struct ScalePoint { double x, y, scale; int octave, level; int16_t px, py; float response; };

// ---- F/alg/feature/detect/interest/FastHessianFeatureDetector.java:156-350 ----
float poly_peak(float lower, float middle, float upper);
template <typename T> void fast_hessian_detect(Img<T> ii, const FastHessianConfig& cfg, std::vector<ScalePoint>& found,
                                               int nthreads = 1);

// ---- Gaussian weight tables: IP/factory/filter/kernel/FactoryKernelGaussian.java:120-135,240-262,297-306,388-448,
//      IP/alg/filter/kernel/KernelMath.java:365-383,446-452; pdf = ddogleg UtilGaussian.computePDF ----
std::vector<double> gaussian2d_f64(double sigma, int radius);   // odd width 2r+1, sum-normalised; sigma<=0 => from radius
std::vector<double> gaussian_width(double sigma, int width);    // even or odd width, sum-normalised

// ---- sparse gradient: IP/alg/transform/ii/SparseIntegralGradient_NoBorder.java:41-47,
//      impl/SparseIntegralGradient_NoBorder_F32.java:37-76 (+_I32), TY/struct/sparse/SparseScaleGradient.java:47-50,
//      TY/struct/sparse/SparseGradientSafe.java:56-61 ----
template <typename T> struct SparseGradient {
	Img<T> ii; int r = 1, w = 3;
	bool haar = false;                       // SparseIntegralHaar_NoBorder_F32/_I32 (IP/alg/transform/ii/impl/...Haar_NoBorder_F32.java:37-73)
	void set_width(double width);
	bool in_bounds(int x, int y) const;
	void compute(int x, int y, double& dx, double& dy) const;        // unchecked, like the Java (reads raw indices)
	void compute_safe(int x, int y, double& dx, double& dy) const;   // (0,0) when out of bounds
};

// ---- F/alg/feature/describe/SurfDescribeOps.java:119-159,183-204 ----
bool is_inside_rotated(int width, int height, double X, double Y, int radius_regions, int kernel_size, double scale,
                       double c, double s);
bool is_inside_rect(int width, int height, double tl_x, double tl_y, double region_size, double sample_size);

// ---- orientation: F/alg/feature/orientation/OrientationIntegralBase.java:75-103,
//      impl/ImplOrientationAverageGradientIntegral.java:54-127, impl/ImplOrientationSlidingWindowIntegral.java:81-188,
//      configs F/abst/feature/orientation/ConfigAverageIntegral.java:34-51, ConfigSlidingIntegral.java:34-54 ----
struct OrientationConfig {
	int variant = 0;                       // 0 average (surfFast), 1 sliding window (surfStable)
	double object_radius_to_scale = 0.5;   // 1/BoofDefaults.SURF_SCALE_TO_RADIUS
	int radius = 6; double sample_period = 1; int sample_width = 6; double weight_sigma = -1;
	double window_size = 1.0471975511965976;   // Math.PI/3.0
	static OrientationConfig average() { return OrientationConfig(); }
	static OrientationConfig sliding() { OrientationConfig c; c.variant = 1; c.radius = 8; c.sample_period = 0.65; return c; }
};
template <typename T> struct Orientation {
	OrientationConfig cfg; std::vector<double> weights; double scale = 1; SparseGradient<T> g;
	explicit Orientation(const OrientationConfig& c);
	void set_image(Img<T> ii) { g.ii = ii; }
	void set_object_radius(double radius);
	double compute(double c_x, double c_y);
};

// ---- descriptor: F/alg/feature/describe/DescribePointSurf.java:112-139,169-313, DescribePointSurfMod.java:71-192,
//      F/alg/descriptor/UtilFeature.java:101-114, config F/abst/feature/describe/ConfigSurfDescribe.java:34-79 ----
struct DescribeConfig {
	int variant = 0;                       // 0 DescribePointSurf (surfFast), 1 DescribePointSurfMod (surfStable)
	int width_large_grid = 4, width_sub_region = 5, width_sample = 3; int use_haar = 0;
	double weight_sigma = 4.5;             // Speed
	int overlap = 2; double sigma_large_grid = 2.5, sigma_sub_region = 2.5;   // Stability
	static DescribeConfig speed() { return DescribeConfig(); }
	static DescribeConfig stability() { DescribeConfig c; c.variant = 1; return c; }
};
template <typename T> struct Describe {
	DescribeConfig cfg; std::vector<double> weight, weight_grid, weight_sub; int radius_descriptor; int dof;
	Img<T> ii; SparseGradient<T> g;
	explicit Describe(const DescribeConfig& c);
	void set_image(Img<T> im) { ii = im; g.ii = im; }
	void describe(double x, double y, double angle, double scale, double* desc, bool* white);
	// DescribePointSurf.describe(x,y,angle,scale,TupleDesc_F64) :195-213 — neither normalised nor signed
	void describe_raw(double x, double y, double angle, double scale, double* desc);
	void features_fast(double c_x, double c_y, double c, double s, double scale, bool safe, double* f);
	void features_mod(double c_x, double c_y, double c, double s, double scale, bool safe, double* f);
	bool laplace_sign(int x, int y, double scale);
};
void normalize_l2(double* v, int n);

// ---- F/abst/feature/detdesc/WrapDetectDescribeSurf.java:75-110 (+_MT :44-62) ----
struct SurfResult {
	std::vector<ScalePoint> points; std::vector<double> angle; std::vector<double> desc; std::vector<uint8_t> white;
};
template <typename TIn, typename TII>
void surf_detect_describe(Img<TIn> input, const FastHessianConfig& fh, const OrientationConfig& oc, const DescribeConfig& dc,
                          SurfResult& out, int nthreads = 1);

// ---- colour SURF: core/image/impl/ImplConvertPlanarToGray.java:44-84,213-253,296-336 (average),
//      F/abst/feature/detdesc/SurfPlanar_to_DetectDescribePoint.java:58-74, F/alg/feature/detdesc/DetectDescribeSurfPlanar.java:87-116,
//      F/alg/feature/describe/DescribePointSurfPlanar.java:97-111 ----
// bands: nb dense w x h images back to back.  desc: nb*dof doubles per feature.
template <typename TIn> void planar_average(const TIn* bands, int nb, int w, int h, TIn* gray);
template <typename TIn, typename TII>
void surf_planar_detect_describe(const TIn* bands, int nb, int w, int h, const FastHessianConfig& fh, const OrientationConfig& oc,
                                 const DescribeConfig& dc, SurfResult& out, int nthreads = 1);

// ---- F/alg/descriptor/DescriptorDistance.java:55-64, F/alg/feature/associate/AssociateGreedy.java:65-118 ----
double euclidean_sq(const double* a, const double* b, int n);
// score_type 0 = ScoreAssociateEuclideanSq_F64, 1 = ScoreAssociateEuclidean_F64 (sqrt; used by the reference's 1-d tests)
void associate_greedy(const double* src, int n, const double* dst, int m, int dof, double max_fit_error, bool backwards,
                      int score_type, int32_t* pairs, double* fit, int nthreads = 1);
// same result without the N x M work buffer (for sizes the reference cannot allocate)
void associate_greedy_nobuffer(const double* src, int n, const double* dst, int m, int dof, double max_fit_error,
                               bool backwards, int32_t* pairs, double* fit, int nthreads = 1);

}  // namespace orc
