// bs_common.cuh — shared declarations of libboofsurf's CUDA translation units (sm_100a only).
// All device arithmetic that must match the reference's Java (strict IEEE-754, no FMA contraction, left-to-right) relies
// on this library being compiled with -fmad=false.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>

#include "../../include/boofsurf.h"

namespace bs {

// ---------------------------------------------------------------- errors
struct Error : std::runtime_error {
	int code;
	Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
void set_last_error(const std::string& m);
[[noreturn]] inline void fail(int code, const std::string& m) { throw Error(code, m); }

#define BS_CUDA(expr)                                                                                         \
	do {                                                                                                      \
		cudaError_t _e = (expr);                                                                              \
		if (_e != cudaSuccess)                                                                                \
			::bs::fail(BS_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + ":" + \
			                            std::to_string(__LINE__) + ")");                                      \
	} while (0)

// every kernel launch goes through this so that bs_launch_count() is an honest count
void count_launch(int n = 1);
#define BS_LAUNCH(kernel, grid, block, smem, stream, ...)            \
	do {                                                             \
		kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);  \
		::bs::count_launch();                                        \
		BS_CUDA(cudaGetLastError());                                 \
	} while (0)

inline int div_up(int a, int b) { return (a + b - 1) / b; }
// SM count of the CURRENT device, cached per device: a process may hold handles on several GPUs (the C ABI takes a device on
// every create call), so nothing about a device may live in a process-wide static keyed to whichever device came first.
// Function attributes (opt-in shared memory sizes) are per device too and are simply set before every launch that needs one.
inline int device_sm_count() {
	static int cache[64] = {0};
	int dev = 0;
	BS_CUDA(cudaGetDevice(&dev));
	if (dev >= 0 && dev < 64 && cache[dev] > 0) return cache[dev];
	int n = 0;
	BS_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
	if (dev >= 0 && dev < 64) cache[dev] = n;
	return n;
}
inline long div_up_l(long a, long b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------- scale-space plan
constexpr int MAX_OCTAVES = 8;
constexpr int MAX_SCALES = 8;
constexpr int MAX_LAYERS = MAX_OCTAVES * MAX_SCALES;
constexpr int MAX_LEVELS = MAX_OCTAVES * (MAX_SCALES - 2);

// One box-filter layer of the Fast-Hessian scale space (one `size` at one `skip`).
// Geometry per DerivativeIntegralImage.kernelDerivXX/YY/XY (IP/alg/transform/ii/DerivativeIntegralImage.java:102-158).
struct Layer {
	int size, skip;
	int w, h;             // intensity dimensions = W/skip, H/skip
	long offset;          // float offset of this layer inside one frame's intensity block
	int r1, r2, r3;       // XX/YY block radii: r1 = (size/3)/2, r2 = size/3 + r1, r3 = (size - size/3 - 1)/2
	int b;                // XY block = size/3
	float norm;           // 1f/(size*size)
	int octave, index;    // index inside the octave
};

// One detection level: NMS on `mid`, 3x3 checks in `lower`/`upper`
// (FastHessianFeatureDetector.findLocalScaleSpaceMax, F/alg/feature/detect/interest/FastHessianFeatureDetector.java:230-298)
struct Level {
	int lower, mid, upper;     // layer ids
	int octave, level;         // level = index of mid inside the octave
	int border;                // ignoreBorder = size[level]/(2*skip)
	int nbx, nby;              // NMS blocks of (radius+1)^2 starting at `border`
	int cta_begin;             // first CTA (of NMS_CTA_BLOCKS blocks each) of this level inside a frame
	int level_size, level_step;  // sizes[level], sizes[level]-sizes[level-1]
};

struct Plan {
	int W, H;
	int n_layers, n_levels;
	int radius;                // extractRadius
	float threshold;           // detectThreshold
	int max_per_level;         // maxFeaturesPerScale (<=0: off)
	long inten_per_frame;      // floats
	int ctas_per_frame;        // NMS CTAs per frame
	Layer layers[MAX_LAYERS];
	Level levels[MAX_LEVELS];
};
#ifndef BS_NMS_CTA
#define BS_NMS_CTA 64
#endif
constexpr int NMS_CTA_BLOCKS = BS_NMS_CTA;   // NMS blocks (=threads) per CTA; also the staging slots per CTA

Plan make_plan(const bs_fh_cfg& cfg, int W, int H);
void check_fh_cfg(const bs_fh_cfg& c);
void check_ori_cfg(const bs_ori_cfg& c);
void check_desc_cfg(const bs_desc_cfg& c);

// Detected point as it leaves the detector kernel (ScalePoint + provenance, F/struct/feature/ScalePoint.java:28-34)
struct alignas(8) RawPoint {
	double x, y, scale;
	float response;
	int16_t px, py;
	int16_t octave, level;
	int32_t pad;
};

// ---------------------------------------------------------------- weight tables (host, double precision)
// FactoryKernelGaussian / KernelMath, see bs_plan.cu
void gaussian2d_odd(double sigma, int radius, double* out);          // (2r+1)^2, sum-normalised
void gaussian_width_table(double sigma, int width, double* out);     // width^2, sum-normalised

// ---------------------------------------------------------------- describe parameters (device constant block)
constexpr int MAX_ORI_W = 33;     // orientation sample grid up to 33x33 (radius <= 16)
constexpr int MAX_DESC_GRID = 32; // descriptor sample grid up to 32x32

struct DescribeParams {
	// orientation
	int ori_variant, ori_radius, ori_sample_width;
	double ori_radius_to_scale, ori_period, ori_window;
	int ori_weighted;
	// descriptor
	int variant, grid, sub, sample_width, overlap, dof;
	int radius_descriptor;
	int use_haar;      // ConfigSurfDescribe.useHaar: the descriptor samples the Haar wavelet instead of the gradient kernel
	// WrapDetectDescribeSurf sets the orientation's object radius to scale*SURF_SCALE_TO_RADIUS (2.0),
	// DetectDescribeSurfPlanar (F/alg/feature/detdesc/DetectDescribeSurfPlanar.java:108) to the scale itself (1.0)
	double ori_object_radius_mult;
	// where a feature's `dof` values go: pool.desc + f*out_stride + out_offset.  finish = 1: L2-normalise them and compute
	// the Laplacian sign (grey SURF); 0: raw sums of one band (colour SURF, finished by launch_planar_finish)
	int out_stride, out_offset, finish;
};

// ---------------------------------------------------------------- stage launchers (implemented in the .cu files)
template <typename TI, typename TO>
void launch_integral(const TI* d_in, long in_frame_stride, int in_stride, TO* d_out, int W, int H, int batch, cudaStream_t st);
// parallel (any-association) scan: exact for integer images, opt-in for float; d_scratch >= integral_scan_scratch_bytes()
size_t integral_scan_scratch_bytes(int W, int H, int batch);
template <typename TI, typename TO>
bool launch_integral_scan(const TI* d_in, long in_frame_stride, int in_stride, TO* d_out, void* d_scratch, int W, int H, int batch, cudaStream_t st);

template <typename T>
void launch_hessian_plan(const T* d_ii, const Plan& plan, const Plan* d_plan, float* d_inten, int batch, cudaStream_t st);
template <typename T>
void launch_hessian_single(const T* d_ii, int W, int H, int skip, int size, float* d_out, cudaStream_t st);

// Feature pool of one detect call (structure of arrays, capacity `cap` features)
struct Pool {
	double* xy;        // [cap][2]   ScalePoint.x, .y
	double* scale;     // [cap]      ScalePoint.scale
	int32_t* meta;     // [cap][4]   octave, level, px, py
	float* response;   // [cap]
	int32_t* frame;    // [cap]      frame index inside the batch
	double* angle;     // [cap]
	double* desc;      // [cap][dof]
	uint8_t* white;    // [cap]
	long cap;
};

// d_keys / d_keep: [batch][ctas_per_frame * NMS_CTA_BLOCKS] scratch, needed only when plan.max_per_level > 0
void launch_nonmax_single(const float* d_img, int stride, int W, int H, int radius, float thr, int border, int* d_cand, int cap, int16_t* d_out,
                          int* d_total, cudaStream_t st);
void launch_nms(const Plan& plan, const Plan* d_plan, const float* d_inten, int batch, RawPoint* d_stage, int* d_cta_counts,
                unsigned* d_keys, unsigned char* d_keep, cudaStream_t st);
// Exclusive scan of the per-CTA counts of the chunk's frames (frame0 .. frame0+batch), continuing from
// d_frame_offset[frame0]; writes d_frame_offset[frame0+1 .. frame0+batch] and the per-CTA pool offsets, then gathers the
// staged points into the pool in (frame, octave, level, block-row, block-column) order.
// d_cta_offsets needs batch * ctas_per_frame + batch ints (the frame totals live behind the per-CTA offsets).
void launch_compact(const Plan& plan, int batch, int frame0, const RawPoint* d_stage, const int* d_cta_counts, int* d_cta_offsets,
                    int* d_frame_offset, const Pool& pool, cudaStream_t st);

struct DescribeTables {
	const double* d_ori_w;     // (2r+1)^2 or null
	const double* d_w;         // fast: (grid*sub)^2 ; mod: weight_sub (sub+2*overlap)^2
	const double* d_wgrid;     // mod: grid^2
	// [DESC_PERM_BINS][G*G] visiting order of the descriptor's sample grid for features whose orientation falls into the
	// bin (angle mod pi): samples sorted into thin bands of equal image row, left to right, so that the 32 lanes of a
	// warp read a few cache lines instead of one each.  The order only decides which lane fetches which sample.
	const uint16_t* d_perm;
	// sliding-window orientation: optional global scratch for the sorted samples of the first walk_cap features of a launch,
	// [walk_cap][n] records {w*dx, w*dy, angle, -} of 32 bytes.  With it the window walks run one per THREAD in their own
	// kernel (thousands in flight per SM) instead of one per warp inside describe_kernel; features beyond walk_cap, and
	// launches without scratch (the stage entry points), walk in describe_kernel.  One scratch per handle, i.e. per stream.
	double* d_walk;
	long walk_cap;
};
constexpr int DESC_PERM_BINS = 32;
// Orientation + descriptor for pool entries [*d_begin, *d_end) (device-side range, persistent warps loop over it).
// ii of frame f is at d_ii + (f - frame0) * W * H.  !do_orientation: use the angles already in pool.angle.
template <typename T>
void launch_describe(const T* d_ii, int W, int H, const DescribeParams& p, const DescribeTables& t, const Pool& pool,
                     const int* d_begin, const int* d_end, int frame0, bool do_orientation, bool do_describe, cudaStream_t st);

// Colour SURF (DescribePointSurfPlanar.describe :97-111): L2-normalise the `total_dof` concatenated band sums of every pool
// entry in [*d_begin, *d_end) and set its Laplacian sign from the grey integral image.
template <typename T>
void launch_planar_finish(const T* d_gray_ii, int W, int H, int total_dof, const Pool& pool, const int* d_begin, const int* d_end,
                          int frame0, cudaStream_t st);
// ConvertImage.average(Planar, gray): gray[f][y][x] = (sum over bands, left to right) / bands, in the pixel type's arithmetic
template <typename T>
void launch_average_bands(const T* d_in, long frame_stride, long band_stride, int stride, int bands, T* d_gray, int W, int H, int batch,
                          cudaStream_t st);

// ---------------------------------------------------------------- exact association (bs_assoc_exact.cu)
struct RowminScratch { double* p_min; int* p_first; int* p_last; long cap; };
// tensor-core candidate filter (bs_assoc_tc.cu)
bool tc_supported(int dof);
int tc_pad_rows(int n);
size_t tc_operand_bytes(int n);
size_t tc_norm_bytes(int n);
int tc_cand_cap();
void launch_tc_split(const double* d_x, int n, void* d_operand, float* d_norm, unsigned* d_max_norm_bits, cudaStream_t st);
void launch_tc_filter(const void* d_opA, const float* d_normA, int nA, const void* d_opB, const float* d_normB,
                      const unsigned* d_max_norm_b_bits, int nB, int* d_cand_count, int* d_cand, float* d_dbg, int sm_count,
                      cudaStream_t st, const double* d_row_bound = nullptr);
void launch_tc_row_bounds(const int32_t* d_pairs, const double* d_fit, int n, int m, double* d_bound, cudaStream_t st);
size_t tc_cand_bytes(size_t rows);
void launch_tc_rescore(const double* d_A, int nA, const double* d_B, const int* d_cand_count, const int* d_cand, const float* d_norm_a,
                       const unsigned* d_max_norm_b_bits, int mode, double max_fit,
                       int index_offset, int32_t* pairs, double* fit, double* o_min, int32_t* o_first, int32_t* o_last, int* d_overflow,
                       cudaStream_t st);
int rowmin_splits(int nA, int nB, int sm_count);
void launch_rowmin(const double* A, int nA, const double* B, int nB, int dof, int splits, const RowminScratch& s, cudaStream_t st);
void launch_rowmin_finalize(const RowminScratch& s, int nA, int splits, int mode, double max_fit, int index_offset, int32_t* pairs,
                            double* fit, double* o_min, int32_t* o_first, int32_t* o_last, cudaStream_t st);
void launch_validate(int n, int row_offset, int32_t* pairs, double* fit, const double* col_min, const int32_t* col_first,
                     const int32_t* col_last, cudaStream_t st);
void launch_merge_columns(int ranks, int m, const double* a_min, const int32_t* a_first, const int32_t* a_last, double* o_min,
                          int32_t* o_first, int32_t* o_last, cudaStream_t st, long stride_min = 0, long stride_idx = 0);

// AssociateSurfBasic helpers (bs_assoc_surf.cu)
void launch_partition_white(const uint8_t* d_white, int n, int32_t* d_idx, int32_t* d_count, cudaStream_t st);
void launch_gather_rows(const double* d_in, const int32_t* d_idx, int n, int dof, double* d_out, cudaStream_t st);
void launch_scatter_pairs(const int32_t* d_pairs_c, const double* d_fit_c, int n_c, const int32_t* d_src_idx, const int32_t* d_dst_idx,
                          int32_t* d_pairs, double* d_fit, cudaStream_t st);

const char* last_error_cstr();
long launch_count();
Layer make_layer(int size, int skip, int W, int H);

}  // namespace bs
