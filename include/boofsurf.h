/*
 * boofsurf.h — C ABI of libboofsurf.so: BoofCV's SURF detect/describe/associate path on one B200 (sm_100a).
 *
 * This is the drop-in boundary.  Every entry point names the reference interface it replaces; paths are
 * relative to the reference tree:
 *   F  = main/boofcv-feature/src/main/java/boofcv        IP = main/boofcv-ip/src/main/java/boofcv
 * A JNI / Panama binding for the two Java interfaces is shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain C, opaque handles, int status (BS_OK = 0).  No exception crosses the boundary; bs_last_error() returns the
 *     message of the last failure on the calling thread.
 *   - caller owns every input/output array.  The library owns device workspaces inside a handle and frees them in
 *     *_destroy.  Calls on ONE handle must be serialised by the caller (the Java objects are not thread safe either,
 *     F/abst/feature/detdesc/DetectDescribePoint.java:38-40); different handles are independent.
 *   - pointers named d_* are DEVICE pointers on the handle's device; all others are host pointers.
 *   - `stream` arguments are cudaStream_t passed as void*.  NULL = the handle's own (non-blocking) stream; to run on the
 *     CUDA default stream pass cudaStreamLegacy ((void*)0x1) or cudaStreamPerThread ((void*)0x2).
 *   - there is no CPU fallback: without a CUDA device every *_create fails with BS_ERR_CUDA.
 */
#ifndef BOOFSURF_H
#define BOOFSURF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BS_OK 0
#define BS_ERR_ARG 1      /* IllegalArgumentException on the Java side */
#define BS_ERR_CUDA 2     /* CUDA runtime failure (message has the cudaError string) */
#define BS_ERR_CAPACITY 3 /* an output pool was too small; counts are still valid, see bs_surf_total */
#define BS_ERR_STATE 4    /* e.g. associate() before setSource/setDestination */

/* pixel types accepted by detect (GIntegralImageOps.getIntegralType, IP/alg/transform/ii/GIntegralImageOps.java:35-46) */
#define BS_PIXEL_F32 0 /* GrayF32 -> GrayF32 integral */
#define BS_PIXEL_U8 1  /* GrayU8  -> GrayS32 integral */
#define BS_PIXEL_S32 2 /* GrayS32 -> GrayS32 integral */

/* ConfigFastHessian (F/abst/feature/detect/interest/ConfigFastHessian.java:33-70) */
typedef struct bs_fh_cfg {
	float detect_threshold;     /* detectThreshold = 1 */
	int extract_radius;         /* extractRadius = 2 */
	int max_features_per_scale; /* maxFeaturesPerScale = -1 */
	int initial_sample_size;    /* initialSampleSize = 1 */
	int initial_size;           /* initialSize = 9 */
	int scales_per_octave;      /* numberScalesPerOctave = 4 */
	int octaves;                /* numberOfOctaves = 4 */
	int scale_step_size;        /* scaleStepSize = 6 */
} bs_fh_cfg;

/* ConfigAverageIntegral / ConfigSlidingIntegral (F/abst/feature/orientation/ConfigAverageIntegral.java:34-51,
 * ConfigSlidingIntegral.java:34-54) */
typedef struct bs_ori_cfg {
	int variant;                   /* 0 = average gradient (surfFast), 1 = sliding window (surfStable) */
	double object_radius_to_scale; /* objectRadiusToScale = 1/2 */
	int radius;                    /* 6 (average) / 8 (sliding) */
	double sample_period;          /* samplePeriod = 1 / 0.65 */
	int sample_width;              /* sampleWidth = 6 */
	double weight_sigma;           /* weightSigma = -1 (derive from radius); 0 = unweighted */
	double window_size;            /* windowSize = PI/3 (sliding only) */
} bs_ori_cfg;

/* ConfigSurfDescribe.Speed / .Stability (F/abst/feature/describe/ConfigSurfDescribe.java:34-79) */
typedef struct bs_desc_cfg {
	int variant;             /* 0 = DescribePointSurf (Speed), 1 = DescribePointSurfMod (Stability) */
	int width_large_grid;    /* widthLargeGrid = 4 */
	int width_sub_region;    /* widthSubRegion = 5 */
	int width_sample;        /* widthSample = 3 */
	int use_haar;            /* useHaar = false; true: SparseIntegralHaar_NoBorder_F32/_I32 samples the descriptor (IP/alg/transform/ii/impl/SparseIntegralHaar_NoBorder_F32.java:37-73) */
	double weight_sigma;     /* Speed.weightSigma = 4.5 */
	int overlap;             /* Stability.overLap = 2 */
	double sigma_large_grid; /* Stability.sigmaLargeGrid = 2.5 */
	double sigma_sub_region; /* Stability.sigmaSubRegion = 2.5 */
} bs_desc_cfg;

/* defaults == passing null configs to FactoryDetectDescribe.surfFast / surfStable */
void bs_fh_cfg_default(bs_fh_cfg* c);
void bs_ori_cfg_default(bs_ori_cfg* c, int variant);
void bs_desc_cfg_default(bs_desc_cfg* c, int variant);

const char* bs_last_error(void);
const char* bs_version(void);
int bs_device_count(int* n);

/* ------------------------------------------------------------------------------------------------------------------
 * DetectDescribePoint<T,BrightFeature>: FactoryDetectDescribe.surfFast / surfStable
 * (F/factory/feature/detdesc/FactoryDetectDescribe.java:118-135, 209-226) -> WrapDetectDescribeSurf
 * (F/abst/feature/detdesc/WrapDetectDescribeSurf.java:75-141).
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct bs_surf bs_surf;

/* Constructor.  Validates the configs the way Configuration.checkValidity does (BS_ERR_ARG). */
int bs_surf_create(int device, const bs_fh_cfg* fh, const bs_ori_cfg* ori, const bs_desc_cfg* desc, int pixel_type,
                   bs_surf** out);
/* Colour SURF: FactoryDetectDescribe.surfColorFast / surfColorStable (F/factory/feature/detdesc/FactoryDetectDescribe.java:
 * 153-178, 247-268) = SurfPlanar_to_DetectDescribePoint (F/abst/feature/detdesc/SurfPlanar_to_DetectDescribePoint.java:40-74)
 * around DetectDescribeSurfPlanar (F/alg/feature/detdesc/DetectDescribeSurfPlanar.java:87-116) and DescribePointSurfPlanar
 * (F/alg/feature/describe/DescribePointSurfPlanar.java:97-111).  Features are detected and oriented on the band average;
 * the descriptor is the concatenation of the per-band sums, normalised once: bs_surf_dof() = 64 * num_bands.
 * pixel_type is the band type (Planar<GrayF32|GrayU8|GrayS32>). */
int bs_surf_create_planar(int device, const bs_fh_cfg* fh, const bs_ori_cfg* ori, const bs_desc_cfg* desc, int pixel_type,
                          int num_bands, bs_surf** out);
/* DetectDescribePoint<Planar<T>>.detect: bands[b] = Planar.bands[b].data (host), all with the same startIndex/stride.
 * A band count other than the handle's is an error (IllegalArgumentException in the reference, :59-61). */
int bs_surf_detect_planar(bs_surf* s, const void* const* bands, int num_bands, int start_index, int stride, int width,
                          int height);
/* Device-resident batch: band b of frame f starts at d_pixels + f*frame_stride + b*band_stride (elements). */
int bs_surf_detect_planar_device(bs_surf* s, const void* d_pixels, long frame_stride, long band_stride, int stride,
                                 int width, int height, int batch, void* stream);
int bs_surf_destroy(bs_surf* s);

/* Workspace tuning (optional, before the first detect).  chunk_frames = frames processed per kernel wave (bounds the
 * intensity workspace: 4*sum(layers) bytes per frame); pool_features = capacity of the feature pool of one detect call
 * (0 = batch * width*height/64). */
int bs_surf_set_workspace(bs_surf* s, int chunk_frames, long pool_features);
/* surfStable only: cap the sliding-window walk scratch (32 bytes x 289 samples per feature, default one chunk's features, at
 * most 6 GB) at max_features features; features beyond it are walked inside describe_kernel (same results, slower).  0 = default.
 * ImplOrientationSlidingWindowIntegral.java:139-188 is the walk either way; tests use this to run the in-kernel branch. */
int bs_surf_set_walk_scratch(bs_surf* s, long max_features);
/* Streams ("ways", 1..4, default 3) a device-resident batch is spread over, integral group by integral group; 1 = every
 * kernel on the caller's stream (what the per-stage timing of bs_surf_set_timing wants).  Results do not depend on it. */
int bs_surf_set_ways(bs_surf* s, int ways);

/* WrapDetectDescribeSurf.detect(T input) :75-96.  HOST image with BoofCV's {data,startIndex,stride,width,height}
 * (sub-images legal, main/boofcv-types/src/main/java/boofcv/struct/image/ImageBase.java:34-52).  Copies the pixels in,
 * runs the whole pipeline, returns when the results are readable.  Replaces any previous result. */
int bs_surf_detect(bs_surf* s, const void* pixels, int start_index, int stride, int width, int height);

/* Same pipeline on a DEVICE-resident batch: frame f starts at d_pixels + f*frame_stride (elements), rows `stride`
 * elements apart.  Asynchronous on `stream`; the first accessor call synchronises. */
int bs_surf_detect_device(bs_surf* s, const void* d_pixels, long frame_stride, int stride, int width, int height,
                          int batch, void* stream);

/* Batched detect+describe with HOST frames in and HOST feature lists out — what a JVM caller with a list of images uses.
 * Each image is what DetectDescribePoint.detect (DetectDescribePoint.java:32-46, WrapDetectDescribeSurf.java:75-96) would
 * receive; the results are what getLocation/getRadius(scale)/getOrientation/getDescription would return, frame after frame.
 * The batch is cut into sub-batches of 32 frames that alternate between two internal workers, each with its own stream:
 * the host->device copy and the device->host copy of one sub-batch overlap the kernels of the other.  Pinned host memory
 * (bs_host_register) makes those copies truly asynchronous; pageable memory works but serialises.
 *   pixels: frame f starts at element f*frame_stride, rows are `stride` elements apart.
 *   frame_offset[batch+1]: features of frame f are [frame_offset[f], frame_offset[f+1]).
 *   xy[cap][2], scale[cap], angle[cap], desc[cap][dof], white[cap]: any of them may be NULL (not copied).
 *   *total: features found.  BS_ERR_CAPACITY if total > cap (the first cap entries are valid). */
int bs_surf_detect_host_batch(bs_surf* s, const void* pixels, long frame_stride, int stride, int width, int height,
                              int batch, long cap, int32_t* frame_offset, double* xy, double* scale, double* angle,
                              double* desc, uint8_t* white, long* total);
/* cudaHostRegister / cudaHostUnregister for callers without a CUDA binding of their own (JNI direct buffers) */
int bs_host_register(void* p, size_t bytes);
int bs_host_unregister(void* p);

/* getNumberOfFeatures() :112-114 for one frame of the last detect; bs_surf_total = sum over frames (and the number the
 * pool would have needed when BS_ERR_CAPACITY was returned). */
int bs_surf_count(bs_surf* s, int frame, int* n);
int bs_surf_total(bs_surf* s, long* n);

/* getLocation/getRadius/getOrientation/getDescription :116-141 for all features of `frame`, in the reference's
 * single-thread emission order.  Any output pointer may be NULL.  radius = scale*BoofDefaults.SURF_SCALE_TO_RADIUS (=2).
 *   xy[2n] scale[n] angle[n] desc[dof*n] white[n]; dof = bs_surf_dof().
 *   meta[4n] = {octave, level, px, py} of the detection (for parity tests), response[n] = Fast-Hessian value. */
int bs_surf_read(bs_surf* s, int frame, double* xy, double* scale, double* angle, double* desc, uint8_t* white,
                 int32_t* meta, float* response);
int bs_surf_dof(const bs_surf* s); /* DescriptorInfo.createDescription().size() */

/* Per-stage device time of the detect calls since bs_surf_set_timing(s, 1): CUDA events recorded on the launching stream
 * around each stage of every chunk.  stage_ms[BS_N_STAGES] = accumulated milliseconds of
 * {integral, fast-hessian, nms+refine, compaction, orientation+describe}; stage_calls = how many chunks were timed. */
#define BS_N_STAGES 5
int bs_surf_set_timing(bs_surf* s, int on);
int bs_surf_get_timing(bs_surf* s, double* stage_ms, long* stage_calls);

/* Device views of the last result (valid until the next detect on this handle): feature i of frame f is at pool index
 * frame_offset[f] + i.  d_desc is [total][dof] doubles — feed it to bs_assoc_run_device without leaving the GPU. */
typedef struct bs_surf_device_view {
	const double* d_xy;      /* [total][2] */
	const double* d_scale;   /* [total] */
	const double* d_angle;   /* [total] */
	const double* d_desc;    /* [total][dof] */
	const uint8_t* d_white;  /* [total] */
	const int32_t* d_frame_offset; /* [batch+1] exclusive prefix of per-frame counts */
	long total;
	int batch;
	int dof;
} bs_surf_device_view;
int bs_surf_view(bs_surf* s, bs_surf_device_view* v);

/* Retained feature sets: a device-resident copy of ONE frame's features that outlives later detect calls on the handle.
 * The frame-at-a-time callers of the reference keep the previous frame's descriptors while the next frame is detected —
 * DetectDescribeAssociate / the DDA point tracker built by FactoryPointTracker.dda_FH_SURF_Fast/Stable
 * (main/boofcv-geo/src/main/java/boofcv/factory/feature/tracker/FactoryPointTracker.java:169-175, 204-210) swap a "source" and a
 * "destination" FastQueue<BrightFeature> every frame.  Here: detect frame t, bs_surf_retain it, detect frame t+1, retain it,
 * bs_assoc_features_device(previous, current) — nothing leaves HBM.  The copy is ~545 bytes per feature (device to device). */
typedef struct bs_features bs_features;
int bs_surf_retain(bs_surf* s, int frame, void* stream, bs_features** out);
int bs_features_count(const bs_features* f, int* n, int* dof);
/* device pointers of the set (batch = 1, total = n, d_frame_offset = {0, n}); valid until bs_features_destroy */
int bs_features_view(const bs_features* f, bs_surf_device_view* v);
int bs_features_destroy(bs_features* f);
/* (bs_assoc_features_device, declared with the associator below, associates two retained sets.) */

/* ------------------------------------------------------------------------------------------------------------------
 * Stage-level entry points (parity tests call these; each runs exactly the kernel the full pipeline uses).
 * Host pointers in, host pointers out.
 * ------------------------------------------------------------------------------------------------------------------ */
/* GIntegralImageOps.transform, IP/alg/transform/ii/GIntegralImageOps.java:56-70 -> ImplIntegralImageOps.transform
 * (IP/alg/transform/ii/impl/ImplIntegralImageOps.java:42-66 F32, :94-118 U8->S32, :120-144 S32).  out is dense
 * width*height (float for F32, int32 otherwise).  in_stride in elements. */
int bs_integral(int device, int pixel_type, const void* in, int in_stride, int width, int height, void* out);

/* NonMaxSuppression.process for the strict rule on maxima: NonMaxBlock.process (F/alg/feature/detect/extract/NonMaxBlock.java:
 * 69-94) with NonMaxBlockSearchStrict.Max (:56-79, :196-221) — what FactoryFeatureExtractor.nonmax(new ConfigExtract(radius,
 * threshold, ignoreBorder, true)) returns and the one plug-in hook the reference has on this path
 * (BOverrideFactoryFeatureExtractor.nonmax, F/factory/feature/detect/extract/FactoryFeatureExtractor.java:69-71).
 * intensity: host float image (stride in elements).  xy[2*max_out] receives (x, y) as int16 like QueueCorner, in the
 * reference's single-thread order (block rows, then blocks); *n = number found (BS_ERR_CAPACITY if > max_out). */
int bs_nonmax(int device, const float* intensity, int stride, int width, int height, int radius, float threshold_max,
              int ignore_border, int max_out, int16_t* xy, int* n);

/* IntegralImageFeatureIntensity.hessian (F/alg/feature/detect/intensity/IntegralImageFeatureIntensity.java:43-79).
 * ii dense width*height of integral type (ii_type: BS_PIXEL_F32 -> float, else int32); out dense (w/skip)*(h/skip). */
int bs_hessian(int device, int ii_type, const void* ii, int width, int height, int skip, int size, float* out);

/* The whole Fast-Hessian scale space of one image through the batch path's own kernels (FastHessianFeatureDetector's
 * intensity[] layers, F/alg/feature/detect/interest/FastHessianFeatureDetector.java:198-221, each layer =
 * IntegralImageFeatureIntensity.hessian(integral, skip, size, intensity)).  layers[5*k..] = {size, skip, w, h, offset} of
 * layer k (offset in floats into out); *n_layers / *n_floats are always set; out may be NULL to query the sizes, else it must
 * hold max_floats >= *n_floats floats. */
int bs_fh_intensity(int device, int ii_type, const void* ii, int width, int height, const bs_fh_cfg* cfg, int* n_layers,
                    long* layers, long* n_floats, float* out, long max_floats);
/* FastHessianFeatureDetector.detect on a given integral image
 * (F/alg/feature/detect/interest/FastHessianFeatureDetector.java:156-188).  Returns the number found in *n and fills
 * up to max_out entries: xys[3k] = {x, y, scale}, meta[4k] = {octave, level, px, py}, response[k]. */
int bs_fh_detect(int device, int ii_type, const void* ii, int width, int height, const bs_fh_cfg* cfg, int max_out,
                 int* n, double* xys, int32_t* meta, float* response);

/* OrientationIntegral.compute after setObjectRadius(2*scale) (WrapDetectDescribeSurf.java:98-110;
 * F/alg/feature/orientation/impl/ImplOrientationAverageGradientIntegral.java:54-127,
 * ImplOrientationSlidingWindowIntegral.java:81-188).  xys[3n] = {x, y, scale}. */
int bs_orientation(int device, int ii_type, const void* ii, int width, int height, const bs_ori_cfg* cfg,
                   const double* xys, int n, double* angle);

/* DescribePointSurf(.Mod).describe (F/alg/feature/describe/DescribePointSurf.java:169-213,
 * DescribePointSurfMod.java:121-192).  desc[dof*n], white[n]. */
int bs_describe(int device, int ii_type, const void* ii, int width, int height, const bs_desc_cfg* cfg,
                const double* xys, const double* angle, int n, double* desc, uint8_t* white);

/* ------------------------------------------------------------------------------------------------------------------
 * AssociateDescription<BrightFeature>: FactoryAssociation.greedy(ScoreAssociateEuclideanSq_F64, maxError, backwards)
 * (F/factory/feature/associate/FactoryAssociation.java:51-65) -> WrapAssociateGreedy
 * (F/abst/feature/associate/WrapAssociateGreedy.java:58-123) -> AssociateGreedy.associate
 * (F/alg/feature/associate/AssociateGreedy.java:65-118) with DescriptorDistance.euclideanSq
 * (F/alg/descriptor/DescriptorDistance.java:55-64).
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct bs_assoc bs_assoc;

int bs_assoc_create(int device, double max_fit_error, int backwards_validation, bs_assoc** out);
int bs_assoc_destroy(bs_assoc* a);
int bs_assoc_set_max_fit_error(bs_assoc* a, double max_fit_error); /* Associate.setMaxScoreThreshold */

/* Filter mode: 0 = exact fp64 distances for every pair; 1 = tensor-core candidate filter (tcgen05, fp16 operands scaled per row, fp16 accumulators)
 * followed by an exact fp64 re-score of the candidates (same pairs/fit, much faster for large N*M).  Default -1 = auto:
 * mode 1 when n*m >= 2^22 and dof == 64 (SURF-64), else mode 0.  Mode 1 synchronises the stream once per call (it has to
 * learn whether a candidate list overflowed, in which case the call is redone with the exact kernel). */
int bs_assoc_set_mode(bs_assoc* a, int mode);

/* Diagnostic for the tensor-core filter (mode 1): runs only the fp16 split and the tcgen05 pass for src against dst
 * and returns what the exact re-score would start from.  scores[n][m] (host, may be NULL) = the filter's fp32 estimate of
 * every squared distance; cand_count[n] (host, may be NULL) = candidates emitted per source row; stats[2] (may be NULL) =
 * {runs of the tensor-core path, runs that fell back to the exact kernel because a candidate list overflowed}.
 * Used by the tests to measure the estimate's real error against the bound the filter assumes. */
int bs_assoc_filter_scores(bs_assoc* a, const double* src, int n, const double* dst, int m, int dof, float* scores,
                           int32_t* cand_count, long* stats);

/* AssociateGreedy.associate: src[n][dof], dst[m][dof] row-major doubles (the TupleDesc_F64.value arrays back to back).
 * pairs[n] = index into dst or -1 (AssociateGreedyBase.getPairs), fit[n] = score of the match
 * (getFitQuality; Double.MAX_VALUE where backward validation rejected, the running best otherwise). */
int bs_assoc_run(bs_assoc* a, const double* src, int n, const double* dst, int m, int dof, int32_t* pairs,
                 double* fit);
/* Same on device pointers, asynchronous on `stream`; d_pairs/d_fit are device arrays. */
int bs_assoc_run_device(bs_assoc* a, const double* d_src, int n, const double* d_dst, int m, int dof,
                        int32_t* d_pairs, double* d_fit, void* stream);

/* Forward greedy association on device pointers that ALLOCATES NOTHING (no cudaMalloc / cudaFree, no handle): for callers
 * that cannot allocate — code under CUDA graph capture, tight serving loops, harnesses that forbid it.  This entry point
 * has no counterpart in the reference.
 * Results: exactly those of bs_assoc_run_device on a handle in exact mode without backward validation
 * (d_pairs[i] = the LAST index of d_dst attaining row i's minimum exact fp64 score if that score <= max_fit_error, else
 * -1; d_fit[i] = that score, else max_fit_error).  Asynchronous on `stream` (NULL = the legacy default stream).
 * The split partials of the tile kernel live in a file-scope device array of 4 MiB that every GPU loading the library
 * holds until exit; hence the limits: n <= BS_ASSOC_STATIC_MAX_ROWS source rows (any m, 1 <= dof <= 256), BS_ERR_ARG with
 * a message naming the limit beyond it.  Every call shares that array and it keeps the last call's contents (nothing in it
 * needs clearing: each call overwrites what it reads): calls that OVERLAP on the GPU race on it — issue them on one
 * stream or order them with events.  The first call on a device resolves the array's address and opts the kernel into its
 * shared-memory size (no allocation either); make it outside a graph capture. */
#define BS_ASSOC_STATIC_MAX_ROWS 131072
int bs_assoc_forward_static(int device, const double* d_src, int n, const double* d_dst, int m, int dof,
                            double max_fit_error, int32_t* d_pairs, double* d_fit, void* stream);

/* detect -> associate on the device: source = the features of frame_src, destination = those of frame_dst of the last
 * batch `s` detected (same device as `a`), i.e. the frame-to-frame use of the pair of interfaces by the reference's callers
 * (ExampleAssociatePoints.java:125-160, DetectDescribeAssociate): no descriptor leaves the GPU.  surf_basic != 0 splits by
 * the Laplacian sign first (AssociateSurfBasic).  d_pairs / d_fit hold bs_surf_count(s, frame_src) entries; the call is
 * ordered after the detect on the host (it waits for the batch) and runs on `stream`. */
int bs_assoc_frames_device(bs_assoc* a, bs_surf* s, int frame_src, int frame_dst, int surf_basic, int32_t* d_pairs,
                           double* d_fit, void* stream);

/* The same for a list of frame pairs of the last batch (a stereo stream: pair k = frames 2k, 2k+1), queued back to back on
 * `stream` with ONE host synchronisation at the end instead of one per pair.  offsets[n_pairs+1] (host, written by the
 * call): the results of pair k are d_pairs / d_fit [offsets[k], offsets[k+1]), offsets[k+1]-offsets[k] = the number of
 * features of frame_src[k].  cap = entries d_pairs / d_fit hold; BS_ERR_CAPACITY if the batch needs more.  Plain greedy
 * association only (AssociateGreedy); results are those of n_pairs bs_assoc_frames_device calls. */
int bs_assoc_frames_batch_device(bs_assoc* a, bs_surf* s, int n_pairs, const int32_t* frame_src, const int32_t* frame_dst,
                                 int32_t* d_pairs, double* d_fit, long cap, long* offsets, void* stream);

/* AssociateDescription.associate on two retained sets (surf_basic != 0: WrapAssociateSurfBasic's split by Laplacian sign):
 * d_pairs[n_src] / d_fit[n_src] as in bs_assoc_run_device. */
int bs_assoc_features_device(bs_assoc* a, const bs_features* src, const bs_features* dst, int surf_basic, int32_t* d_pairs,
                             double* d_fit, void* stream);

/* Sharded association: the src rows of one rank against ALL dst rows.
 * Step 1 (bs_assoc_forward_device): forward pass for the local rows + per-column partial minima over the local rows:
 *   d_col_min[m] (score), d_col_first[m], d_col_last[m] (GLOBAL row index = row_offset + local row; -1 if none).
 * The caller combines the partials of all ranks (min of col_min; first = min / last = max over the ranks attaining it)
 * with NCCL all-gather + bs_assoc_merge_columns_device, then
 * Step 2 (bs_assoc_validate_device): applies AssociateGreedy.java:98-117 to the local rows. */
int bs_assoc_forward_device(bs_assoc* a, const double* d_src, int n, int row_offset, const double* d_dst, int m,
                            int dof, int32_t* d_pairs, double* d_fit, double* d_col_min, int32_t* d_col_first,
                            int32_t* d_col_last, void* stream);
/* The forward pass alone (AssociateGreedy.java:76-96, no backward validation whatever the handle was created with): what a
 * rank of a sharded association runs when backwardsValidation is off — no column statistics, no exchange step. */
int bs_assoc_forward_only_device(bs_assoc* a, const double* d_src, int n, const double* d_dst, int m, int dof,
                                 int32_t* d_pairs, double* d_fit, void* stream);
int bs_assoc_merge_columns_device(bs_assoc* a, int ranks, int m, const double* d_all_min, const int32_t* d_all_first,
                                  const int32_t* d_all_last, double* d_col_min, int32_t* d_col_first,
                                  int32_t* d_col_last, void* stream);
/* The same merge for ONE packed record per rank, which is what a single all-gather moves: rank r's record starts at
 * d_packed + r*16*m bytes and holds col_min[m] (double), then col_first[m], then col_last[m] (int32) — i.e. the three outputs
 * of bs_assoc_forward_device written into one 16*m-byte buffer. */
int bs_assoc_merge_packed_device(bs_assoc* a, int ranks, int m, const void* d_packed, double* d_col_min,
                                 int32_t* d_col_first, int32_t* d_col_last, void* stream);
int bs_assoc_validate_device(bs_assoc* a, int n, int row_offset, int32_t* d_pairs, double* d_fit,
                             const double* d_col_min, const int32_t* d_col_first, const int32_t* d_col_last,
                             void* stream);

/* WrapAssociateGreedy.associate :85-92 + FindUnassociated.checkDestination
 * (F/alg/feature/associate/FindUnassociated.java:56-72): turn pairs/fit into the match list (ascending src) and the two
 * unassociated index lists.  Host-side helper (pure index bookkeeping).  Outputs sized n, n, n, n, m; counts returned. */
int bs_assoc_matches(const int32_t* pairs, const double* fit, int n, int m, int32_t* match_src, int32_t* match_dst,
                     double* match_score, int* n_matches, int32_t* unassoc_src, int* n_unassoc_src,
                     int32_t* unassoc_dst, int* n_unassoc_dst);

/* AssociateSurfBasic (F/alg/feature/associate/AssociateSurfBasic.java:83-143) behind WrapAssociateSurfBasic
 * (F/abst/feature/associate/WrapAssociateSurfBasic.java:35-90), the associator FactoryPointTracker.dda_FH_SURF_Fast/Stable
 * build around FactoryAssociation.greedy: BrightFeature lists are split by `white` (sign of the Laplacian), each class is
 * associated with the handle's greedy settings, indices refer to the callers' lists.
 * src_white[n] / dst_white[m] = BrightFeature.white (0/1).  pairs[n] = index into dst or -1, fit[n] as bs_assoc_run. */
int bs_assoc_surf_run(bs_assoc* a, const double* src, const uint8_t* src_white, int n, const double* dst,
                      const uint8_t* dst_white, int m, int dof, int32_t* pairs, double* fit);
/* Same on device pointers (e.g. straight from bs_surf_view); synchronises `stream` once (class sizes). */
int bs_assoc_surf_run_device(bs_assoc* a, const double* d_src, const uint8_t* d_src_white, int n, const double* d_dst,
                             const uint8_t* d_dst_white, int m, int dof, int32_t* d_pairs, double* d_fit, void* stream);
/* Match list in AssociateSurfBasic's order (white class in ascending source index, then the other class), its
 * unassociated-source list (same class order; empty when either input list is empty, :89-92) and
 * FindUnassociated.checkDestination over all m destinations.  Host-side index bookkeeping; outputs sized n,n,n,n,m. */
int bs_assoc_surf_matches(const int32_t* pairs, const double* fit, const uint8_t* src_white, int n, int m,
                          int32_t* match_src, int32_t* match_dst, double* match_score, int* n_matches,
                          int32_t* unassoc_src, int* n_unassoc_src, int32_t* unassoc_dst, int* n_unassoc_dst);

/* kernel launch counter (all kernels this library launched on the calling process since load) — bench.py's
 * gpu_launches */
long bs_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* BOOFSURF_H */
