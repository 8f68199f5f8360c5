/*
 * mvstereo.h -- C ABI of libmvstereo: the B200 (sm_100a) implementation of multiview's per-frame
 * stereo reconstruction hot path (rectify -> SGM disparity -> point cloud -> floor histogram).
 *
 * The reference has no FFI for this path; the operator boundary is the four static call sites that
 * its pipeline nodes make.  Each entry point below names the reference
 * interface it replaces.  Paths are relative to multiview/multiview_cpp/src/perceive/.
 *
 * Conventions
 *   - every function returns 0 on success, a negative mvs_status otherwise; nothing throws or aborts
 *     (the reference's nodes are noexcept and return nullptr on failure, pipeline-task.hpp:266-272);
 *   - a context owns one CUDA stream and a grow-only device arena; contexts are independent, so the
 *     pipeline's pool threads use one context each (execute() "MUST BE REENTRANT", pipeline-task.hpp:292-295);
 *     a single context must not be used from two threads at once;
 *   - pointers are HOST pointers unless the name ends in _dev; steps are in bytes (cv::Mat::step);
 *   - there is no CPU fallback: without a CUDA device mvs_ctx_create fails.
 */
#ifndef MVSTEREO_H
#define MVSTEREO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mvs_ctx mvs_ctx;

typedef enum {
    MVS_OK = 0,
    MVS_ERR_ARG = -1,         /* bad argument (null pointer, non-positive size, unsupported enum) */
    MVS_ERR_CUDA = -2,        /* a CUDA runtime call failed; see mvs_last_error */
    MVS_ERR_UNSUPPORTED = -3, /* parameters outside the supported range (see mvs_sgbm_check) */
    MVS_ERR_NOMEM = -4
} mvs_status;

/* cv::INTER_* values accepted by mvs_remap_bgr8 (calc_rectified.hpp:17-22). */
enum { MVS_INTER_NEAREST = 0, MVS_INTER_LINEAR = 1, MVS_INTER_CUBIC = 2, MVS_INTER_LANCZOS4 = 4 };

/* cv::StereoSGBM modes (stereosgbm_disp.cpp:55-80): 0 = 5-path, 1 = 8-path full DP, 3 = 4-path HH4. */
enum { MVS_MODE_SGBM = 0, MVS_MODE_HH = 1, MVS_MODE_HH4 = 3 };

/* FloorHistogram::WeightingMethod (cost-functions/floor-histogram.hpp:20). */
enum { MVS_WGT_NONE = 0, MVS_WGT_NORM = 1, MVS_WGT_QUADRANCE = 2 };

/* -- context ------------------------------------------------------------------------------------ */
int mvs_ctx_create(int device, mvs_ctx** out);
void mvs_ctx_destroy(mvs_ctx* ctx);
const char* mvs_last_error(const mvs_ctx* ctx); /* never NULL */
int mvs_sync(mvs_ctx* ctx);                     /* wait for the context's stream */
void* mvs_stream(mvs_ctx* ctx);                 /* the main cudaStream_t (stage entry points, first half of the frame path) */
void* mvs_tail_stream(mvs_ctx* ctx);            /* the cudaStream_t the frame path finishes a batch on (left-right check ..
                                                   histogram, the histogram sums): work that consumes frame results is
                                                   enqueued here */
const char* mvs_version(void);
long long mvs_launch_count(const mvs_ctx* ctx); /* kernels launched through this context so far */

/* -- a1: rectification -----------------------------------------------------------------------------
 * Replaces cv::remap(in, rectified, mapx, mapy, interp, BORDER_CONSTANT, Scalar(255,255,255)) in
 * calc_rectified::Task::execute, pipeline/nodes/calc_rectified.cpp:109-115.
 * src: BGR 8UC3 sw x sh; mapx/mapy: CV_32FC1 w x h; dst: BGR 8UC3 w x h.  Bit-exact with OpenCV's fixed-point path for
 * the four modes the node offers (calc_rectified.hpp:17-22): INTER_NEAREST, INTER_LINEAR, INTER_CUBIC (the default,
 * calc_rectified.hpp:34; the tuned kernel) and INTER_LANCZOS4; any other value returns MVS_ERR_UNSUPPORTED. */
int mvs_remap_bgr8(mvs_ctx* ctx, const uint8_t* src, int sw, int sh, size_t sstep, const float* mapx,
                   const float* mapy, size_t mstep, int w, int h, int interp, const uint8_t border_bgr[3],
                   uint8_t* dst, size_t dstep);

/* The per-video half of get_xy_mapping (pipeline/nodes/get_xy_mapping.cpp:59, utils/create-cv-remap.cpp:10-68
 * produce mapx / mapy once per video): register a sensor's maps once and the INTER_CUBIC remap of the frame path reads
 * 6 bytes per pixel of precomputed fixed-point coordinates instead of converting 8 bytes of float maps in every frame.
 * mapx / mapy: CV_32FC1 w x h, byte step mstep, host (on_device = 0) or device pointers; they are not needed after the
 * call.  *handle (> 0) goes into mvs_frame_desc.maps_handle[i].  Results are identical (the same conversion, done once). */
int mvs_maps_register(mvs_ctx* ctx, const float* mapx, const float* mapy, size_t mstep, int w, int h, int on_device, int* handle);
int mvs_maps_release(mvs_ctx* ctx, int handle);

/* -- a2: grey + pre-smoothing ----------------------------------------------------------------------
 * Replaces the prologue of Disparity::calculate, cost-functions/disparity/disparity.cpp:106-155:
 * cv::cvtColor(BGR2GRAY) when channels == 3, then cv::GaussianBlur(31x31, sigma) applied TWICE when
 * sigma > 0.  dst is 8UC1 w x h.  Bit-exact. */
int mvs_presmooth(mvs_ctx* ctx, const uint8_t* src, int channels, int w, int h, size_t step, double sigma,
                  uint8_t* dst, size_t dstep);

/* -- a3/a4: semi-global matching -------------------------------------------------------------------
 * Field order and meaning follow StereoSGBMDisparity::Params,
 * cost-functions/disparity/strategies/stereosgbm_disp.hpp:23-33 (sg_SAD_window_size, sg_12_max_diff,
 * sg_min_disparity, sg_num_disparities, sg_speckle_window_size, sg_speckle_range, sg_prefilter_cap,
 * sg_uniqueness_ratio, sg_mode, sg_p1, sg_p2). */
typedef struct {
    int block, disp12_max_diff, min_disp, num_disp, speckle_win, speckle_range, prefilter_cap, uniqueness,
        mode, p1, p2;
} mvs_sgbm_params;

/* Returns MVS_OK when the parameter set is inside the bit-exact domain for a w-pixel-wide image:
 * num_disp a positive multiple of 16 and <= 256, min_disp >= 0, odd block in [1, 13] such that
 * block^2 * (2*ftzero + 63) + P2 <= 32767 (16-bit costs cannot wrap), mode in {0, 1, 3}. */
int mvs_sgbm_check(const mvs_sgbm_params* p, int w, int h);

/* The reference pins OpenCV 4.4.0 (build/install-files/build-scripts/build-opencv.sh:29); the OpenCV importable where this
 * library was validated is 4.13.0, whose path recurrence subtracts min(Lp) where older sources subtract min(Lp) + P2.
 * subtract_p2 = 0 (default): the 4.13 form, bit-exact with cv2 4.13.  subtract_p2 = 1: the older form: every path cost is
 * lower by P2, which changes only the uniqueness test; bit-exact with the CPU restatement built with
 * -DORC_SGBM_SUBTRACT_P2 wherever no 16-bit cost sum saturates (a saturated sum makes the next synchronising call return
 * MVS_ERR_UNSUPPORTED), and limited to directions * P2 <= 32767.  Applies to all later SGBM calls of the context. */
int mvs_set_sgbm_recurrence(mvs_ctx* ctx, int subtract_p2);

/* Replaces cv::StereoSGBM::compute(left, right, disp16S) as called from
 * StereoSGBMDisparity::calculate, strategies/stereosgbm_disp.cpp:121-125 (left is the reference image).
 * l, r: 8UC1 w x h; disp16: CV_16S w x h, Q4 fixed point, invalid = (min_disp - 1) * 16.  Bit-exact. */
int mvs_sgbm_compute16(mvs_ctx* ctx, const mvs_sgbm_params* p, const uint8_t* l, const uint8_t* r, int w, int h,
                       size_t step, int16_t* disp16, size_t dstep);

/* Replaces Disparity::calculate(p, l_is_ref = true, STEREO_SGBM, image0, image1, out, confidence),
 * cost-functions/disparity/disparity.cpp:68-182 with strategies/stereosgbm_disp.cpp:84-156:
 * grey + double blur (a2), StereoSGBM (a3), then out = |d16| / 16, values > 150 -> 0 (:141-150).
 * l, r: 8UC1 or 8UC3 (channels = 1 | 3); out: CV_32FC1 w x h.  Bit-exact. */
int mvs_disparity_sgbm(mvs_ctx* ctx, const mvs_sgbm_params* p, double presmooth_sigma, const uint8_t* l,
                       const uint8_t* r, int channels, int w, int h, size_t step, float* out, size_t ostep);

/* -- a5: the CUDA_SGM strategy ---------------------------------------------------------------------------
 * Replaces SGMDisparity::calculate(p, l_is_ref = true, image0, image1, out), strategies/sgm_disp.cpp:102-211, i.e. libSGM's
 * sgm::StereoSGM(w, h, disparity_size, depth_bits, 8 | 16).execute plus the float conversion (:195-207): census 9x7,
 * Hamming cost, 8 paths with P1 = 20 / P2 = 100, left / right winner with uniqueness 0.95, 3x3 median, left-right check;
 * disparities are stored + 1, 0 = invalid, values > 512 -> 0.  l, r: grey images of depth_bits 8 or 16 (use_16bit_greyscale,
 * sgm_disp.hpp:19-20), byte step `step`; disparity_size 64, 128 or 256; w and h even; out: CV_32FC1.  disp_tap (may be
 * NULL): the 16-bit disparity image before the float conversion.  Bit-exact with the CPU restatement of the vendored libSGM
 * kernels (oracle/census_oracle.c, which lists where those kernels' border artefacts are not reproduced); the restatement is
 * pinned against those kernels themselves, compiled from the reference tree and run on the GPU, at disparity sizes 64 and 128
 * (tests/test_gpu_census_ref.py); at 256 -- un-vendored libSGM 2.x in the reference -- nothing runnable exists to pin it. */
int mvs_disparity_census(mvs_ctx* ctx, const void* l, const void* r, int depth_bits, int w, int h, size_t step,
                         int disparity_size, float* out, size_t ostep, uint16_t* disp_tap);

/* -- a8: merge with the scene's reference disparity ------------------------------------------------------
 * Replaces the loop of disp_map_update::Task::execute, pipeline/nodes/disp_map_update.cpp:98-113: in place,
 * disp[x] = ref[x] where ref[x] is finite, labels[x] >= 0 and (the super-pixel is not flagged still_cuttoff, or
 * disp[x] is not finite, or disp[x] <= 0).  disp, ref: CV_32FC1 w x h; labels: int32 w x h (the SLIC lookup);
 * still_cutoff: one byte per super-pixel (ImageFeatures2d::slic_info[].still_cuttoff); labels >= n_labels are
 * skipped.  The caller then rebuilds the cloud with mvs_point_cloud (disp_map_update.cpp:115-119).  Bit-exact. */
int mvs_disp_merge_ref(mvs_ctx* ctx, float* disp, size_t dstep, const float* ref, size_t rstep, const int32_t* labels,
                       size_t lstep, const uint8_t* still_cutoff, int n_labels, int w, int h);

/* -- a6/a7: point cloud -----------------------------------------------------------------------------
 * The 24 doubles of BinocularCamera that the device needs: KRR_inv_[2] (row-major 3x3) and C_[2],
 * geometry/projective/binocular-camera.hpp (members filled by binocular-camera.cpp:682-785). */
typedef struct {
    double krr_inv[2][9];
    double c[2][3];
} mvs_bcam;

/* Replaces PointCloud::init(disparity, left_is_reference, bino_cam), geometry/point-cloud.cpp:10-72
 * (solve3d: geometry/projective/binocular-camera.cpp:908-927).  disp: CV_32FC1 w x h.
 * Xs: capacity w*h*3 doubles (std::vector<Vector3>), xy: capacity w*h*2 int32 (std::vector<Point2>),
 * lookup: w*h int32, row stride w (IntImage).  *n_out receives N.  xy/lookup/N bit-exact; Xs and
 * centroid within 1e-4 relative (fp64 evaluated without FMA, centroid summed in a different order). */
int mvs_point_cloud(mvs_ctx* ctx, const mvs_bcam* bcam, const float* disp, size_t step, int w, int h,
                    int left_is_ref, double* Xs, int32_t* xy, int32_t* lookup, int32_t* n_out, double centroid[3]);

/* The same point cloud for callers that must not allocate at frame time.  This entry point has no
 * counterpart in the reference; it exists for hosts where the library may not call cudaMalloc once frames
 * are flowing: a real-time camera loop (cudaMalloc synchronises the whole device and would stall the other
 * streams), or an application whose own pool already owns all GPU memory (cudaMalloc would fail).
 *   - All pointers are DEVICE pointers supplied by the caller: disp_dev (w*h floats, tightly packed), Xs_dev
 *     (capacity w*h*3 doubles), xy_dev (capacity w*h*2 int32), lookup_dev (w*h int32), n_dev (one int32),
 *     centroid_dev (3 doubles).  The call is asynchronous on the context's MAIN stream, mvs_stream() (mvs_sync to
 *     wait).  That stream is not ordered behind the frame path's second half: a disparity taken from
 *     mvs_frame_result.disp_dev is produced on mvs_tail_stream(), so make the main stream wait for it first
 *     (an event recorded on the tail stream, or mvs_sync).
 *   - It performs no cudaMalloc/cudaFree and calls nothing that allocates.  Its compaction scratch is a fixed
 *     288 KiB device array inside the library, resident on every GPU that loads the library until exit.
 *   - Limit: w*h <= MVS_STATIC_CLOUD_MAX_PIXELS (4096 x 2304); larger images return MVS_ERR_UNSUPPORTED and
 *     mvs_last_error says so -- use mvs_point_cloud for those.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress) and, with CUDA's
 *     default lazy module loading, is also the call that loads the four point-cloud kernels onto the device (a
 *     driver-side allocation outside this library's control); make that first call before starting any stream
 *     capture on that device, and before the phase in which nothing may allocate.
 *   - All calls on a device share the one array.  Its contents are whatever the previous call left; a call
 *     overwrites every entry before reading it, so nothing needs clearing, but two calls whose kernels overlap
 *     on the GPU (two contexts / streams on the same device) race on it: serialise them. */
#define MVS_STATIC_CLOUD_MAX_PIXELS (4096 * 2304)
int mvs_point_cloud_static_dev(mvs_ctx* ctx, const mvs_bcam* bcam, const float* disp_dev, int w, int h,
                               int left_is_ref, double* Xs_dev, int32_t* xy_dev, int32_t* lookup_dev,
                               int32_t* n_dev, double* centroid_dev);

/* -- a9/a10: floor histogram ------------------------------------------------------------------------
 * Subset of FloorHistogram::Params used per point, cost-functions/floor-histogram.hpp:28-36. */
typedef struct {
    float hist_sz, min_z, max_z;
    int weighting; /* MVS_WGT_* */
    int use_still; /* apply_slic_still_weights */
    int ordered;   /* 1: sum every bin in point order like the reference's sequential loop (weighted histograms bit-exact;
                      a stable sort by bin, slower); 0: float atomics (counts exact, weighted sums <= 1e-5 relative) */
    float super_pixel_min_z; /* FloorHistogram::Params::super_pixel_min_z (floor-histogram.hpp:39, default 0.25f): only
                                used by the super-pixel table */
} mvs_hist_params;

/* SparseHistCell, geometry/sparse-hist-cell.hpp:8-17: histogram cell of a super-pixel's centroid and its "count". */
typedef struct {
    int32_t x, y;  /* {-1, -1} when no point of the super-pixel landed in the histogram */
    float count;
} mvs_sparse_hist_cell;

/* ensure_floor_hist_dims, cost-functions/floor-histogram.cpp:19-35.  aabb = {left, top, right, bottom}. */
int mvs_hist_dims(const double aabb[4], float hist_sz, int* hw, int* hh);

/* Replaces process_camera_i inside FloorHistogram::calculate, cost-functions/floor-histogram.cpp:197-275
 * (histogram part; mvs_floor_hist_camera_cells adds the super-pixel centroid table).
 * et = {tx,ty,tz, qx,qy,qz,qw, scale} of EuclideanTransform (geometry/euclidean-transform.hpp:121-124);
 * cam_center = translations[cam] (pipeline/nodes/floor_hist.cpp:122); labels (int32 lw x lh, row stride lw)
 * and still_score (n_labels doubles) may be NULL.  cam_hist: float hw x hh, row stride hw, overwritten.
 * Counts (MVS_WGT_NONE without still weights) are bit-exact; weighted sums are float accumulations in a
 * different order (<= 1e-5 relative) unless p->ordered is set, which reproduces the reference's point order. */
int mvs_floor_hist_camera(mvs_ctx* ctx, const mvs_hist_params* p, const double et[8], const double aabb[4],
                          const double cam_center[3], const double* Xs, const int32_t* xy, int32_t n,
                          const int32_t* labels, int lw, int lh, const double* still_score, int n_labels,
                          float* cam_hist, int hw, int hh);

/* The same, plus camera_info[cam].slic_to_hist_xy (floor-histogram.cpp:222-225,269-286): per super-pixel, in point order,
 * sC += Vector4f(to_vec3f(Y) * float(weight), sC.w + weight) over the points that landed in the histogram with
 * label >= 0 and Y.z >= super_pixel_min_z -- exactly as written there, so the fourth component roughly doubles at every
 * point and overflows to +inf beyond ~128 points (the centroid then comes out as the world origin and count as inf) --
 * then xy = round(project_to_histogram(sC.xyz / sC.w)), count = round(sC.w).  slic_to_hist_xy: n_labels cells, bit-exact
 * with the CPU restatement (labels >= n_labels are skipped; the reference would index past its vector).  NULL skips it. */
int mvs_floor_hist_camera_cells(mvs_ctx* ctx, const mvs_hist_params* p, const double et[8], const double aabb[4],
                                const double cam_center[3], const double* Xs, const int32_t* xy, int32_t n,
                                const int32_t* labels, int lw, int lh, const double* still_score, int n_labels,
                                float* cam_hist, int hw, int hh, mvs_sparse_hist_cell* slic_to_hist_xy);

/* Replaces `hist += camera_info[cam].hist` over cameras, cost-functions/floor-histogram.cpp:296-300.
 * cam_hists: n_cams consecutive hw x hh float images, summed in camera order. */
int mvs_hist_sum(mvs_ctx* ctx, const float* cam_hists, int n_cams, int hw, int hh, float* hist);

/* Replaces the tail of FloorHistogram::calculate, cost-functions/floor-histogram.cpp:411-434: in place,
 * cv::GaussianBlur(CV_32F, ksize = ceil(sigma)*4 | 1, sigma, BORDER_DEFAULT) when gaussian_sigma is finite and > 0,
 * then clip (values < min_hist_count -> 0, values > max_hist_count -> max_hist_count).  Applied to the summed
 * histogram (after the cross-GPU sum).  Float arithmetic: within 1e-6 relative of OpenCV (whose own result varies by
 * an ulp with its SIMD dispatch).  mvs_hist_post takes a host image, mvs_hist_post_dev a device image (asynchronous, on
 * mvs_tail_stream()).  Limit: gaussian_sigma <= 8 (a 33-tap kernel); larger values return MVS_ERR_UNSUPPORTED (the
 * reference accepts any finite sigma; its default is 0.5). */
int mvs_hist_post(mvs_ctx* ctx, float* hist, int hw, int hh, float gaussian_sigma, float min_hist_count, float max_hist_count);
int mvs_hist_post_dev(mvs_ctx* ctx, float* hist_dev, int hw, int hh, float gaussian_sigma, float min_hist_count,
                      float max_hist_count);

/* -- the step upstream: input_images::Task::execute, pipeline/nodes/input_images.cpp:65-76:
 *   equal = raw_image_equalize ? hist_equalize(raw) : raw     (graphics/hist-equalize.cpp:9-30: BGR -> YCrCb,
 *                                                               cv::equalizeHist on Y, YCrCb -> BGR)
 *   grey  = cvtColor(equal, BGR2GRAY)
 * raw: BGR 8UC3 w x h.  equal (8UC3) and grey (8UC1) may each be NULL.  Bit-exact with OpenCV.  The _dev variant takes
 * device images (so a decoded frame can go straight on to mvs_frame_run with inputs_on_device) and is asynchronous. */
int mvs_input_images(mvs_ctx* ctx, const uint8_t* raw, int w, int h, size_t step, int raw_image_equalize, uint8_t* equal,
                     size_t estep, uint8_t* grey, size_t gstep);
int mvs_input_images_dev(mvs_ctx* ctx, const uint8_t* raw_dev, int w, int h, size_t step, int raw_image_equalize,
                         uint8_t* equal_dev, size_t estep, uint8_t* grey_dev, size_t gstep);

/* -- next step downstream: the front half of LocalizationData::calculate,
 * cost-functions/localization/localization-data.cpp:845-850 with the pose-less branch :855-858: from the floor histogram
 *   motion = hist > n_deviations ? hist : 0                                   (make_motion_histogram :201-233)
 *   rhist  = disc-weighted average of motion, radius person_diameter / 2      (make_radius_filtered_hist :147-187)
 *   bg     = min(prob_bg_max, phi(-(rhist - background_median) / background_stddev))   (:237-253)
 *   fg0    = phi((rhist - track_median) / track_stddev)                        (:264-285)
 *   fg     = max(fg0, 0.1)^2                                                   (:298-315 with fg1 = fg0)
 * Field names and defaults as LocalizationData::Params (localization-data.hpp:27-39); use_median is ignored as in the
 * reference (:213-219).  Results within 1e-6 relative of the reference arithmetic (fp64 erfc differs by <= 1 ulp).
 * Output pointers of mvs_localization_hists may be NULL; the _dev variant takes device images and is asynchronous. */
typedef struct {
    double n_deviations, person_diameter, background_median, background_stddev, track_median, track_stddev, prob_bg_max;
} mvs_loc_params;
int mvs_localization_hists(mvs_ctx* ctx, const mvs_loc_params* p, const float* hist, int hw, int hh, double hist_sz,
                           float* motion, float* rhist, float* bg, float* fg0, float* fg);
int mvs_localization_hists_dev(mvs_ctx* ctx, const mvs_loc_params* p, const float* hist_dev, int hw, int hh, double hist_sz,
                               float* motion_dev, float* rhist_dev, float* bg_dev, float* fg0_dev, float* fg_dev);

/* -- device-resident frame path ----------------------------------------------------------------------
 * One camera-frame through a1 -> a10 without leaving the GPU: what calc_rectified[2c],[2c+1] ->
 * disp_init[c] -> disp_map_update[c] (pass-through, disp_map_update.cpp:89-93) -> floor_hist compute
 * for one camera (pipeline/frame-results.cpp:206-255).  Inputs are host or device pointers (flag);
 * results stay on the device inside the context until fetched. */
typedef struct {
    /* sensor images: BGR 8UC3 sw x sh, byte step sstep; rectification maps CV_32FC1 w x h, byte step mstep */
    const uint8_t* sensor[2];
    const float* mapx[2];
    const float* mapy[2];
    int sw, sh, w, h;
    size_t sstep, mstep;
    int inputs_on_device; /* sensor images: 0 = host pointers (copied in), 1 = device pointers (used in place) */
    int maps_on_device;   /* mapx/mapy: same meaning; the maps are per-camera constants (get_xy_mapping.cpp:59 computes
                             them once per video), so a caller uploads them once and passes device pointers */
    mvs_sgbm_params sgbm;
    double presmooth_sigma;
    mvs_bcam bcam;
    mvs_hist_params hist;
    double et[8];
    double aabb[4];
    double cam_center[3];
    /* Per-call knobs of the nodes (set them: a zeroed struct means NEAREST / right-is-reference):
     * interp = calc_rectified::Params::interpolation_method (calc_rectified.hpp:17-22,34; the default is MVS_INTER_CUBIC),
     * left_is_ref = the argument of PointCloud::init (geometry/point-cloud.cpp:48-58; the pipeline passes true). */
    int interp;
    int left_is_ref;
    /* Inputs of the reference's DEFAULT histogram weighting (apply_slic_still_weights, floor-histogram.hpp:34-35,
     * floor-histogram.cpp:239-251): the camera's SLIC label image (int32 w x h, byte step lstep; create_slic_lookup's result)
     * and ImageFeatures2d::slic_info[].still_score (n_labels doubles).  NULL / 0 when hist.use_still is 0. */
    const int32_t* labels;
    size_t lstep;
    const double* still_score;
    int n_labels;
    int labels_on_device; /* labels and still_score: 0 = host pointers (copied in), 1 = device pointers */
    int maps_handle[2];   /* > 0: the sensor's maps were registered with mvs_maps_register; mapx / mapy of that sensor are
                             then ignored (may be NULL).  Used with interp = MVS_INTER_CUBIC, w x h must match. */
} mvs_frame_desc;

/* Device-side results of the last mvs_frame_run on this context slot (valid until the next run on it). */
typedef struct {
    const uint8_t* rectified_dev[2]; /* BGR 8UC3 w x h, step 3*w */
    const int16_t* disp16_dev;       /* CV_16S w x h */
    const float* disp_dev;           /* CV_32FC1 w x h */
    const double* Xs_dev;            /* N x 3 */
    const int32_t* xy_dev;           /* N x 2 */
    const int32_t* lookup_dev;       /* w x h */
    const float* cam_hist_dev;       /* hw x hh */
    const int32_t* n_dev;            /* N (device scalar) */
    const double* centroid_dev;      /* 3 doubles */
    int hw, hh;
    const mvs_sparse_hist_cell* slic_to_hist_xy_dev; /* n_labels cells (NULL when the frame had no label image) */
    int n_labels;
} mvs_frame_result;

/* Enqueue the whole path for one camera-frame on the context's stream (asynchronous).  `slot` selects
 * one of the context's result sets (0 <= slot < 64) so that several camera-frames can be in flight. */
int mvs_frame_run(mvs_ctx* ctx, int slot, const mvs_frame_desc* desc, mvs_frame_result* out);

/* The camera-frames of one step together (1 <= n_frames <= 32, slots first_slot .. first_slot+n_frames-1; all
 * frames must share image size, SGBM parameters and histogram bounds).  Equivalent to n_frames calls of
 * mvs_frame_run, but the SGM aggregation sweeps the whole batch in one launch per direction set, which is what
 * fills the GPU -- the reference's thread pool likewise runs the nodes of different cameras concurrently
 * (pipeline/pipeline-task.cpp:181-200).  descs / out are arrays of n_frames elements (out may be NULL). */
int mvs_frame_run_batch(mvs_ctx* ctx, int first_slot, const mvs_frame_desc* descs, int n_frames, mvs_frame_result* out);

/* Copy results of a slot to host buffers (any may be NULL) and wait for them.  The copies run on the context's
 * copy stream behind the slot's last kernel, so a batch enqueued on OTHER slots keeps the GPU busy meanwhile:
 * alternate two slot ranges (run batch k+1, then fetch batch k) to hide the PCIe transfers behind the kernels.
 * Host-pointer inputs of mvs_frame_run(_batch) are uploaded on the same copy stream. */
int mvs_frame_fetch(mvs_ctx* ctx, int slot, float* disp, size_t dstep, double* Xs, int32_t* xy, int32_t* lookup,
                    int32_t* n_out, double centroid[3], float* cam_hist);

/* The same for n slots at once (first_slot .. first_slot + n - 1) with ONE blocking point per batch instead of two per
 * frame: all copies are enqueued on the copy stream, the host waits once, and only when a point cloud is requested does it
 * wait a second time (the clouds' sizes N are known after the first wait).  d[i] holds the destination pointers of slot
 * first_slot + i (any may be NULL). */
typedef struct {
    float* disp;
    size_t dstep;
    double* Xs;
    int32_t* xy;
    int32_t* lookup;
    int32_t* n_out;
    double* centroid; /* 3 doubles */
    float* cam_hist;
    mvs_sparse_hist_cell* slic_to_hist_xy; /* n_labels cells of the slot's frame (skipped when it had no label image) */
} mvs_fetch_desc;
int mvs_frame_fetch_batch(mvs_ctx* ctx, int first_slot, int n, const mvs_fetch_desc* d);

/* Sum the per-camera histograms of slots [first_slot, first_slot + n_slots) in slot order into a device buffer owned by the
 * context (floor-histogram.cpp:296-300); returns its device address (hw*hh floats).  This buffer is what
 * a multi-GPU caller hands to ncclAllReduce / ncclAllGather. */
int mvs_frame_hist_sum(mvs_ctx* ctx, int first_slot, int n_slots, float** hist_dev, int* hw, int* hh);

/* -- test taps (kernel-level parity; not used by the reference-facing path) ---------------------------
 * Run cost volume + aggregation for grey images and copy the 16-bit volumes back ([h][w1][num_disp]).
 * Either output may be NULL. */
int mvs_sgbm_volumes(mvs_ctx* ctx, const mvs_sgbm_params* p, const uint8_t* l, const uint8_t* r, int w, int h,
                     size_t step, int16_t* C, int16_t* S, int16_t* raw_disp16);

/* Raise the error flag that an expired wait of the aggregation sweeps raises (the sweeps' waits are bounded: two
 * seconds).  The next mvs_sync / mvs_frame_fetch(_batch) / host-buffer SGBM call then fails with MVS_ERR_CUDA and clears
 * the flag.  Device-only callers (mvs_frame_run_batch + NCCL on mvs_tail_stream) learn of a failed sweep from mvs_sync. */
int mvs_test_raise_wave_error(mvs_ctx* ctx, int code);

/* Time (ms, CUDA events on the context stream) spent in the named stage during the last sgbm call:
 * "cost", "aggregate", "wta", "post".  Returns < 0 when unknown. */
float mvs_stage_ms(mvs_ctx* ctx, const char* stage);

#ifdef __cplusplus
}
#endif
#endif /* MVSTEREO_H */
