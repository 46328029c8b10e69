/*
 * b2stereo.h -- C ABI of libb2stereo.so: the B200 (sm_100a) stereo disparity engine.
 *
 * Drop-in boundary for the ONE hot path of ocean1100/stereo_match: what the
 * reference hands to OpenCV.  Each entry point names the reference call site it
 * replaces (paths under the reference repository):
 *
 *   b2s_sgbm_*      cv2.StereoSGBM_create(...).compute(l, r)
 *                       stereo_vision/stereo_vision.py:153-163,178
 *                       disparity_test.py:165-177, try_try.py:69-81
 *   b2s_bm_*        cv2.StereoBM_create(numDisparities, blockSize).compute(l, r)
 *                       stereo_vision/stereo_vision.py:165-166,178
 *   b2s_reproject   cv2.reprojectImageTo3D(disp, Q)
 *                       stereo_vision/stereo_vision.py:209, disparity_calculation.py:302
 *   b2s_remap_u8    cv2.remap(img, mapx, mapy, cv2.INTER_LINEAR)
 *                       rectified_img_cal.py:253-254, disparity_calculation.py:207-208
 *   b2s_bgr2gray    cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
 *                       disparity_calculation.py:286-287
 *   b2s_remap_gray_u8   the two above fused (BGR in, rectified gray out)
 *   b2s_pipeline    remap x2 -> (gray) -> StereoSGBM.compute -> reprojectImageTo3D, one call
 *                       disparity_calculation.py:207-208, 286-289, 302
 *   b2s_compact_points  points_3d[mask], colors[mask] in front of io_functions.write_ply
 *                       disparity_calculation.py:309-321, try_try.py:91-94
 *   b2s_median3_i16, b2s_filter_speckles
 *                   the cv2.medianBlur / cv2.filterSpeckles that StereoSGBM.compute
 *                   runs internally, exposed for stage-level parity
 *
 * Conventions
 *   - every data pointer is a DEVICE pointer on the current CUDA device unless the
 *     name ends in _host; images are dense row-major, batch-major ([batch][H][W]).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *     Calls are asynchronous on that stream.
 *   - return value: 0 = ok, otherwise a B2S_E_* code; b2s_last_error() returns a
 *     thread-local message that names the violated condition.
 *   - no CPU fallback exists: without a usable CUDA device every compute entry
 *     point fails with B2S_E_CUDA.
 */
#ifndef B2STEREO_H
#define B2STEREO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B2S_API __attribute__((visibility("default")))
#else
#define B2S_API
#endif

#define B2S_OK 0
#define B2S_E_INVALID 1      /* bad argument (message mirrors the cv2 assertion) */
#define B2S_E_UNSUPPORTED 2  /* valid for cv2 but outside this engine's contract */
#define B2S_E_CUDA 3         /* CUDA runtime error */
#define B2S_E_NOMEM 4

/* cv2.STEREO_SGBM_MODE_* */
#define B2S_MODE_SGBM 0
#define B2S_MODE_HH 1
#define B2S_MODE_SGBM_3WAY 2 /* 3 paths (-> , down, <-) over cv2's four fixed image stripes */
#define B2S_MODE_HH4 3       /* 4 paths: -> , <- , down, up (with cv2 4.13.0's bottom-row rule) */

/* matcher parameters exactly as cv2.StereoSGBM_create receives them (raw, before
   cv2's silent fallbacks -- those are applied inside) */
typedef struct b2s_sgbm_params {
    int32_t minDisparity;
    int32_t numDisparities;
    int32_t blockSize;
    int32_t P1;
    int32_t P2;
    int32_t disp12MaxDiff;
    int32_t preFilterCap;
    int32_t uniquenessRatio;
    int32_t speckleWindowSize;
    int32_t speckleRange;
    int32_t mode;
} b2s_sgbm_params;

/* cv2.StereoBM parameters that the reference path can reach (prefilter XSOBEL).  minDisparity is 0 in
   the reference's own calls and -(numDisparities - 1) in the right-view matcher that cv2.ximgproc.createRightMatcher
   derives from it (stereo_vision/stereo_vision.py:165-171,179); any value is bit-exact with cv2 4.13.0. */
typedef struct b2s_bm_params {
    int32_t numDisparities;   /* >0, multiple of 16 */
    int32_t blockSize;        /* odd, 5..255 */
    int32_t preFilterCap;     /* 1..63, cv2 default 31 */
    int32_t textureThreshold; /* cv2 default 10 */
    int32_t uniquenessRatio;  /* cv2 default 15 */
    int32_t speckleWindowSize;/* cv2 default 0 */
    int32_t speckleRange;     /* cv2 default 0 */
    int32_t minDisparity;     /* cv2 default 0 */
    int32_t disp12MaxDiff;    /* cv2 default -1 (off).  >= 4096 can never reject (the ximgproc factories set 1000000)
                                 and is accepted; 0..4095 is B2S_E_UNSUPPORTED (cv2's validateDisparity is not restated) */
} b2s_bm_params;

typedef struct b2s_matcher* b2s_handle;

/* derived geometry of an SGBM call, for callers that want to inspect volumes */
typedef struct b2s_sgbm_geometry {
    int32_t minX1, maxX1, width1; /* computed columns [minX1, maxX1) */
    int32_t Dp;                   /* padded disparity stride of the cost volumes */
    int32_t invalid;              /* (minDisparity-1)*16 */
    int32_t P1, P2, ftzero, uniquenessRatio, disp12MaxDiff; /* effective values */
    int32_t ndirs;                /* 5 (MODE_SGBM), 8 (MODE_HH), 4 (MODE_HH4) or 3 (MODE_SGBM_3WAY) */
    int32_t R, G;                 /* Dp = 2*R*G: G lanes per pixel, R packed registers per lane */
    int32_t swizzled;             /* 1: cell d of a [Dp] record sits at ((o % (R/4))*G + o/(R/4))*8 + d%8, o = d/8 */
} b2s_sgbm_geometry;

B2S_API const char* b2s_last_error(void);
B2S_API const char* b2s_version(void);
/* number of this library's kernels launched by the calling thread so far */
B2S_API uint64_t b2s_launch_count(void);

/* ---- StereoSGBM ---------------------------------------------------------- */
B2S_API int b2s_sgbm_create(const b2s_sgbm_params* params, b2s_handle* out);
B2S_API int b2s_sgbm_set_params(b2s_handle h, const b2s_sgbm_params* params);
B2S_API int b2s_sgbm_get_geometry(b2s_handle h, int W, b2s_sgbm_geometry* out);
/* bytes of workspace the handle will hold for (H, W, batch); allocated lazily */
B2S_API size_t b2s_sgbm_workspace_bytes(b2s_handle h, int H, int W, int batch);
/* left/right: [batch][H][W] uint8 ; disp: [batch][H][W] int16 (Q4 fixed point) */
B2S_API int b2s_sgbm_compute(b2s_handle h, const uint8_t* left, const uint8_t* right,
                     int H, int W, int batch, int16_t* disp, void* stream);
/* same with cn = 1 or 3 interleaved channels per pixel ([batch][H][W][cn]): cv2.StereoSGBM.compute on colour images
   (reference: try_try.py:69-81) sums the matching cost of the three channels */
B2S_API int b2s_sgbm_compute_cn(b2s_handle h, const uint8_t* left, const uint8_t* right,
                     int H, int W, int cn, int batch, int16_t* disp, void* stream);
/* stage taps for parity tests: copies the last compute's volumes (device to device, on
   `stream`) into caller buffers of `elems` uint16 each (either may be NULL); layout
   [batch][H][width1][Dp] (cell order inside a record: see b2s_sgbm_geometry.swizzled).  cost = raw block cost (cv2's C minus P2, padded cells 0);
   S = min(32767, sum of path costs) (padded cells 32767).  The production pipeline never
   materialises the full S (the last sweep feeds the winner-take-all from registers); call
   b2s_sgbm_keep_volumes(h, 1) before compute to have it written as well. */
B2S_API int b2s_sgbm_keep_volumes(b2s_handle h, int on);
B2S_API int b2s_sgbm_debug_volumes(b2s_handle h, uint16_t* cost_out, uint16_t* S_out,
                                   size_t elems, void* stream);
/* 1 if the last compute left the domain in which bit-exact parity with cv2 is
   guaranteed: 2*max(raw block cost) + 3*P2 > 32767.  Beyond it cv2's saturating SIMD
   lanes and its scalar tail disagree with each other, so results there
   are unpinned (they may still match).  Synchronises the stream. */
B2S_API int b2s_sgbm_overflow_flag(b2s_handle h, int* flag_host);

/* opt-in stage timing for bench.py: when enabled, compute records CUDA events on its
   stream between stages; b2s_sgbm_stage_ms waits for the last compute and writes the
   B2S_NSTAGES stage durations (milliseconds) to a HOST array:
   0 descriptors, 1 cost volume, 2 path aggregation (all directions), 3 WTA + LR check,
   4 median, 5 speckle filter. */
#define B2S_NSTAGES 6
B2S_API int b2s_sgbm_enable_timing(b2s_handle h, int on);
B2S_API int b2s_sgbm_stage_ms(b2s_handle h, float* ms_host);
/* The "aggregate" stage split into its two sweep launches (forward sweep; mirrored sweep with the fused winner-take-all):
   two durations in milliseconds to a HOST array.  Same conditions as b2s_sgbm_stage_ms; both values are 0 when the
   matcher ran the one-launch-per-direction fallback (numDisparities > 512). */
B2S_API int b2s_sgbm_sweep_ms(b2s_handle h, float* ms2_host);

/* ---- StereoBM ------------------------------------------------------------ */
B2S_API int b2s_bm_create(const b2s_bm_params* params, b2s_handle* out);
B2S_API int b2s_bm_compute(b2s_handle h, const uint8_t* left, const uint8_t* right,
                   int H, int W, int batch, int16_t* disp, void* stream);

B2S_API void b2s_destroy(b2s_handle h);

/* ---- the whole per-pair chain in one call ---------------------------------
   remap (cn = 1) or remap + BGR2GRAY (cn = 3) of both raw images -> StereoSGBM.compute ->
   reprojectImageTo3D(float(disp) * disp_scale, Q)      [skipped when xyz == NULL]
   Replaces disparity_calculation.py:207-208, 286-289, 302 (and stereo_vision.py:126-127, 178, 209).
   left_raw/right_raw: [batch][sH][sW][cn] uint8; maps: [H][W] float32 per view, shared by the batch;
   rect_l/rect_r: caller scratch, [batch][H][W] uint8 each (the rectified gray pair, left for the caller);
   disp: [batch][H][W] int16; xyz: [batch][H][W][3] float32; Q_host: 16 doubles on the HOST;
   scratch: >= batch floats, only for handle_missing.  Everything is enqueued on `stream`. */
B2S_API int b2s_pipeline(b2s_handle sgbm, const uint8_t* left_raw, const uint8_t* right_raw,
                 int sH, int sW, int cn, const float* mapx_l, const float* mapy_l,
                 const float* mapx_r, const float* mapy_r, int H, int W, int batch,
                 uint8_t* rect_l, uint8_t* rect_r, int16_t* disp, const double* Q_host,
                 float disp_scale, int handle_missing, float* scratch, float* xyz, void* stream);

/* ---- stand-alone stages -------------------------------------------------- */
#define B2S_DT_U8 0
#define B2S_DT_I16 1
#define B2S_DT_I32 2
#define B2S_DT_F32 3
#define B2S_DT_U16 4   /* b2s_map2dto3d, b2s_wls_filter only */
#define B2S_DT_F64 5   /* b2s_map2dto3d only */
/* xyz: [batch][H][W][3] float32.  Q_host: 16 doubles, row-major, HOST pointer.
   disp_scale multiplies the disparity after conversion to float32 (1.0 = cv2's
   behaviour; 1/16 fuses project_points_3D's "/16.0", stereo_vision.py:204).
   handle_missing needs `scratch` = device buffer of >= batch floats. */
B2S_API int b2s_reproject(const void* disp, int dtype, int H, int W, int batch,
                  const double* Q_host, float disp_scale, int handle_missing,
                  float* scratch, float* xyz, void* stream);
/* map2Dto3D of the reference (mapTo3D.py:35-56, called at :116): pinhole back-projection of a depth image in float64,
   u = -(col - cx) * z / fx, v = (row - cy) * z / fy, z, for the pixels with pt2d > threshold.  pt2d, depth:
   [batch][H][W] of the given B2S_DT_* type; xyz: [batch][H][W][3] float64 (NaN where the pixel fails the threshold);
   npass: one device uint64, receives the number of passing pixels (the reference only works when all pass). */
B2S_API int b2s_map2dto3d(const void* pt2d, int pt2d_dtype, const void* depth, int depth_dtype,
                  int H, int W, int batch, double cx, double cy, double fx, double fy,
                  double threshold, double* xyz, unsigned long long* npass, void* stream);
/* src: [batch][sH][sW][cn] uint8 ; maps: [H][W] float32 shared by the batch */
B2S_API int b2s_remap_u8(const uint8_t* src, int sH, int sW, int cn, const float* mapx,
                 const float* mapy, int H, int W, int batch, uint8_t* dst, void* stream);
B2S_API int b2s_bgr2gray(const uint8_t* bgr, size_t npixels, uint8_t* gray, void* stream);
/* cv2.remap of a BGR image followed by cv2.cvtColor(BGR2GRAY), fused (same bytes as the two
   calls; disparity_calculation.py:207-208 + :286-287).  src: [batch][sH][sW][3], dst: [batch][H][W] */
B2S_API int b2s_remap_gray_u8(const uint8_t* src_bgr, int sH, int sW, const float* mapx,
                      const float* mapy, int H, int W, int batch, uint8_t* dst_gray, void* stream);
/* Point cloud behind reprojectImageTo3D (disparity_calculation.py:309-321, try_try.py:91-94): keeps the pixels with
   lo < disp < hi in row-major order (numpy's points_3d[mask]), their colours as RGB bytes from the BGR image (bgr and
   rgb may both be NULL), NaN / inf coordinates replaced by 0.  Images of a batch are packed one after the other:
   offsets[b] .. offsets[b+1] are the vertices of image b (device array of batch + 1 int64; offsets[batch] = total).
   verts / rgb must hold batch*H*W vertices (worst case); scratch: b2s_compact_points_scratch_bytes bytes. */
B2S_API size_t b2s_compact_points_scratch_bytes(int H, int W, int batch);
B2S_API int b2s_compact_points(const float* xyz, const void* disp, int disp_dtype, double lo, double hi,
                       const uint8_t* bgr, int H, int W, int batch, float* verts, uint8_t* rgb,
                       long long* offsets, void* scratch, void* stream);
/* ---- disparity WLS filter (stereo_vision/stereo_vision.py:171-182, disparity_test.py:182-199) -------------------
   Replaces cv2.ximgproc.createDisparityWLSFilter(...).filter(displ, gray_l, None, dispr).  PARITY UNPINNED: no
   cv2.ximgproc exists in this image; the kernels are bit-exact with the restatement in oracle/wls_oracle.c.
   The parameter rules of createDisparityWLSFilter(matcher) live in the host layer (stereo_match_b200/wls.py). */
typedef struct b2s_wls* b2s_wls_handle;
typedef struct {
    int32_t roi_x, roi_y, roi_w, roi_h;  /* valid-disparity rectangle of the left view */
    int32_t min_disp;                    /* pixels outside the rectangle get 16 * (min_disp - 1) */
    int32_t use_confidence;              /* 1: LRC / depth-discontinuity confidence (needs disp_right) */
    int32_t lrc_thresh;                  /* 24 */
    int32_t radius;                      /* depth-discontinuity radius */
    float roll_off;                      /* 0.001 */
    float lambda, sigma;                 /* setLambda / setSigmaColor */
    int32_t num_iter;                    /* 3 */
    float attenuation;                   /* 0.25 */
} b2s_wls_params;
B2S_API int b2s_wls_create(b2s_wls_handle* out);
B2S_API void b2s_wls_destroy(b2s_wls_handle h);
/* disp_left, disp_right (may be NULL without confidence), out: [batch][H][W] int16; guide: [batch][H][W][cn] uint8,
   cn = 1 or 3; conf_out (may be NULL): [batch][H][W] float32, the confidence map (0..255, 0 outside the rectangle).
   The handle owns the guide-weight table and a workspace grown on demand; everything is enqueued on `stream`. */
B2S_API int b2s_wls_filter(b2s_wls_handle h, const int16_t* disp_left, const int16_t* disp_right,
                   const uint8_t* guide, int cn, int H, int W, int batch, const b2s_wls_params* params,
                   int16_t* out, float* conf_out, void* stream);
B2S_API int b2s_median3_i16(const int16_t* src, int H, int W, int batch, int16_t* dst, void* stream);
/* in place; labels: device scratch of batch*H*W int32, counts likewise */
B2S_API int b2s_filter_speckles(int16_t* img, int H, int W, int batch, int newVal,
                        int maxSpeckleSize, int maxDiff, int32_t* labels,
                        int32_t* counts, void* stream);

/* No counterpart in the reference.  b2s_filter_speckles for callers that
   cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the labels / counts
   scratch is a static device array inside the library (8 MiB per GPU that loads it, held until exit), so the call
   allocates nothing.  Limit: batch*H*W <= B2S_SPECKLE_NOALLOC_MAX_PIXELS, larger calls return B2S_E_UNSUPPORTED.
   All calls on a device share that scratch: calls that overlap on the GPU race -- keep them ordered (one stream, or
   events between streams).  The first call on a device resolves the array's address (do it before a capture starts). */
#define B2S_SPECKLE_NOALLOC_MAX_PIXELS (1 << 20)
B2S_API int b2s_filter_speckles_noalloc(int16_t* img, int H, int W, int batch, int newVal,
                        int maxSpeckleSize, int maxDiff, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2STEREO_H */
