/*
 * vslam.h -- C ABI of the B200-native KeyPointDetection hot path (libvslam_b200.so).
 *
 * The reference (JacobYoung115/VisualSLAM, KeyPointDetection/) has no FFI layer: its surface is a C++ class and
 * free functions compiled into the executables.  This header is the boundary a maintainer binds
 * instead of OpenCV; each entry point cites the reference interface it replaces (paths relative to
 * KeyPointDetection/).  The C++17 mirror of the reference's own names lives in visualslam_b200/host/.
 *
 * Conventions
 *   - plain pointers and sizes only; the caller owns every host buffer; the library owns device memory inside
 *     opaque handles released by the matching *_destroy.
 *   - every function returns 0 (VSLAM_OK) or a negative vslam_status; vslam_last_error() gives the text.
 *   - a context is bound to one GPU and one CUDA stream and is not thread-safe: one context per (thread, GPU).
 *   - images are single-channel uint8, row-major, `pitch` bytes between rows (pitch >= width).
 *   - there is no CPU fallback: without a CUDA device vslam_ctx_create fails with VSLAM_E_CUDA.
 */
#ifndef VSLAM_H_
#define VSLAM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define VSLAM_API
#else
#define VSLAM_API __attribute__((visibility("default")))
#endif

typedef enum vslam_status {
    VSLAM_OK = 0,
    VSLAM_E_INVALID = -1,  /* bad argument (null pointer, non-positive size, unsupported kernel width ...) */
    VSLAM_E_CUDA = -2,     /* a CUDA call failed; see vslam_last_error */
    VSLAM_E_CAPACITY = -3, /* output list too small; counts[] hold the true sizes and the list holds the first `capacity`
                            * entries of the ordered result.  One launch batch stages at most max(capacity, 2^20) entries:
                            * if it emits more, the positions are still those of the full ordered result, but entries that
                            * found no staging slot are returned as markers with row = -1 (never silently as garbage) */
    VSLAM_E_NOMEM = -4     /* device or host allocation failed */
} vslam_status;

typedef struct vslam_ctx vslam_ctx;         /* one GPU + one stream + reusable workspaces */
typedef struct vslam_pyramid vslam_pyramid; /* device-resident GaussPyramid of a batch of frames */

/* SLAM::point (Diff_of_Gauss.cpp:27-35): six int32, same order. row/col are the PADDED coordinates the reference
 * stores (Diff_of_Gauss.cpp:289). */
typedef struct vslam_keypoint {
    int32_t row, col, value, padding, octave, level;
} vslam_keypoint;

/* One NMS2 corner (Harris_corners.cpp:116-118): position and response value. */
typedef struct vslam_corner {
    int32_t row, col;
    float value;
} vslam_corner;
/* The same corner in 8 bytes, for callers that move large lists across PCIe (no counterpart in the reference; lossless for
 * images below 65536 rows / columns): OR VSLAM_COMPACT into vslam_harris's out_space and pass a vslam_corner8 buffer.
 * An entry the staging buffer had no room for is marked row = col = 0xffff (the 12-byte list marks it row = -1). */
typedef struct vslam_corner8 {
    uint16_t row, col;
    float value;
} vslam_corner8;

/* GaussPyramid ctor arguments + constants (GaussPyramid.hpp:17-21,65-69). */
typedef struct vslam_pyramid_params {
    int32_t num_octaves;   /* 4 in Diff_of_Gauss.cpp:742; <=0 -> floor(log2(min(rows,cols)))-4 (GaussPyramid.cpp:150-152) */
    int32_t scale_samples; /* scaleSamples_ = 3 -> numLevels_ = 6 (only 3 is supported) */
    double sigma0;         /* 1.6 in Diff_of_Gauss.cpp:743 */
    uint32_t materialise;  /* bit mask of vslam_map_kind to keep in HBM; gauss and DoG are always kept */
} vslam_pyramid_params;

/* initialKeypointDetection / FeaturePointLocalization / filterKeypoints constants
 * (Diff_of_Gauss.cpp:772 windowSize=3, :245 0.03f, :332 r=10). */
typedef struct vslam_dog_params {
    int32_t window;
    float contrast_thr;
    float edge_r;
} vslam_dog_params;

/* HarrisCorner / NMS constants (Harris_corners.cpp:34-36 window 3, k 0.04; :178 window 3; :179 window 5). */
typedef struct vslam_harris_params {
    float k;
    int32_t nms2_window;
} vslam_harris_params;

typedef enum vslam_map_kind {
    VSLAM_MAP_BASE = 0,   /* octaveImage / imagePyramid           u8   (GaussPyramid.hpp:29,38) level ignored  */
    VSLAM_MAP_GAUSS = 1,  /* octaveBlur / pyramidGauss            u8   (:31,39)  level 0..5                     */
    VSLAM_MAP_DOG = 2,    /* octaveDiff / pyramidDiff             u8   (:32,40)  level 0..4                     */
    VSLAM_MAP_GRADX = 3,  /* octaveGradX / pyramidGradX           f32  (:33,41)  level 0..5                     */
    VSLAM_MAP_GRADY = 4,  /* octaveGradY / pyramidGradY           f32  (:34,42)                                 */
    VSLAM_MAP_MAG = 5,    /* octaveGradMag / pyramidGradMag       f32  (:35,43)                                 */
    VSLAM_MAP_ORIENT = 6  /* octaveGradOrient / pyramidGradOrient f32  (:36,44)  degrees [0,360)                */
} vslam_map_kind;

#define VSLAM_MATERIALISE_GRAD (1u << VSLAM_MAP_GRADX | 1u << VSLAM_MAP_GRADY)
#define VSLAM_MATERIALISE_MAGORIENT (1u << VSLAM_MAP_MAG | 1u << VSLAM_MAP_ORIENT)
#define VSLAM_MATERIALISE_ALL (VSLAM_MATERIALISE_GRAD | VSLAM_MATERIALISE_MAGORIENT)

/* memory space of a caller buffer */
#define VSLAM_HOST 0
#define VSLAM_DEVICE 1
#define VSLAM_COMPACT 0x100 /* vslam_harris / vslam_dog_harris: `corners` receives vslam_corner8 entries (8 bytes) instead of vslam_corner */
#define VSLAM_DEFER 0x200   /* vslam_dog_harris with host lists: return once the work is enqueued; vslam_finish completes it */

/* ---- context ------------------------------------------------------------------------------------------------ */
/* Stream contract.  Every call enqueues its work on the context's stream and returns; host-space outputs are complete on
 * return, device-space outputs when the stream reaches that point (vslam_sync, or later work on the same stream).
 * vslam_ctx_create makes a private NON-BLOCKING stream: device buffers produced or consumed by the caller's other streams
 * are NOT ordered with it -- finish that work before the call and vslam_sync before reading device outputs, or create the
 * context on the caller's stream with vslam_ctx_create_on_stream, which orders everything with the caller's own work.
 * The GaussPyramid materialise mask (vslam_pyramid_params.materialise, VSLAM_MATERIALISE_*) makes the chosen fp32 gradient
 * maps of every level resident at build time; maps outside the mask are formed on demand by vslam_pyramid_get.
 * Kernel widths: octaves whose 8U Gaussian kernels exceed 247 taps (sigma > ~41: the fifth octave and beyond at sigma0 = 1.6,
 * i.e. vslam_calculate_num_octaves on images whose short side is >= 512) run on an exact but slow global-memory path; they are tiny,
 * so a pyramid that has one is built frame by frame (vslam_dog_keypoints / vslam_dog_harris fall back to materialising each frame;
 * VSLAM_DEFER is refused).  Four octaves -- what ./DoG uses (Diff_of_Gauss.cpp:742) -- stay on the fused kernels. */
VSLAM_API int vslam_ctx_create(int device, vslam_ctx **out);
/* Same, but work is enqueued on a caller-owned CUDA stream (cudaStream_t / CUstream passed as void*). */
VSLAM_API int vslam_ctx_create_on_stream(int device, void *cuda_stream, vslam_ctx **out);
VSLAM_API void vslam_ctx_destroy(vslam_ctx *ctx);
VSLAM_API const char *vslam_last_error(const vslam_ctx *ctx); /* ctx may be NULL: last process-wide error */
VSLAM_API int vslam_sync(vslam_ctx *ctx);
VSLAM_API const char *vslam_version(void);
/* number of kernel launches this context has enqueued since creation (bench.py's gpu_launches) */
VSLAM_API long long vslam_launch_count(const vslam_ctx *ctx);

/* Per-kernel device timing with CUDA events on the context's stream (used by bench.py's roofline leg).
 * vslam_profile(ctx, 1) clears the accumulators and starts bracketing every launch; vslam_profile_read syncs the
 * stream and returns the summed duration and launch count of one kernel class. */
typedef enum vslam_kernel_id {
    VSLAM_K_UPSAMPLE = 0, VSLAM_K_PAD = 1 /* reflect-101 margins of the octave bases */, VSLAM_K_VBLUR = 2, VSLAM_K_DETECT = 3, VSLAM_K_COMPACT = 4,
    VSLAM_K_HARRIS = 5, VSLAM_K_GRAD = 6, VSLAM_K_FUSED = 7 /* octave_fused_k: dp4a / dp2a, materialising launches */,
    VSLAM_K_DESCRIPTOR = 8 /* orientation + SIFT */, VSLAM_K_OCTAVE_MMA = 9 /* octave_mma_k: mma.sync */, VSLAM_K_OCTAVE_UMMA = 10 /* octave_umma_k: tcgen05 + TMA */,
    VSLAM_KERNEL_COUNT = 11
} vslam_kernel_id;
VSLAM_API int vslam_profile(vslam_ctx *ctx, int enable);
VSLAM_API int vslam_profile_read(vslam_ctx *ctx, int kernel_id, double *total_ms, long long *launches);

/* ---- sigma table: GaussPyramid::calculateSigma / calculateNumOctaves (GaussPyramid.cpp:150-162) --------------- */
VSLAM_API double vslam_calculate_sigma(int octave, int level, double sigma0, int scale_samples);
VSLAM_API int vslam_calculate_num_octaves(int rows, int cols);
/* OpenCV's 8U Gaussian kernel as GaussianBlur(Size(0,0), sigma) builds it (GaussPyramid.cpp:177): ksize and the
 * 8.8 fixed-point taps (sum 256).  taps must hold vslam_gauss_ksize(sigma) ints. */
VSLAM_API int vslam_gauss_ksize(double sigma);
VSLAM_API int vslam_gauss_taps(double sigma, int ksize, int32_t *taps);

/* ---- GaussPyramid: createPyramid (GaussPyramid.cpp:106-131) ---------------------------------------------------- */
/* Replaces the GaussPyramid constructors (GaussPyramid.hpp:17-21) for a batch of n_frames equally sized frames.
 * frames: n_frames images, frame f at frames + f*frame_stride.  `space` says where `frames` lives. */
VSLAM_API int vslam_pyramid_build(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols,
                                  size_t pitch, size_t frame_stride, int space,
                                  const vslam_pyramid_params *params, vslam_pyramid **out);
/* The same pyramid object on a new batch of frames of the same number, size and parameters: every map is rebuilt in place and
 * nothing is allocated (a stream of batches otherwise pays a multi-GB cudaMalloc / cudaFree per batch: 7.6 of 14.7 ms per eight
 * 1080p frames with every gradient map resident).  Maps handed out by vslam_pyramid_device_map keep their addresses. */
VSLAM_API int vslam_pyramid_rebuild(vslam_ctx *ctx, vslam_pyramid *pyr, const uint8_t *frames, size_t pitch, size_t frame_stride,
                                    int space);
VSLAM_API void vslam_pyramid_destroy(vslam_pyramid *pyr);
VSLAM_API int vslam_pyramid_num_octaves(const vslam_pyramid *pyr);
VSLAM_API int vslam_pyramid_num_frames(const vslam_pyramid *pyr);
VSLAM_API int vslam_pyramid_octave_size(const vslam_pyramid *pyr, int octave, int *rows, int *cols);
/* getSigmaAt (GaussPyramid.cpp:187-189) */
VSLAM_API int vslam_pyramid_sigma(const vslam_pyramid *pyr, int octave, int level, double *sigma);
/* octaveBlur/octaveDiff/... getters (GaussPyramid.hpp:29-36): copy one map to a caller buffer (host or device).
 * dst_pitch in bytes.  Maps outside `materialise` (gradients) are computed on demand from the Gaussian level. */
VSLAM_API int vslam_pyramid_get(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int kind, int octave,
                                int level, void *dst, size_t dst_pitch, int space);
/* device pointer + pitch of a resident u8 map (BASE, GAUSS, DOG) for zero-copy consumers; NULL if not resident */
VSLAM_API const uint8_t *vslam_pyramid_device_map(const vslam_pyramid *pyr, int frame, int kind, int octave,
                                                  int level, size_t *pitch);

/* ---- DoG keypoints: initialKeypointDetection + FeaturePointLocalization + filterKeypoints' edge test
 *      (Diff_of_Gauss.cpp:254-297, 223-251, 79-109, 326-338) ----------------------------------------------------- */
/* All frames and octaves of `pyr`; keypoints in the reference's emission order (frame, octave, level, row, col).
 * out:        capacity entries (host or device per `space`)
 * counts:     n_frames ints (host): keypoints per frame
 * stage_counts: optional n_frames*num_octaves*4 int64 (host): tested, candidates, after contrast, after edge */
VSLAM_API int vslam_dog_detect(vslam_ctx *ctx, const vslam_pyramid *pyr, const vslam_dog_params *params,
                               vslam_keypoint *out, int capacity, int space, int *counts,
                               long long *stage_counts);

/* Throughput path of ./DoG up to Diff_of_Gauss.cpp:336: frames -> keypoints with a per-context workspace that is
 * reused frame after frame (nothing but keypoints leaves the GPU).  Same results and order as
 * vslam_pyramid_build + vslam_dog_detect. */
VSLAM_API int vslam_dog_keypoints(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols,
                                  size_t pitch, size_t frame_stride, int space,
                                  const vslam_pyramid_params *pparams, const vslam_dog_params *dparams,
                                  vslam_keypoint *out, int capacity, int out_space, int *counts);

/* ---- frame upload shared by several calls ------------------------------------------------------------------------ */
/* Starts the asynchronous, chunked upload of a batch of HOST frames (pinned memory for full PCIe speed) into the
 * context's ingest buffer and returns its device address, pitch and frame stride.  Pass those with VSLAM_DEVICE to
 * vslam_dog_keypoints / vslam_harris / vslam_pyramid_build: each call waits per chunk, so kernels on the first frames
 * overlap the upload of the rest, and one upload serves every call.  The context holds TWO ingest buffers and alternates between
 * them: the returned frames stay valid until the SECOND next call that uploads (vslam_upload_frames, or any call given HOST
 * frames), so a caller can start the upload of batch i+1 before it calls the detectors on batch i -- the copy then runs under
 * batch i's kernels, and a batch that has already landed runs in full-size launches.  An upload into a buffer first waits for
 * the compute-stream work enqueued so far (the consumers of the batch the buffer held). */
VSLAM_API int vslam_upload_frames(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols, size_t pitch,
                                  size_t frame_stride, const uint8_t **d_frames, size_t *d_pitch,
                                  size_t *d_frame_stride);

/* JPEG streams in host memory -> grey frames (imread(path, IMREAD_COLOR) + cvtColor(COLOR_BGR2GRAY), Diff_of_Gauss.cpp:730-739,
 * Harris_corners.cpp:148-149): nvJPEG decodes to BGR on the device, the bit-exact cvtColor kernel turns it
 * grey.  gray: n frames of rows x cols in `space`; pass VSLAM_DEVICE frames straight to the detectors.  Every stream must decode to
 * rows x cols (vslam_jpeg_info).  NOT a parity target: a decoder's output differs from libjpeg's by a few grey levels.  libnvjpeg is
 * opened at run time; without it both calls return VSLAM_E_CUDA and everything else works. */
VSLAM_API int vslam_jpeg_info(vslam_ctx *ctx, const uint8_t *stream, size_t length, int *rows, int *cols, int *channels);
VSLAM_API int vslam_decode_jpeg_gray(vslam_ctx *ctx, const uint8_t *const *streams, const size_t *lengths, int n, int rows,
                                     int cols, uint8_t *gray, size_t gray_pitch, size_t gray_stride, int space);

/* ---- Harris: main's pre-blur + Sobel, HarrisCorner, convertScaleAbs + NonMaximumSuppression, NMS2
 *      (Harris_corners.cpp:158-164, 31-68, 176-178, 83-129) ------------------------------------------------------ */
/* Any of response / r8 / mask3 may be NULL (not materialised).  Maps are dense n_frames x rows x cols in `space`.
 * corners: capacity entries in row-major order per frame, frames concatenated; counts: n_frames ints (host).
 * out_space: VSLAM_HOST or VSLAM_DEVICE, optionally | VSLAM_COMPACT (corners is then a vslam_corner8 buffer: a third fewer
 * bytes to bring back, the dominant device-to-host transfer of the end-to-end path). */
VSLAM_API int vslam_harris(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols, size_t pitch,
                           size_t frame_stride, int space, const vslam_harris_params *params, float *response,
                           uint8_t *r8, uint8_t *mask3, vslam_corner *corners, int capacity, int out_space,
                           int *counts);

/* Both detectors over the same batch in one call: what a caller who wants ./DoG's keypoints and ./Harris's corners of every frame
 * does (vslam_dog_keypoints + vslam_harris with response / r8 / mask3 = NULL), with the Harris launch of every chunk of frames
 * issued in front of the chunk's DoG launches, so that the (much larger) corner list leaves the GPU chunk by chunk under the DoG
 * kernels instead of after them.  Lists, order and counts are exactly those of the two separate calls.  out_space: VSLAM_HOST or
 * VSLAM_DEVICE for both lists, | VSLAM_COMPACT for the corner list.  Returns VSLAM_E_CAPACITY if either list overflowed. */
VSLAM_API int vslam_dog_harris(vslam_ctx *ctx, const uint8_t *frames, int n_frames, int rows, int cols, size_t pitch,
                               size_t frame_stride, int space, const vslam_pyramid_params *pparams,
                               const vslam_dog_params *dparams, const vslam_harris_params *hparams,
                               vslam_keypoint *keypoints, int kp_capacity, int *kp_counts, vslam_corner *corners,
                               int co_capacity, int *co_counts, int out_space);

/* Deferred completion, for callers that process a stream of batches (no counterpart in the reference): with VSLAM_HOST |
 * VSLAM_DEFER vslam_dog_harris returns as soon as its work is enqueued -- nothing in the host lists or the count arrays is valid
 * yet.  vslam_finish waits for the OLDEST pending call, fills its count arrays, copies its lists into the host buffers given to
 * that call and returns its status (VSLAM_OK, VSLAM_E_CAPACITY ...).  At most two calls may be pending, so the usual loop is
 *   vslam_dog_harris(batch i, buffers[i & 1], ... | VSLAM_DEFER); if (i) vslam_finish(ctx); upload(i+1);      (vslam_finish completes batch i-1)
 * and the lists of batch i-1, then the frames of batch i+1, cross PCIe under the kernels of batch i (one direction at a time: a
 * host shared by several GPUs moves far fewer bytes when both directions are busy at once).  Buffers and count arrays of a pending call must stay
 * alive and untouched until its vslam_finish.  vslam_sync does not complete pending calls; vslam_ctx_destroy drops them. */
VSLAM_API int vslam_finish(vslam_ctx *ctx);

/* ---- per-function entry points: the reference's own function granularity, for the C++17 mirror in
 *      visualslam_b200/host/ (HOST buffers, dense unless a pitch is given; each call synchronises).  The throughput
 *      paths above (vslam_dog_keypoints, vslam_harris) fuse these stages; results are identical. ------------------- */
/* cvtColor(img_color, img, COLOR_BGR2GRAY) on 8-bit BGR frames (Diff_of_Gauss.cpp:739, tests/GaussPyramid_Test.cpp:81):
 * gray = (3735 B + 19235 G + 9798 R + 2^14) >> 15.  bgr: n_frames x rows x cols x 3 bytes (pitch in bytes per row),
 * gray: n_frames x rows x cols. */
VSLAM_API int vslam_bgr_to_gray(vslam_ctx *ctx, const uint8_t *bgr, int n_frames, int rows, int cols, size_t pitch,
                                size_t frame_stride, uint8_t *gray, size_t gray_pitch, size_t gray_stride);
/* SLAM::Rotation::rotate_mat_CCW(I, center, angles) (include/src/Rotation/rotation.cpp:66-84) on an 8U image: every
 * output pixel gathers I at rotate_pt_CW(pixel) (:19-28, int offsets x float cos/sin, truncated) when the rotated point
 * has 0 < x < cols and 0 < y < rows, else 0.  (cos_a, sin_a) = Rotation::cos_sin_of_angle(theta), computed by the caller. */
VSLAM_API int vslam_rotate_mat_ccw(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, int center_x,
                                   int center_y, float cos_a, float sin_a, uint8_t *dst, size_t dst_pitch);
/* SLAM::Rotation::getRotatedWindow(I, center, windowSize, theta) (rotation.cpp:87-108): dst[i][j] = I at
 * rotate_pt_CW(center - windowSize/2 + (j, i)), window x window (the reference's loops run one row / column further
 * and write past its ROI; its reads outside the image are undefined: 0 here). */
VSLAM_API int vslam_rotated_window(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, int center_x,
                                   int center_y, int window, float cos_a, float sin_a, uint8_t *dst, size_t dst_pitch);
/* SLAM::Rotation::doubleCrop's two OpenCV calls (include/src/Rotation/rotation.cpp:152,185 and :198); the function itself is
 * host geometry in visualslam_b200/host/src/rotation.cpp.
 * cv::getRotationMatrix2D(Point2f(center_x, center_y), angle, scale): m = 6 doubles, row-major 2 x 3.  Host arithmetic only. */
VSLAM_API int vslam_get_rotation_matrix_2d(float center_x, float center_y, double angle_deg, double scale, double *m);
/* cv::warpAffine(src, dst, M, Size(dst_cols, dst_rows)) with the defaults the reference uses: INTER_LINEAR, BORDER_CONSTANT 0,
 * 8UC1, M the forward 2 x 3 matrix (inverted inside, as OpenCV does).  Bit-exact with OpenCV's fixed-point path (10-bit
 * coordinates, 5 fractional bits, 15-bit weights). */
VSLAM_API int vslam_warp_affine(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, const double *m,
                                uint8_t *dst, int dst_rows, int dst_cols, size_t dst_pitch);
/* GaussianBlur(img, blurred, Size(3,3), 0, 0, BORDER_DEFAULT) on 8U (Harris_corners.cpp:157-158) */
VSLAM_API int vslam_gaussian_blur3(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, uint8_t *dst,
                                   size_t dst_pitch);
/* Sobel(src, CV_32F, 1,0 / 0,1, ksize=1, BORDER_DEFAULT) (Harris_corners.cpp:161-164, GaussPyramid.cpp:72-75); gx or gy may be NULL */
VSLAM_API int vslam_sobel1(vslam_ctx *ctx, const uint8_t *src, int rows, int cols, size_t pitch, float *gx, float *gy);
/* HarrisCorner(Ix, Iy) incl. StructureMatrix (Harris_corners.cpp:10-68): window 3, k 0.04 in the reference */
VSLAM_API int vslam_harris_response(vslam_ctx *ctx, const float *ix, const float *iy, int rows, int cols, int window,
                                    float k, float *response);
/* No counterpart in the reference: added for callers that cannot allocate (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The same map as vslam_harris's
 * `response` (= vslam_gaussian_blur3 -> vslam_sobel1 -> vslam_harris_response) of ONE image, but with DEVICE pointers for
 * the image (pitch in bytes) and the dense rows x cols float response, and with the intermediates (blurred image, Ix, Iy,
 * one counter) in a file-scope static __device__ array instead of cudaMalloc'd context buffers: the call allocates
 * nothing, enqueues its work on the context's stream and returns without synchronising.  d_positive: optional DEVICE int
 * that receives the number of responses > 0.  Limits: rows * cols <= VSLAM_STATIC_MAX_PIXELS (larger images return
 * VSLAM_E_INVALID with a message; use vslam_harris); the array (256 + 9 bytes per pixel of the maximum = 18 MiB) is held by
 * every GPU that loads the library until the process exits.  Every call on a device shares it: calls that overlap on the
 * GPU (different streams or contexts) race; calls on one stream are ordered and safe. */
#define VSLAM_STATIC_MAX_PIXELS 2097152 /* 2 Mi pixels: covers 1920 x 1080 */
VSLAM_API int vslam_harris_response_static(vslam_ctx *ctx, const uint8_t *d_image, int rows, int cols, size_t pitch,
                                           int window, float k, float *d_response, int *d_positive);
/* convertScaleAbs(src, dst) with alpha 1, beta 0 (Harris_corners.cpp:176) */
VSLAM_API int vslam_convert_scale_abs(vslam_ctx *ctx, const float *src, int rows, int cols, uint8_t *dst);
/* NonMaximumSuppression(response8u, 3) (Harris_corners.cpp:70-81): 255 where strictly above the 8 neighbours */
VSLAM_API int vslam_nms3(vslam_ctx *ctx, const uint8_t *r8, int rows, int cols, uint8_t *mask);
/* NMS2(response, window) (Harris_corners.cpp:83-129) as a proper float map: window maximum at the kept positions, else 0 */
VSLAM_API int vslam_nms2(vslam_ctx *ctx, const float *response, int rows, int cols, int window, float *out_map);
/* initialKeypointDetection (+ FeaturePointLocalization) of ONE octave of one frame (Diff_of_Gauss.cpp:254-297);
 * apply_edge != 0 also applies filterKeypoints' edge threshold (:326-338).  Host list, reference order. */
VSLAM_API int vslam_dog_detect_octave(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave,
                                      const vslam_dog_params *params, int apply_edge, vslam_keypoint *out,
                                      int capacity, int *count);
/* computeEdgeResponse (Diff_of_Gauss.cpp:79-109) of n keypoints of one octave (padded coordinates as stored) */
VSLAM_API int vslam_edge_responses(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave,
                                   const vslam_keypoint *keypoints, int n, float *responses);
/* FeaturePointLocalization's dog_zhat (Diff_of_Gauss.cpp:223-243) for n candidates given as int32 quadruples
 * (d_x, d_y, d_scale, value); the caller applies `> 0.03f` and `(int)(dog_zhat * 255.0f)` (:245-246) */
VSLAM_API int vslam_localize(vslam_ctx *ctx, const int32_t *candidates, int n, float *dog_zhat);
/* computeEdgeResponse's arithmetic (Diff_of_Gauss.cpp:93-107) on n windows of window_len (gx, gy) pairs:
 * responses[i] = tr^2 / det (or NULL); sums[3i..3i+2] = Ix2, Iy2, IxIy = StructureMatrix (Harris_corners.cpp:10-29) (or NULL) */
VSLAM_API int vslam_edge_from_windows(vslam_ctx *ctx, const float *gx, const float *gy, int n, int window_len,
                                      float *responses, float *sums);

/* ---- the stage after the hot path in ./DoG ------------------------------------------
 * fp32 Gaussian kernel of cv::GaussianBlur on a CV_32F image: size cvRound(sigma*4*2+1)|1, taps = the normalised double
 * Gaussian rounded to float (getGaussianKernel(n, sigma, CV_32F)). */
VSLAM_API int vslam_gauss_ksize_f32(double sigma);
VSLAM_API int vslam_gauss_taps_f32(double sigma, int ksize, float *taps);
/* filterKeypoints' orientation stage (Diff_of_Gauss.cpp:340-369, orientationHistogram :112-133) for n keypoints of ONE
 * octave of one frame that passed the edge test: 16x16 window of the level's magnitude map blurred with sigma =
 * 1.5 * getSigmaAt(octave, level), 36-bin histogram, one output point {row, col, angle = bin * 10, 0, octave, level}
 * per bin above 0.8 of the peak, in the reference's order.  *count = number found (may exceed capacity; only the first
 * `capacity` are stored).  histograms: optional n x 36 floats. */
VSLAM_API int vslam_assign_orientations(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave,
                                        const vslam_keypoint *keypoints, int n, vslam_keypoint *out, int capacity,
                                        int *count, float *histograms);
/* SIFT (Diff_of_Gauss.cpp:561-693, rotateImageSection :528-559) for n oriented keypoints of one octave: descriptors =
 * n x 128 floats.  valid[i] = 0 where the reference's swapped (x, y) indexing leaves the padded map (it then reads
 * memory outside the image: undefined there; here such samples come from the nearest end of the map). */
VSLAM_API int vslam_sift_descriptors(vslam_ctx *ctx, const vslam_pyramid *pyr, int frame, int octave,
                                     const vslam_keypoint *oriented, int n, float *descriptors, uint8_t *valid);
/* Both stages for every frame of a pyramid in a handful of launches (one CTA per keypoint, all frames and octaves together):
 * keypoints = the hot path's lists of frames 0..n_frames-1 back to back (counts[f] each, every point read in its own
 * .octave); oriented / descriptors / valid receive up to `capacity` points in the same frame order and
 * oriented_counts[f] their number per frame (the totals are reported even when VSLAM_E_CAPACITY is returned). */
VSLAM_API int vslam_describe_keypoints(vslam_ctx *ctx, const vslam_pyramid *pyr, const vslam_keypoint *keypoints,
                                       const int *counts, vslam_keypoint *oriented, int capacity, int *oriented_counts,
                                       float *descriptors, uint8_t *valid);
/* featureDescriptors.dat as ./DoG writes it (Diff_of_Gauss.cpp:837-863): int32 count, int32 128, int32
 * sizeof(std::vector<float>) (24 on LP64), then count x 128 float32.  Host only. */
VSLAM_API int vslam_write_descriptor_file(const char *path, const float *descriptors, int n);

#ifdef __cplusplus
}
#endif
#endif /* VSLAM_H_ */
