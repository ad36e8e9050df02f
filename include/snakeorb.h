/*
 * snakeorb.h — C ABI of libsnakeorb.so: the B200 (sm_100a) implementation of Snake-SLAM's ORB front end.
 *
 * Plain C, opaque handles, caller-owned buffers, int status codes (0 = OK, <0 = snk_status), no exceptions
 * and no aborts across the boundary. One handle per host thread; handles are not thread-safe, the library is
 * re-entrant. There is no CPU fallback: every entry point fails with SNK_ERR_CUDA if no sm_100 device is usable.
 *
 * Each entry point names the reference interface it replaces. Reference paths:
 *   saiga/  = /root/reference/External/saiga/src/saiga/
 *   Snake/  = /root/reference/Snake/
 * INTEGRATION.md shows the C++ binding (snakeorb_saiga_adapter.h) a maintainer adds on the reference side.
 */
#ifndef SNAKEORB_H
#define SNAKEORB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#    define SNK_API
#else
#    define SNK_API __attribute__((visibility("default")))
#endif

typedef enum snk_status
{
    SNK_OK              = 0,
    SNK_ERR_INVALID     = -1, /* bad argument (null pointer, size mismatch, unsupported parameter) */
    SNK_ERR_CUDA        = -2, /* CUDA runtime error or no usable device; snk_last_error() has the text */
    SNK_ERR_CAPACITY    = -3, /* an output or internal buffer is too small for this input */
    SNK_ERR_UNSUPPORTED = -4
} snk_status;

/* Text of the last error raised on the calling thread (never NULL). No reference counterpart: the reference aborts
 * through SAIGA_ASSERT (saiga/core/util/assert.h) where this library returns a status. snk_version likewise. */
SNK_API const char* snk_last_error(void);
SNK_API const char* snk_version(void);

/* ------------------------------------------------------------------------------------------------------------
 * Wire types
 * ---------------------------------------------------------------------------------------------------------- */

/* == Saiga::KeyPoint<float>, saiga/vision/features/Features.h:21-67 (24 bytes, align 4) */
typedef struct snk_keypoint_f32
{
    float x, y;
    float size;
    float angle;    /* degrees, [0,360) */
    float response; /* FAST score */
    int32_t octave;
} snk_keypoint_f32;

/* == Saiga::KeyPoint<double> (Snake::KeyPoint, Snake/System/SnakeGlobal.h:53), 48 bytes, align 16 */
typedef struct snk_keypoint_f64
{
    double x, y;
    double size;
    double angle;
    double response;
    int32_t octave;
    int32_t _pad;
} snk_keypoint_f64;

/* A descriptor is 4 x uint64_t == Saiga::DescriptorORB (Features.h:70); byte i holds pattern pairs 8i..8i+7, LSB first. */
#define SNK_DESC_WORDS 4

/* ------------------------------------------------------------------------------------------------------------
 * ORB extractor   — replaces Saiga::ORBExtractor (saiga/vision/features/ORBExtractor.h:44-90, .cpp:82-289) and
 *                   Saiga::ORBExtractorGPU (saiga/cuda/imageProcessing/ORBExtractorGPU.h:38-112) as called from
 *                   Snake/Preprocess/FeatureDetector.cpp:31-41,119-124,149-154.
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct snk_orb snk_orb;

typedef struct snk_orb_params
{
    int32_t nfeatures;    /* fd_features      (ORBExtractor ctor arg 1) */
    float scale_factor;   /* fd_scale_factor  (arg 2; float, widened to double like the reference) */
    int32_t nlevels;      /* fd_levels        (arg 3), 1..16 */
    int32_t th_fast;      /* fd_iniThFAST     (arg 4) */
    int32_t th_fast_min;  /* fd_minThFAST     (arg 5) */
    int32_t max_batch;    /* images per call the handle is sized for (>=1); the reference processes 1 */
    int32_t device;       /* CUDA device ordinal */
} snk_orb_params;

/* ORBExtractor::ORBExtractor (ORBExtractor.cpp:82-88). Device buffers are allocated lazily for the first image
 * size seen (ORBExtractor::AllocatePyramid, :236-259) and re-allocated if the size changes. */
SNK_API int snk_orb_create(const snk_orb_params* params, snk_orb** out);
SNK_API void snk_orb_destroy(snk_orb* h);

/* Run all work of this handle on `cuda_stream` (a cudaStream_t; NULL = the handle's own stream). The reference's GPU
 * extractor owns its streams (ORBExtractorGPU.h:81,103-105); here the caller may supply one to order the extractor
 * against its own copies and matchers. */
SNK_API int snk_orb_set_stream(snk_orb* h, void* cuda_stream);

/* ORBExtractor::Detect (ORBExtractor.cpp:178-233): one image, results in host memory. `img` is rows x cols u8 with
 * `pitch` bytes per row, in host memory (img_is_device = 0) or device memory (1). At most `capacity` keypoints are
 * written (SNK_ERR_CAPACITY if there are more; snk_orb_max_keypoints() is always enough). Empty image => *n_out = 0. */
SNK_API int snk_orb_detect(snk_orb* h, const uint8_t* img, int rows, int cols, size_t pitch, int img_is_device,
                           snk_keypoint_f32* kps, uint64_t* desc, int capacity, int* n_out);

/* Batched Detect: n_imgs (<= max_batch) images of identical size, image i at imgs + i*img_stride. Results for image
 * i are written at kps + i*capacity / desc + 4*i*capacity, counts in n_out[i]. Frames are independent, exactly as if
 * ORBExtractor::Detect had been called n_imgs times (FeatureDetector.cpp:124,154 calls it twice per stereo frame). */
SNK_API int snk_orb_detect_batch(snk_orb* h, const uint8_t* imgs, int n_imgs, int rows, int cols, size_t pitch,
                                 size_t img_stride, int imgs_are_device, snk_keypoint_f32* kps, uint64_t* desc,
                                 int capacity, int* n_out);

/* Asynchronous halves of the call above, for pipelines that keep results on the device (the upload + kernels /
 * download split of ORBExtractorGPU::Detect, saiga/cuda/imageProcessing/ORBExtractorGPU.cu:100-192):
 * enqueue = copy-in (if host) + all kernels on the handle's stream; no host synchronisation. */
SNK_API int snk_orb_enqueue(snk_orb* h, const uint8_t* imgs, int n_imgs, int rows, int cols, size_t pitch,
                            size_t img_stride, int imgs_are_device);
/* Same, for 8-bit RGBA input (4 bytes per pixel, pitch / img_stride in bytes): Input::computeGrayscaleImage
 * (Snake/Preprocess/Input.cpp:394-413, cv::cvtColor COLOR_RGBA2GRAY) fused in front of the pyramid — the grey image
 * is written straight into the level-0 buffer: gray = (R*9798 + G*19235 + B*3735 + 16384) >> 15, alpha ignored. */
SNK_API int snk_orb_enqueue_rgba(snk_orb* h, const uint8_t* rgba, int n_imgs, int rows, int cols, size_t pitch,
                                 size_t img_stride, int imgs_are_device);
/* No reference counterpart (its extractors only return host vectors, ORBExtractor.cpp:223-232): device-resident
 * results of the last enqueue for the matchers that consume them in place: keypoints [max_batch][stride] (24-byte records), descriptors
 * [max_batch][stride][4] u64, counts [max_batch] i32. Valid until the next enqueue on this handle. */
SNK_API int snk_orb_device_results(snk_orb* h, const snk_keypoint_f32** kps_dev, const uint64_t** desc_dev,
                                   const int32_t** counts_dev, int* stride);
/* Copy the results of the last enqueue to the host (synchronises the stream): the output half of
 * ORBExtractor::Detect (ORBExtractor.cpp:223-232: keypoints and descriptors appended level by level). */
SNK_API int snk_orb_fetch(snk_orb* h, snk_keypoint_f32* kps, uint64_t* desc, int capacity, int* n_out);

/* Same copies as snk_orb_fetch, enqueued on the handle's stream without waiting for them: keypoints, descriptors and
 * counts of the last enqueue go to the caller's (pinned) host buffers and are valid once the stream has reached
 * this point (cudaStreamSynchronize / an event). Lets a caller keep several batches in flight — the asynchronous
 * counterpart of the download step of ORBExtractorGPU::Detect (saiga/cuda/imageProcessing/ORBExtractorGPU.cu:178-192).
 * capacity must be >= snk_orb_max_keypoints(). */
SNK_API int snk_orb_fetch_async(snk_orb* h, snk_keypoint_f32* kps, uint64_t* desc, int capacity, int* n_out);
/* Error check for the asynchronous path (snk_orb_enqueue + snk_orb_fetch_async never look at the kernels' error flag:
 * capacity guards of the FAST / quadtree kernels, the bounded wait on the tile loads). Synchronises the handle's stream,
 * returns SNK_ERR_CAPACITY (and clears the flag) if any batch since the last check hit one, SNK_OK otherwise. The
 * reference aborts through SAIGA_ASSERT at the equivalent places (ORBExtractor.cpp:184). */
SNK_API int snk_orb_check(snk_orb* h);
/* Test hook, no reference counterpart: writes `code` into the device-side fault flag the extraction kernels set when a
 * capacity guard or the bounded wait on a tile load trips (they cannot be provoked with valid input: the capacities are
 * proven bounds), so that the reporting paths — snk_orb_fetch, snk_orb_check — can be exercised. */
SNK_API int snk_orb_debug_set_error(snk_orb* h, int code);
/* Number of images of the last enqueue (0 before the first one / after an empty one). No reference counterpart. */
SNK_API int snk_orb_last_batch(const snk_orb* h);
/* Stream ordering for consumers of snk_orb_device_results (no reference counterpart: its extractors are synchronous).
 * snk_orb_order_stream_after: work queued on `cuda_stream` after this call starts only when everything queued on the
 * handle's stream so far (the last enqueue) has finished. snk_orb_order_after_stream: the handle's NEXT enqueue starts
 * only when everything queued on `cuda_stream` so far has finished (the consumer has read the results). Events only,
 * the host does not wait. */
SNK_API int snk_orb_order_stream_after(snk_orb* h, void* cuda_stream);
SNK_API int snk_orb_order_after_stream(snk_orb* h, void* cuda_stream);

/* Upper bound of keypoints per image for this handle (sum over levels of quota+3): what a caller
 * sizes its arrays with, where the reference lets std::vector grow (ORBExtractor.cpp:219-221 reserve, :223 insert). */
SNK_API int snk_orb_max_keypoints(const snk_orb* h);

/* ORBExtractor::getPyramid (ORBExtractor.h:61) / ScalePyramid accessors (saiga/vision/util/ScalePyramid.h:78-84):
 * per-level scale (double) and feature quota. Arrays of nlevels entries; either may be NULL. */
SNK_API int snk_orb_pyramid(const snk_orb* h, double* scale, int32_t* features);

/* ORBExtractor::GetImage(level) (ORBExtractor.h:55-59): device pointer to level `level` of image `img` of the last
 * batch (blurred = 0) or to its Gaussian-blurred copy (blurred = 1). Valid until the next enqueue. */
SNK_API int snk_orb_get_level(snk_orb* h, int img, int level, int blurred, const uint8_t** dev_ptr, int* rows,
                              int* cols, size_t* pitch);
/* Same, copied to host memory with `dst_pitch` bytes per row (test/debug helper; the reference's GetImage returns
 * a host ImageView directly, ORBExtractor.h:55-59). */
SNK_API int snk_orb_read_level(snk_orb* h, int img, int level, int blurred, uint8_t* dst, size_t dst_pitch);
/* Raw FAST candidates of one level before distribution (ORBExtractor.cpp:133-155 order), as x | y<<12 | score<<24
 * with coordinates relative to (16,16). Test/debug helper. */
SNK_API int snk_orb_read_candidates(snk_orb* h, int img, int level, uint32_t* packed, int capacity, int* n_out);

/* Per-stage device time accumulated since the last reset, in milliseconds (cudaEvents on the handle's stream).
 * stage: 0 copy-in, 1 pyramid, 2 FAST, 3 distribute, 4 blur, 5 orientation + descriptors + output records,
 * 6 reserved, 7 total.
 * Profiling costs two event records per stage; it is off by default. Counterpart of the per-module timers the
 * reference prints at exit (Snake/System/Module.cpp:78-100, FeatureDetector.cpp:75,81), per kernel stage here. */
#define SNK_ORB_NUM_STAGES 8
SNK_API int snk_orb_set_profiling(snk_orb* h, int enabled);
SNK_API int snk_orb_stage_ms(snk_orb* h, float* ms /*[SNK_ORB_NUM_STAGES]*/, int* launches /*[SNK_ORB_NUM_STAGES]*/, int reset);

/* ------------------------------------------------------------------------------------------------------------
 * Matching context (device, stream, growable scratch) shared by all matchers below. No reference counterpart: the
 * reference's matchers are stateless objects on the host (Features.h:299, SnakeORBMatcher.h); a GPU library needs
 * a place for its device scratch and its stream.
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct snk_ctx snk_ctx;
SNK_API int snk_ctx_create(int device, snk_ctx** out);
SNK_API void snk_ctx_destroy(snk_ctx* c);
SNK_API int snk_ctx_set_stream(snk_ctx* c, void* cuda_stream);

/* Saiga::distance(DescriptorORB, DescriptorORB) (saiga/vision/features/Features.h:124-138) for n pairs
 * (a[i] vs b[i]); host pointers. */
SNK_API int snk_hamming_pairs(snk_ctx* c, const uint64_t* a, const uint64_t* b, int n, int32_t* dist);

/* BruteForceMatcher<DescriptorORB>::matchKnn2 / matchKnn2_omp (Features.h:323-388): for each of the n descriptors
 * d1[i], the best and second-best of the m descriptors d2 (strict <, first minimum wins; {1000,-1} if none).
 * == the public `knn2` member (Features.h:420). Host pointers. */
typedef struct snk_knn2
{
    int32_t dist0, idx0, dist1, idx1;
} snk_knn2;
SNK_API int snk_match_knn2(snk_ctx* c, const uint64_t* d1, int n, const uint64_t* d2, int m, snk_knn2* out);
/* Batched form of the same matchKnn2 (Features.h:323-388) on device memory, one problem per keyframe pair (the
 * loops over candidate keyframes at Snake/LoopClosing/LoopORBMatcher.cpp:104-106 and
 * Initialization/MonoInitializer.cpp:587-590 call it once per pair): problem p matches d1 + 4*p*stride1 (n[p] rows) against d2 + 4*p*stride2 (m[p]);
 * n, m are device int arrays; out is device [n_problems][stride1]. Asynchronous on the context stream. */
SNK_API int snk_match_knn2_batch_device(snk_ctx* c, const uint64_t* d1_dev, const int32_t* n_dev, int stride1,
                                        const uint64_t* d2_dev, const int32_t* m_dev, int stride2, int n_problems,
                                        snk_knn2* out_dev);
/* kNN-2 matching without any device allocation. This entry point has no counterpart in the reference.
 *
 * Who it is for: callers that are not allowed to allocate GPU memory at the moment they need a brute-force match.
 * In Snake-SLAM that is the tracking thread's relocalisation / brute-force fallback (TrackBruteForce), which runs in
 * the middle of the real-time loop: snk_match_knn2 may call cudaMalloc to grow its context scratch, and cudaMalloc /
 * cudaFree stall the whole device (they wait for every running kernel, including the extractor's, on every stream)
 * and can fail outright when the rest of the system — dense mapper, viewer, other processes — has already reserved
 * the card's memory. Integrations that hand the library a fixed memory budget at start-up have the same constraint.
 *
 * What it does: the same result as snk_match_knn2 (same kernel, same tie-breaking), with the device-side staging of
 * d1, d2 and the result records placed in a fixed array inside the library instead of allocated memory. The array is
 * SNK_KNN2_NOALLOC_BYTES (320 KiB) and is held by every GPU that loads the library until the process exits.
 * Inputs and outputs are the caller's host arrays, exactly as for snk_match_knn2. It calls no cudaMalloc, cudaFree or
 * any other allocating routine.
 *
 * Limits and rules:
 *  - at most SNK_KNN2_NOALLOC_MAX (4096) descriptors per side; larger n or m is rejected with SNK_ERR_INVALID and a
 *    message naming the limit (use snk_match_knn2 there);
 *  - the first call on a device looks up the array's device address (a one-time set-up for that device); make that
 *    first call before starting any CUDA stream capture;
 *  - all calls on a device share the one array and it keeps whatever the previous call left in it. Nothing in it
 *    needs to be zero at entry (each call overwrites what it reads), but two calls whose GPU work overlaps — from
 *    two threads, or on two streams — race on it: serialise them;
 *  - `cuda_stream` is the cudaStream_t the copies and the kernel run on (NULL = the default stream); the call returns
 *    after that stream has finished this call's work. */
#define SNK_KNN2_NOALLOC_MAX 4096
#define SNK_KNN2_NOALLOC_BYTES (SNK_KNN2_NOALLOC_MAX * (32 + 32 + 16))
SNK_API int snk_match_knn2_noalloc(int device, void* cuda_stream, const uint64_t* d1, int n, const uint64_t* d2, int m,
                                   snk_knn2* out);

/* BruteForceMatcher::filterMatches (Features.h:394-414): threshold + float ratio test over knn2; writes (i, j) pairs
 * in ascending i == the public `matches` member; returns the count in *n_matches. Host pointers. */
SNK_API int snk_filter_matches(const snk_knn2* knn2, int n, int threshold, float ratio, int32_t* matches_ij,
                               int* n_matches);

/* Preprocess::StereoMatching (Snake/Preprocess/Preprocess.cpp:122-242) after the two Rectification::Forward loops
 * (:140-150): left/right are the *rectified* keypoints. */
typedef struct snk_stereo_params
{
    double bf;             /* rect_left.bf */
    int32_t relaxed;       /* settings.fd_relaxed_stereo: thresholds 75 / 0.9 / 25 deg, else 40 / 0.7 / 5 deg */
    int32_t nlevels;       /* entries of scale[] */
    double scale[16];      /* scalePyramid.Scale(level) */
} snk_stereo_params;
/* right_points / depth: nl floats each, -1000 where unmatched (Frame::allocateTmp, Snake/Map/Frame.cpp:25-26). */
SNK_API int snk_stereo_match(snk_ctx* c, const snk_stereo_params* p, const snk_keypoint_f64* left, const uint64_t* dl,
                             int nl, const snk_keypoint_f64* right, const uint64_t* dr, int nr, float* right_points,
                             float* depth, int* n_matches);
/* Batched, device-resident form of the same Preprocess::StereoMatching body (Preprocess.cpp:152-239), fed directly by
 * an extractor handle: images 2p and 2p+1 of the last enqueue are the
 * left / right image of stereo pair p (identity rectification, i.e. EuRoC/KITTI rectified input). Outputs are device
 * arrays [n_pairs][stride] (stride from snk_orb_device_results) and n_matches_dev[n_pairs]; 2 * n_pairs must not
 * exceed the image count of the last enqueue. Asynchronous: the kernel runs on the CONTEXT's stream, ordered by events
 * after the extractor stream's last enqueue and before its next one (no host synchronisation); the outputs are valid
 * once the context's stream has reached this point. Device-side faults show up in snk_ctx_check. */
SNK_API int snk_stereo_match_orb_device(snk_ctx* c, snk_orb* h, int n_pairs, const snk_stereo_params* p,
                                        float* right_points_dev, float* depth_dev, int32_t* n_matches_dev);

/* The same two orderings for a context's stream: call snk_ctx_wait_orb before handing snk_orb_device_results pointers
 * to snk_match_knn2_batch_device / snk_bow_transform_batch_device / snk_bow_score_batch (they run on the context's
 * stream and take raw pointers, so they cannot order themselves), and snk_orb_wait_ctx before the handle's next
 * enqueue. Both are no-ops when the two handles share one stream (snk_orb_set_stream / snk_ctx_set_stream). */
SNK_API int snk_ctx_wait_orb(snk_ctx* c, snk_orb* h);
SNK_API int snk_orb_wait_ctx(snk_orb* h, snk_ctx* c);
/* Synchronises the context's stream and reports (then clears) device-side faults of the asynchronous matcher entry
 * points (SNK_ERR_CAPACITY), so that they do not surface in a later unrelated call. */
SNK_API int snk_ctx_check(snk_ctx* c);

/* Preprocess::ComputeStereoFromRGBD (Snake/Preprocess/Preprocess.cpp:79-120): per undistorted keypoint, re-distort
 * with the depth camera's model (saiga/vision/cameraModel/Distortion.h:128-170), look the depth up at the nearest
 * pixel (int(v + 0.5)) and derive the virtual right-image coordinate x - bf/depth. right_points / depth get -1 where
 * the depth image holds no measurement (<= 0). The reference asserts that the lookup lands inside the depth image;
 * here such keypoints are reported like "no measurement". depth_img is a host float image, pitch in bytes. */
typedef struct snk_rgbd_params
{
    double fx, fy, cx, cy;        /* K (colour camera; unproject2 ignores the skew, Intrinsics4.h:80) */
    double dfx, dfy, dcx, dcy, ds; /* rgbd_intrinsics.depthModel.K */
    double dist[8];               /* rgbd_intrinsics.depthModel.dis: k1 k2 k3 k4 k5 k6 p1 p2 */
    double bf;                    /* rgbd_intrinsics.bf */
} snk_rgbd_params;
SNK_API int snk_stereo_from_rgbd(snk_ctx* c, const snk_rgbd_params* p, const snk_keypoint_f64* kp_un, int n,
                                 const float* depth_img, int rows, int cols, size_t pitch, float* right_points,
                                 float* depth, int* n_matches);

/* Preprocess::undistortKeypoints (Snake/Preprocess/Preprocess.cpp:55-77): per keypoint K_src.unproject2
 * (saiga/vision/cameraModel/Intrinsics4.h:78), undistortPointGN with the point as its own start (5 Gauss-Newton
 * steps on distortNormalizedPoint and its Jacobian, 2x2 Eigen LDLT solve; Distortion.h:128-285), rotation by the
 * rectification quaternion (Rectify.h:15-33), dehomogenisation -> normalized_points, K_dst.normalizedToImage ->
 * undistorted keypoint. kp_un[i] is kps[i] with the point replaced. Double arithmetic in the reference's operation
 * order, unfused. */
typedef struct snk_rectify
{
    double fx, fy, cx, cy;         /* rect_left.K_src (unproject2 ignores the skew) */
    double dist[8];                /* rect_left.D_src: k1 k2 k3 k4 k5 k6 p1 p2 */
    double quat[4];                /* rect_left.R as x y z w (identity: 0 0 0 1) */
    double dfx, dfy, dcx, dcy, ds; /* rect_left.K_dst */
} snk_rectify;
SNK_API int snk_undistort_keypoints(snk_ctx* c, const snk_rectify* r, const snk_keypoint_f64* kps, int n,
                                    snk_keypoint_f64* kp_un, double* normalized_points /* 2n */);

/* The whole of Preprocess::StereoMatching (Snake/Preprocess/Preprocess.cpp:122-242) including its two
 * Rectification::Forward loops (saiga/vision/cameraModel/Rectify.cpp:18-25: unproject, undistortPointGN, rotate,
 * project with K_dst — the per-point math of snk_undistort_keypoints) for the RAW left / right keypoints; everything
 * stays on the device between the steps. snk_rectify as in snk_undistort_keypoints, one per camera. */
SNK_API int snk_stereo_matching(snk_ctx* c, const snk_rectify* rect_left, const snk_rectify* rect_right,
                                const snk_stereo_params* p, const snk_keypoint_f64* left, const uint64_t* dl, int nl,
                                const snk_keypoint_f64* right, const uint64_t* dr, int nr, float* right_points,
                                float* depth, int* n_matches);

/* Camera, grid and frame views used by the projection searches. */
typedef struct snk_camera
{
    double fx, fy, cx, cy, s; /* Saiga::IntrinsicsPinholed (saiga/vision/cameraModel/Intrinsics4.h) — global K */
    double bf;                /* stereo_cam.bf (Intrinsics4.h:214) */
} snk_camera;

/* Saiga::FeatureGridBounds2<double,20> + FeatureGrid2 (saiga/vision/util/FeatureGrid2.h:10-211): cells[row*cols+col]
 * = (begin,end) index range into the (already grid-permuted) keypoint arrays. */
typedef struct snk_grid_view
{
    double bmin_x, bmin_y, bmax_x, bmax_y;
    double cell_size_inv_x, cell_size_inv_y;
    int32_t rows, cols;
    const int32_t* cells; /* 2*rows*cols ints */
} snk_grid_view;

/* The parts of Snake::Frame / Snake::Features the searches read (Snake/Map/Features.h:18-41, Frame.h). */
typedef struct snk_frame_view
{
    int32_t n;
    const snk_keypoint_f64* kp_un;   /* undistorted_keypoints */
    const uint64_t* desc;            /* descriptors */
    const float* right_points;       /* right_points */
    const int32_t* kp_to_point;      /* mvpMapPoints as ids: >=0 already matched, -1 free */
    double pose_q[4];                /* Frame::Pose() unit quaternion x,y,z,w */
    double pose_t[3];
    snk_grid_view grid;
    const float* depth;              /* Features::depth (hasDepth(i) == depth[i] > 0); read by snk_fuse_search only,
                                        may be NULL elsewhere */
} snk_frame_view;

/* Snake::FineTrackingPoint (Snake/Map/LocalMap.h:32-55) */
typedef struct snk_point_fine
{
    uint64_t desc[4];
    double pos[3];
    double normal[3];
    float ref_depth;
    int32_t ref_level;
    int32_t valid;
    int32_t id; /* stands for MapPoint* */
} snk_point_fine;

/* Snake::CoarseTrackingPoint (Snake/Map/LocalMap.h:17-30) */
typedef struct snk_point_coarse
{
    uint64_t desc[4];
    double pos[3];
    double normal[3];
    int32_t octave;
    float angle;
    int32_t id;
    int32_t _pad;
} snk_point_coarse;

typedef struct snk_scale_pyramid
{
    int32_t nlevels;
    int32_t _pad;
    double scale_factor; /* ScalePyramid::scale_factor (float 1.2f widened) */
    double scale[16];
} snk_scale_pyramid;

/* FeatureGridBounds2<double,20>::computeFromIntrinsicsDist (saiga/vision/util/FeatureGrid2.h:63-104), called once
 * at start-up (Snake/Preprocess/Input.cpp:93): the bounding box of the undistorted image border (resolution + 1
 * points per side, undistortAll of Distortion.h:326-338), widened by 0.1, and the 20-px cell grid over it. Fills every
 * field of *out except `cells`. Host-side helper, no GPU involved; dist = k1 k2 k3 k4 k5 k6 p1 p2. */
SNK_API int snk_grid_bounds(int w, int h, const snk_camera* K, const double* dist, int resolution, snk_grid_view* out);

/* Preprocess::computeFeatureGrid (Snake/Preprocess/Preprocess.cpp:244-266) == FeatureGrid2::create
 * (saiga/vision/util/FeatureGrid2.h:130-192): counting sort of the undistorted keypoints into the 20-px cells of
 * `bounds` (its `cells` member is ignored). permutation[i] = new index of keypoint i (keypoints outside the bounds
 * go to the end, input order is kept inside a cell); cells_out[2*(row*cols+col)] = (begin, end) of the cell in the
 * permuted arrays. The caller applies the permutation to its keypoint / descriptor arrays as the reference does
 * (Preprocess.cpp:254-260). n <= 8192. */
SNK_API int snk_feature_grid(snk_ctx* c, const snk_grid_view* bounds, const snk_keypoint_f64* kp_un, int n,
                             int32_t* permutation, int32_t* cells_out);

/* SnakeORBMatcher::SearchByProjection2 (Snake/Tracking/SnakeORBMatcher.cpp:365-526).
 * In/out: points[i].valid is cleared where the reference clears lmp.valid; visible[i] = 1 where it calls
 * IncreaseVisible(). Out: kp_to_point_out[n] = frame.kp_to_point with new assignments (point ids) added,
 * best_idx/best_dist [n_points] = tmp_matches_world_image (-1 if none). Returns matches in *n_matches. */
SNK_API int snk_search_projection_fine(snk_ctx* c, const snk_frame_view* frame, snk_point_fine* points, int n_points,
                                       const snk_camera* cam, const snk_scale_pyramid* pyr, float th, float ratio,
                                       int32_t* kp_to_point_out, int32_t* best_idx, int32_t* best_dist,
                                       uint8_t* visible, int* n_matches);

/* The same SearchByProjection2 (SnakeORBMatcher.cpp:365-526) on data that STAYS on the device between calls. In the
 * reference the frame (Snake/Map/Frame.h) and the local map (LocalMap<FineTrackingPoint>, Snake/Map/LocalMap.h:83-186,
 * rebuilt by TrackingFine::updateLocalMap once per frame) live in host memory and the search runs where they are; the
 * host-struct entry point above uploads both on every call (1.9 MB for 20 000 points), which is most of its time.
 *   snk_frame_dev_*  : a device copy of a frame view (keypoints, descriptors, right_points, assignments, grid); the pose
 *                      can be changed (pose refinement runs between the coarse and the fine search), the assignments
 *                      follow the searches (keep_assignment) or are set from the host.
 *   snk_localmap_*   : capacity points on the device; snk_localmap_upload replaces points [first, first + n) and sets the
 *                      number of points in use (new_size).
 *   snk_search_projection_fine_resident: the search; lm point `valid` flags are updated on the device as the reference
 *   updates lmp.valid. Host outputs (each may be NULL): kp_to_point_out[n keypoints], point_valid_out[n points] (int32),
 *   visible_out[n points] (1 where the reference calls IncreaseVisible). Synchronous. */
typedef struct snk_frame_dev snk_frame_dev;
typedef struct snk_localmap snk_localmap;
SNK_API int snk_frame_dev_create(snk_ctx* c, const snk_frame_view* host_view, snk_frame_dev** out);
SNK_API void snk_frame_dev_destroy(snk_frame_dev* f);
SNK_API int snk_frame_dev_set_pose(snk_frame_dev* f, const double* q_xyzw, const double* t);
SNK_API int snk_frame_dev_set_assignment(snk_frame_dev* f, const int32_t* kp_to_point);
SNK_API int snk_localmap_create(snk_ctx* c, int capacity, snk_localmap** out);
SNK_API void snk_localmap_destroy(snk_localmap* lm);
SNK_API int snk_localmap_upload(snk_localmap* lm, const snk_point_fine* points, int first, int n, int new_size);
SNK_API int snk_search_projection_fine_resident(snk_ctx* c, snk_frame_dev* frame, snk_localmap* lm, const snk_camera* cam,
                                                const snk_scale_pyramid* pyr, float th, float ratio, int keep_assignment,
                                                int32_t* kp_to_point_out, int32_t* point_valid_out, uint8_t* visible_out,
                                                int* n_matches);

/* SnakeORBMatcher::SearchByProjectionFrameFrame2 (Snake/Tracking/SnakeORBMatcher.cpp:191-354).
 * motion: 0 neither, 1 bForward, 2 bBackward (the caller evaluates :209-212). */
SNK_API int snk_search_projection_coarse(snk_ctx* c, const snk_frame_view* frame, const snk_point_coarse* points,
                                         int n_points, const snk_camera* cam, const snk_scale_pyramid* pyr, float th,
                                         int feature_error, int motion, int32_t* kp_to_point_out, int* n_matches);

/* SnakeORBMatcher::SearchByProjectionFrameToKeyframe (Snake/Tracking/SnakeORBMatcher.cpp:71-188): fully serial in
 * the reference (each match removes a keypoint for later points); points with id < 0 are skipped (null MapPoint),
 * ids already present in frame.kp_to_point are skipped (sAlreadyFound). */
typedef struct snk_point_kf
{
    uint64_t desc[4];
    double pos[3];
    int32_t id;
    int32_t _pad;
} snk_point_kf;
SNK_API int snk_search_projection_frame_to_kf(snk_ctx* c, const snk_frame_view* frame, const snk_point_kf* points,
                                              int n_points, const snk_camera* cam, float th, int feature_error,
                                              int32_t* kp_to_point_out, int* n_matches);

/* LoopORBmatcher::MatchBoW (Snake/LoopClosing/LoopORBMatcher.cpp:121-215). A FeatureVector
 * (saiga/vision/slam/MiniBow2.h:98-132) is given as CSR: node_ids[n_nodes] ascending, offsets[n_nodes+1], indices[]. */
typedef struct snk_featvec_view
{
    int32_t n_nodes;
    const int32_t* node_ids;
    const int32_t* offsets;
    const int32_t* indices;
} snk_featvec_view;
/* valid1/valid2: 1 where the keyframe has a good map point at that feature (vpMapPoints[i] && !isBad()).
 * match12[n1] = matched feature index in keyframe 2, or -1. */
SNK_API int snk_match_bow(snk_ctx* c, const uint64_t* d1, int n1, const uint64_t* d2, int n2,
                          const snk_featvec_view* fv1, const snk_featvec_view* fv2, const uint8_t* valid1,
                          const uint8_t* valid2, int th, float ratio, int32_t* match12, int* n_matches);

/* One keyframe against many: the loop LoopDetector::ComputeSim3(new_kf) runs over candidateKFs
 * (Snake/LoopClosing/LoopDetector.cpp:384-391 -> :225) calls MatchBoW(new_kf, candidate, ...) once per candidate, and
 * a call reads nothing an earlier one wrote, so all of them can be issued at once: one upload, one launch (a warp per
 * (node of keyframe 1, candidate)), one download. match12 is n_cand x n1, row k = the match12 snk_match_bow returns
 * for candidate k; n_matches[n_cand]. */
typedef struct snk_bow_candidate
{
    const uint64_t* desc; /* n x 4 */
    int32_t n;
    const snk_featvec_view* fv;
    const uint8_t* valid; /* n */
} snk_bow_candidate;
SNK_API int snk_match_bow_batch(snk_ctx* c, const uint64_t* d1, int n1, const snk_featvec_view* fv1,
                                const uint8_t* valid1, const snk_bow_candidate* cands, int n_cand, int th, float ratio,
                                int32_t* match12, int32_t* n_matches);

/* ---- local-mapping and loop-closing searches -------------------------------------------------
 * All of them are independent per point / feature in the reference (its tmp_flags are never set), so the device
 * evaluates every point at once and the results are listed in the reference's iteration order. */

/* Snake::FusionPoint (Snake/Map/LocalMap.h:57-79) */
typedef struct snk_point_fusion
{
    uint64_t desc[4];
    double pos[3];
    double normal[3];
    float ref_depth;
    int32_t ref_level;
    int32_t observations;
    int32_t id;
} snk_point_fusion;

/* MappingORBMatcher::Fuse(kf, pose, point_mask, LocalMap<FusionPoint>, fuseCandidates, th, obs_factor, feature_th)
 * (Snake/LocalMapping/MappingORBMatcher.cpp:359-482, built with MATCHING_MIN_MAX_DISTANCE2 and
 * MATCHING_CHECK_SCALE_CONSISTENCY2, SnakeGlobal.h:168-169). kf->pose_q/pose_t is the `pose` argument, kf->depth is
 * required, point_mask may be NULL. fuseCandidates = (cand_kp[j], cand_id[j]), j < *n_candidates, in point order;
 * both arrays need n_points entries. */
SNK_API int snk_fuse_search(snk_ctx* c, const snk_frame_view* kf, const uint8_t* point_mask,
                            const snk_point_fusion* points, int n_points, const snk_camera* cam,
                            const snk_scale_pyramid* pyr, float th, float obs_factor, int feature_th, int32_t* cand_kp,
                            int32_t* cand_id, int* n_candidates);

/* The search of LoopORBmatcher::Fuse (Snake/LoopClosing/LoopORBMatcher.cpp:13-72): skip[i] = mp->isBad() ||
 * mp->IsInKeyFrame(pKF) evaluated by the caller; best_idx[i] / best_dist[i] = the most similar keypoint within `th`
 * pixels of the projection (-1 / 256 when the point is skipped, culled or has no candidate). The caller then applies
 * :75-93 in point order to the points with best_dist <= TH_LOW (the map edits there do not feed back into the
 * search: a replaced point is bad and was in the keyframe). ref_depth / ref_level / observations are not read. */
SNK_API int snk_fuse_search_loop(snk_ctx* c, const snk_frame_view* kf, const uint8_t* skip,
                                 const snk_point_fusion* points, int n_points, const snk_camera* cam, float th,
                                 int32_t* best_idx, int32_t* best_dist);

/* MappingORBMatcher::SearchForTriangulation2 (MappingORBMatcher.cpp:14-101): for every feature of kf1 without a map
 * point, the best feature of kf2 under the same FeatureVector node (no map point, distance <= featureDistance and
 * <= the running best starting at TH_LOW = 50 — the LAST minimum wins — and squared epipolar distance below
 * (2*epipolarDistance/fx)^2; Saiga::EpipolarDistanceSquared, saiga/vision/reconstruction/Epipolar.cpp:47-56).
 * kfX->kp_to_point[i] >= 0 stands for GetMapPoint(i) != null; npX = normalized_points (2 doubles per feature);
 * E row-major. pairs = (idx1, idx2) in the reference's order, room for 2*kf1->n ints. Only n, desc, kp_to_point of
 * the frame views are read. */
SNK_API int snk_search_triangulation_bow(snk_ctx* c, const snk_frame_view* kf1, const double* np1,
                                         const snk_featvec_view* fv1, const snk_frame_view* kf2, const double* np2,
                                         const snk_featvec_view* fv2, const double* E, const snk_camera* cam,
                                         float epipolar_distance, int feature_distance, int32_t* pairs, int* n_pairs);

/* MappingORBMatcher::SearchForTriangulationProject (MappingORBMatcher.cpp:167-247): every feature of kf1 without a
 * map point is unprojected with the depth of its coarse grid cell (depth_grid(cell.y/4, cell.x/4), row-major
 * grid_rows x grid_cols), moved into kf2 (kf1->pose inverted, kf2->pose) and searched within 20 px (epipolar gate
 * (epipolarDistance/fx)^2, then distance, last minimum wins). kf1->grid supplies featureGridBounds. */
SNK_API int snk_search_triangulation_project(snk_ctx* c, const snk_frame_view* kf1, const double* np1,
                                             const snk_frame_view* kf2, const double* np2, const double* depth_grid,
                                             int grid_rows, int grid_cols, const double* E12, const snk_camera* cam,
                                             float epipolar_distance, int feature_distance, int32_t* pairs,
                                             int* n_pairs);

/* MonoInitializer::ComputeMatchesPrediction (Snake/Tracking/Initialization/MonoInitializer.cpp:619-697): feature i
 * of the left frame searches the right frame within `radius` (50) of last_positions[i]; first minimum wins; a match
 * needs distance < feature_error (50). matches[i] = index in the right frame or -1; last_positions[i] is moved onto
 * the matched keypoint. Only grid, kp_un and desc of `right` are read. */
SNK_API int snk_match_prediction(snk_ctx* c, const uint64_t* desc_left, int n_left, double* last_positions,
                                 const snk_frame_view* right, float radius, int feature_error, int32_t* matches,
                                 int* n_matches);

/* MiniBow2::TemplatedVocabulary (saiga/vision/slam/MiniBow2.h): the vocabulary tree on the device.
 * snk_voc_load_raw reads the file format of loadRaw/saveRaw (:1020-1075: k, L, two unused ints, u64 node count,
 * per node {i32 id, i32 parent, f64 weight, i32 word_id, 32-byte descriptor}, then the word table, which is
 * redundant and ignored). snk_voc_create takes the same per-node arrays (node 0 = root, node id = array index,
 * children in index order — what loadRaw builds). Fails like loadRaw when the file cannot be opened. */
typedef struct snk_voc snk_voc;
SNK_API int snk_voc_load_raw(int device, const char* path, snk_voc** out);
SNK_API int snk_voc_create(int device, int k, int L, int n_nodes, const int32_t* parent, const double* weight,
                           const int32_t* word_id, const uint64_t* desc, snk_voc** out);
SNK_API void snk_voc_destroy(snk_voc* v);
SNK_API int snk_voc_info(const snk_voc* v, int* k, int* L, int* n_nodes, int* n_words);

/* TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup) (MiniBow2.h:883-983), as
 * Frame::computeBoW calls it. BowVector: (word id, L1-normalised weight) ascending in bow_ids / bow_vals;
 * FeatureVector: CSR as in snk_featvec_view (node ids ascending, offsets[n_nodes + 1], feature indices in the
 * reference's list order — the order its unstable std::sort leaves, reproduced exactly). Features that land on a
 * word of weight <= 0 are in neither. Output arrays need n entries (fv_offsets n + 1). n <= 8192. */
#define SNK_BOW_MAX_FEATURES 8192
SNK_API int snk_bow_transform(snk_ctx* c, const snk_voc* voc, const uint64_t* desc, int n, int levelsup,
                              int32_t* bow_ids, float* bow_vals, int* n_bow, int32_t* fv_node_ids,
                              int32_t* fv_offsets, int32_t* fv_indices, int* n_fv_nodes);

/* Batched, device-resident form of the same transform (MiniBow2.h:883-931; Frame::computeBoW, Snake/Map/Frame.cpp:36-42,
 * once per frame) — e.g. straight from snk_orb_device_results: desc_dev [n_frames][stride][4],
 * counts_dev [n_frames], stride <= 8192. Outputs are device arrays: bow_ids / bow_vals / fv_node_ids / fv_indices
 * [n_frames][stride], fv_offsets [n_frames][stride + 1], out_counts [n_frames][2] = (words, nodes). Asynchronous on
 * the context's stream. */
SNK_API int snk_bow_transform_batch_device(snk_ctx* c, const snk_voc* voc, const uint64_t* desc_dev,
                                           const int32_t* counts_dev, int n_frames, int stride, int levelsup,
                                           int32_t* bow_ids_dev, float* bow_vals_dev, int32_t* fv_node_ids_dev,
                                           int32_t* fv_offsets_dev, int32_t* fv_indices_dev, int32_t* out_counts_dev);

/* FeatureVector::score (MiniBow2.h:134-170; TemplatedVocabulary::score, :283): the L1 similarity of one query
 * BowVector against n_cand others (keyframe-database queries of loop detection and relocalisation). Candidates are
 * concatenated: candidate k owns cand_ids / cand_vals [cand_offsets[k], cand_offsets[k+1]); ids ascending inside
 * every vector, as snk_bow_transform produces them. scores[k] has the reference's float rounding. */
SNK_API int snk_bow_score_batch(snk_ctx* c, const int32_t* q_ids, const float* q_vals, int nq, const int32_t* cand_ids,
                                const float* cand_vals, const int32_t* cand_offsets, int n_cand, float* scores);

/* The reference's feature cache (Snake/Preprocess/FeatureDetector.cpp:97-111 read, :134-139 write; settings
 * fd_bufferToFile), a saiga BinaryFile dump (saiga/core/util/BinaryFile.h:30-78): u64 n, n x KeyPoint<double>
 * (48-byte records), u64 n, n x DescriptorORB (32 bytes). Host-side helpers, no GPU involved. Also the format of
 * saiga/data/vision/{0,50}.features. */
SNK_API int snk_features_write(const char* path, const snk_keypoint_f64* kps, const uint64_t* desc, int n);
/* The read side (FeatureDetector.cpp:101-102, `bf >> frame.keypoints >> frame.descriptors`). Returns
 * SNK_ERR_CAPACITY (and the file's feature count in *n_out) when capacity is too small. */
SNK_API int snk_features_read(const char* path, snk_keypoint_f64* kps, uint64_t* desc, int capacity, int* n_out);
/* KeyPoint<float>::cast<double>() as FeatureDetector.cpp:129-132 applies it to the extractor's output. */
SNK_API int snk_keypoints_widen(const snk_keypoint_f32* in, int n, snk_keypoint_f64* out);

/* Input::computeGrayscaleImage (Snake/Preprocess/Input.cpp:394-413) alone (host RGBA in, host grey out), for callers
 * that keep the grey image. */
SNK_API int snk_rgba_to_gray(snk_ctx* c, const uint8_t* rgba, int rows, int cols, size_t pitch, uint8_t* gray,
                             size_t gray_pitch);

/* Register-resident popc microbenchmark: measured 256-bit Hamming distances per second of this device, the
 * denominator of the matcher roofline. Measurement helper, no reference counterpart. */
SNK_API int snk_hamming_peak_selftest(snk_ctx* c, double* pair_ops_per_s);

#ifdef __cplusplus
}
#endif
#endif /* SNAKEORB_H */
