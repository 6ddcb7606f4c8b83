/*
 * svo_gpu.h -- C ABI of the B200 (sm_100a) feature front-end: pyramid build, FAST-9/16, grid bucketing, pyramidal LK.
 *
 * This is the ONLY way host code reaches the CUDA kernels (libsvo_gpu.so: plain pointers and sizes, no C++ or torch
 * types).  Each entry point names the reference interface (ikryukov/svo, paths relative to the reference tree) it stands
 * in for.  There is no CPU fallback: every call either runs the sm_100a kernels or returns a negative SVO_GPU_E_* code.
 *
 * Conventions
 *   - return value: 0 = ok, negative = error (svo_gpu_last_error(ctx) holds a message).
 *   - host pointers are owned by the caller and never retained after the call returns; outputs are caller-allocated
 *     with a stated capacity; a too-small capacity returns SVO_GPU_E_CAPACITY with the required count written,
 *     never a silently truncated list.
 *   - every call is synchronous (outputs valid on return), like the OpenCV calls the reference makes
 *     (src/tracking.cpp:82, :101-105, :160-165 consume their results immediately).
 *   - a context is bound to one device and is NOT thread-safe; use one context per host thread per device.
 *   - "slot" = a persistent device-side image pyramid (0 <= slot < SVO_GPU_NUM_SLOTS); "lane" = one independent image
 *     stream inside a batched call (0 <= lane < max_batch).  The single-image calls work on lane 0.
 */
#ifndef SVO_GPU_H
#define SVO_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define SVO_GPU_API __declspec(dllexport)
#else
#define SVO_GPU_API __attribute__((visibility("default")))
#endif

#define SVO_GPU_OK 0
#define SVO_GPU_E_ARG (-1)         /* bad argument (null pointer, size out of range, point outside the bucket table ...) */
#define SVO_GPU_E_CAPACITY (-2)    /* output capacity too small; the required count has been written to *count_out */
#define SVO_GPU_E_CUDA (-3)        /* a CUDA runtime call or kernel failed */
#define SVO_GPU_E_UNSUPPORTED (-4) /* valid for OpenCV, not implemented here (OPTFLOW_USE_INITIAL_FLOW, window > 45) */
#define SVO_GPU_E_STATE (-5)       /* slot not uploaded / pyramid not built (or built for a smaller window) */

#define SVO_GPU_NUM_SLOTS 5
#define SVO_GPU_STEP_DEPTH 2   /* steps that may be outstanding between svo_gpu_step_submit and svo_gpu_step_wait */
#define SVO_GPU_MAX_LEVELS 8       /* pyramid levels 0..7 (OpenCV maxLevel <= 7) */
#define SVO_GPU_MAX_WIN 45         /* largest LK window side */

/* cv::TermCriteria::Type and cv::OPTFLOW_* flag values, so callers can pass OpenCV's constants straight through */
#define SVO_GPU_TC_COUNT 1
#define SVO_GPU_TC_EPS 2
#define SVO_GPU_LK_USE_INITIAL_FLOW 4
#define SVO_GPU_LK_GET_MIN_EIGENVALS 8

/* memory kind of an image/points pointer handed to the *_batch / step calls */
#define SVO_GPU_MEM_HOST 0
#define SVO_GPU_MEM_DEVICE 1

typedef struct svo_gpu_ctx svo_gpu_ctx;

/* ---------------------------------------------------------------------------------------------- context ------- */

/* Creates a context on CUDA device `device`: one stream, device buffers for SVO_GPU_NUM_SLOTS pyramids of up to
 * max_w x max_h pixels and max_levels levels per lane, keypoint/track buffers of max_points entries per lane, and
 * pinned staging.  Replaces nothing in the reference (which has no device); it is the state that
 * Tracking::Tracking (src/tracking.cpp:22-59) would create next to mDetector. */
SVO_GPU_API int svo_gpu_create(svo_gpu_ctx** out, int device, int max_w, int max_h, int max_levels, int max_points,
                               int max_batch);
SVO_GPU_API int svo_gpu_destroy(svo_gpu_ctx* ctx);
/* message of the last failing call on this context ("" if none); with ctx == NULL, of the last failed create */
SVO_GPU_API const char* svo_gpu_last_error(const svo_gpu_ctx* ctx);
/* the cudaStream_t all work of this context is enqueued on (as void*), for callers that time with CUDA events */
SVO_GPU_API void* svo_gpu_stream(svo_gpu_ctx* ctx);
/* ABI version of this header */
SVO_GPU_API int svo_gpu_abi_version(void);

/* Page-locked host memory for frames and result arrays that should move while kernels run (cudaMemcpyAsync only overlaps
 * from pinned memory).  Plays the part of the cv::Mat buffers the reference's loader thread fills ahead of time
 * (include/async_image_loader.h:27-33).  Not tied to a context; free with svo_gpu_host_free. */
SVO_GPU_API int svo_gpu_host_alloc(void** out, size_t bytes);
SVO_GPU_API int svo_gpu_host_free(void* p);

/* ---------------------------------------------------------------------------------------------- images -------- */

/* Copies one 8-bit gray image (cv::Mat CV_8UC1: data, cols, rows, step) into `slot`, lane 0.
 * Stands in for handing frame->leftImg()/rightImg() (include/frame.h:32-33) to an OpenCV call. */
SVO_GPU_API int svo_gpu_upload(svo_gpu_ctx* ctx, int slot, const uint8_t* host, int w, int h, size_t stride_bytes);
/* Same for `batch` images (lane b at images + b*image_stride_bytes); mem_kind says whether `images` is host or device. */
SVO_GPU_API int svo_gpu_upload_batch(svo_gpu_ctx* ctx, int slot, const uint8_t* images, int mem_kind, int w, int h,
                                     size_t stride_bytes, size_t image_stride_bytes, int batch);

/* Colour frames: `host` / `images` are interleaved 8-bit BGR (cv::Mat CV_8UC3 as cv::imread returns it; stride_bytes >=
 * 3 w).  The slot receives cv::cvtColor(img, COLOR_BGR2GRAY) of them, bit for bit (OpenCV 4.x 8-bit fixed point), computed
 * on the device.  Replaces the host-side conversion of the reference's loader (include/async_image_loader.h:66-69: imread,
 * then cvtColor when the sequence is the colour one). */
SVO_GPU_API int svo_gpu_upload_bgr(svo_gpu_ctx* ctx, int slot, const uint8_t* host, int w, int h, size_t stride_bytes);
SVO_GPU_API int svo_gpu_upload_bgr_batch(svo_gpu_ctx* ctx, int slot, const uint8_t* images, int mem_kind, int w, int h,
                                         size_t stride_bytes, size_t image_stride_bytes, int batch);
/* Read-back of the 8-bit gray image lane `lane` of `slot` holds (rows stride_bytes apart): what the reference keeps as
 * frame->leftImg() / rightImg() (include/frame.h:32-33) when the frame arrived as BGR. */
SVO_GPU_API int svo_gpu_download_image(svo_gpu_ctx* ctx, int slot, int lane, uint8_t* host, size_t stride_bytes);

/* Builds the Gaussian pyramid of `slot` (all uploaded lanes): level l = pyrDown(level l-1), every level stored with a
 * BORDER_REFLECT_101 margin.  Same result as cv::buildOpticalFlowPyramid(img, pyr, winSize, maxLevel, false), which
 * cv::calcOpticalFlowPyrLK runs on both images on every call (src/tracking.cpp:101-105, :160-165); *levels_out gets the
 * effective maxLevel (construction stops when a level would be <= winSize). */
SVO_GPU_API int svo_gpu_build_pyramid(svo_gpu_ctx* ctx, int slot, int win_w, int win_h, int max_level, int* levels_out);
/* Pure helpers for callers that keep pyramids across calls (the wrappers' frame cache, svo_b200/host/feature.cpp):
 * the effective maxLevel cv::buildOpticalFlowPyramid would return for this image / window / maxLevel, the margin a
 * pyramid built for this window carries, and what `slot` holds now (*levels_out = levels built, 0 = no pyramid).
 * svo_gpu_klt accepts any two pyramids of equal margin that hold the levels it needs: level l of an image does not
 * depend on the window the pyramid was built for. */
SVO_GPU_API int svo_gpu_pyramid_levels(int w, int h, int win_w, int win_h, int max_level);
SVO_GPU_API int svo_gpu_pyramid_margin(int win_w, int win_h);
SVO_GPU_API int svo_gpu_slot_pyramid(const svo_gpu_ctx* ctx, int slot, int* levels_out, int* margin_out);
/* How many images `slot` has received so far (every upload call and every svo_gpu_step ingest into it counts).  A caller that
 * remembers "slot s holds frame F" -- the wrappers' frame cache -- records this number with the tag and treats a different
 * number as "someone else has written the slot since": the context is reachable by other code (svo::defaultContext()). */
SVO_GPU_API int svo_gpu_slot_generation(const svo_gpu_ctx* ctx, int slot, unsigned long long* generation_out);
/* Test/debug read-back of one pyramid level of lane `lane` (tightly packed w*h bytes). */
SVO_GPU_API int svo_gpu_download_level(svo_gpu_ctx* ctx, int slot, int lane, int level, uint8_t* host, int capacity_bytes,
                                       int* w_out, int* h_out);
/* The same level WITH its BORDER_REFLECT_101 margin, as cv::buildOpticalFlowPyramid stores it (the margin here is
 * *margin_out >= the window it was built for): host receives (h + 2 margin) rows of (w + 2 margin) bytes. */
SVO_GPU_API int svo_gpu_download_level_padded(svo_gpu_ctx* ctx, int slot, int lane, int level, uint8_t* host,
                                              int capacity_bytes, int* w_out, int* h_out, int* margin_out);

/* ---------------------------------------------------------------------------------------------- FAST ---------- */

/* FAST-9/16 + optional 3x3 non-max suppression + optional pixel mask on level 0 of `slot`, lane 0; keypoints in raster
 * order.  Replaces mDetector->detect(frame->leftImg(), keypoints, mask) (src/tracking.cpp:82) for the detector built by
 * cv::FastFeatureDetector::create(threshold, nonMaxSuppression) (src/tracking.cpp:54-57).
 *   mask_or_null: host CV_8UC1 mask of the image size (0 = reject), as built at src/tracking.cpp:76-79, or NULL.
 *   xy_out[2*i], xy_out[2*i+1] = KeyPoint::pt; response_out[i] = KeyPoint::response (score with NMS, 0 without).
 *   The other KeyPoint fields are constants: size 7, angle -1, octave 0, class_id -1.
 *   Entries of xy_out / response_out beyond *count_out (up to capacity) are unspecified: the list leaves the device
 *   together with its length, sized by the previous call's result, so that the usual call synchronises once. */
SVO_GPU_API int svo_gpu_fast(svo_gpu_ctx* ctx, int slot, int threshold, int nms, const uint8_t* mask_or_null,
                             size_t mask_stride, float* xy_out, float* response_out, int capacity, int* count_out);

/* The reference's "do not re-detect near tracked features" mask, built on the device from the points themselves:
 * 255 everywhere, 0 in the inclusive box [cvRound(x-half), cvRound(x+half)] x [cvRound(y-half), cvRound(y+half)]
 * around each point (src/tracking.cpp:76-79, half = 10), then FAST as above.  Saves the w*h-byte mask upload. */
SVO_GPU_API int svo_gpu_fast_masked_by_points(svo_gpu_ctx* ctx, int slot, int threshold, int nms, const float* mask_xy,
                                              int n_mask_pts, float half, float* xy_out, float* response_out,
                                              int capacity, int* count_out);

/* ---------------------------------------------------------------------------------------------- bucketing ----- */

/* FeatureSet::bucketingFeatures(image, bucket_size, features_per_bucket) (src/bucket.cpp:24-68, Bucket::add_feature
 * :72-101, get_features :103-106), in place on (points, ages), bit-exact including its two quirks: a full bucket always
 * overwrites its slot 0, and buckets aliased by the row stride are emitted twice.  img_w/img_h are image.cols/rows
 * (src/bucket.cpp:29-30).  A point that maps outside the bucket table (undefined behaviour in the reference) returns
 * SVO_GPU_E_ARG.  capacity = number of entries xy_inout/ages_inout can hold (>= n_in). */
SVO_GPU_API int svo_gpu_bucket(svo_gpu_ctx* ctx, float* xy_inout, int* ages_inout, int n_in, int img_w, int img_h,
                               int bucket_size, int per_bucket, int capacity, int* n_out);

/* Bucketing without any allocation.  No counterpart in the reference (which has no device code); same rule and same result
 * as svo_gpu_bucket, for callers that cannot obtain device memory themselves or through a context:
 *   - a per-frame tracking thread with a latency budget: cudaMalloc / cudaFree take driver-wide locks and can wait for every
 *     stream of the device, so such loops are written to allocate nothing once they run -- svo_gpu_create allocates all of a
 *     context's buffers, and svo_gpu_bucket re-allocates its bucket tables when a call needs a finer grid than any before;
 *   - a plug-in inside an application that has handed all of the GPU's memory to its own pool (an inference server, a
 *     simulator): a cudaMalloc from a guest library fails there, and the plug-in only ever sees buffers it is given.
 * Such a caller passes device buffers it already owns and needs no svo_gpu_ctx.  The per-bucket tables (arrival count,
 * largest index and the per_bucket smallest indices of each bucket -- the scratch svo_gpu_bucket keeps in context buffers)
 * live in ONE static device array of SVO_GPU_BUCKET_STATIC_BYTES that is part of the library image: every GPU on which this
 * library's kernels get loaded holds it until the process exits, whether or not this function is ever called.
 *   limit      ((img_h / bucket_size + 1) * (img_w / bucket_size + 1)) * (2 + per_bucket) * 4 + 256 bytes must fit the array
 *              (the reference's 1241x376 frame with 32-px buckets of 5 needs 13 KB; 3840x2160 with 16-px buckets of 5 needs
 *              0.9 MB); anything larger returns SVO_GPU_E_CAPACITY and says so -- nothing is truncated.  n_in is not limited.
 *   inputs     xy_in_dev [n_in][2] floats and ages_in_dev_or_null [n_in] (NULL = all ages 0, FeatureSet::appendNewFeatures)
 *              in DEVICE memory; not modified.
 *   outputs    xy_out_dev [capacity][2], ages_out_dev [capacity] in DEVICE memory, not overlapping the inputs; *n_out (host)
 *              = bucketed features.  More than `capacity` of them: SVO_GPU_E_CAPACITY, *n_out = the count required, the
 *              first `capacity` entries written.  xy arrays must be 8-byte aligned.
 *   stream     a cudaStream_t (as void*) of the CURRENT device, or NULL for the default stream.  The work is enqueued there,
 *              after whatever the caller enqueued to produce the inputs; the call returns when it is done (outputs valid).
 *   errors     codes as everywhere in this header; the message goes to errbuf_or_null (errbuf_len bytes, NUL-terminated).
 *   first call The first call on each device looks up where the static array lives on that device (the device that is
 *              current at that moment), which may make the CUDA runtime load this library's kernels there.  Make that first
 *              call during start-up, before any stream capture begins on the device; later calls only launch work.
 *   sharing    All calls in the process use the same array on a given device, and the array keeps the previous call's
 *              tables: each call clears the part it is about to use before reading it.  Two calls whose GPU work can overlap
 *              (two host threads, or two streams, on one device) would corrupt each other's tables: serialise them.
 * The function calls no cudaMalloc / cudaFree / cudaHostAlloc and no library routine that does; its only transfers are a
 * 4-byte memset and one 8-byte read-back (count and error flag) into a local variable. */
#define SVO_GPU_BUCKET_STATIC_BYTES (1 << 20)
SVO_GPU_API int svo_gpu_bucket_static(const float* xy_in_dev, const int* ages_in_dev_or_null, int n_in, int img_w, int img_h,
                                      int bucket_size, int per_bucket, float* xy_out_dev, int* ages_out_dev, int capacity,
                                      int* n_out, void* stream, char* errbuf_or_null, size_t errbuf_len);

/* ---------------------------------------------------------------------------------------------- KLT ----------- */

/* cv::calcOpticalFlowPyrLK(prevImg, nextImg, prevPts, nextPts, status, err, winSize, maxLevel, criteria, flags,
 * minEigThreshold) on lane 0 of two slots whose pyramids were built with a window >= (win_w, win_h):
 *   temporal tracking src/tracking.cpp:160-165 (21x21, maxLevel 3, {COUNT+EPS,50,0.001}, OPTFLOW_LK_GET_MIN_EIGENVALS)
 *   stereo matching   src/tracking.cpp:101-105 (11x11, maxLevel 3, {COUNT+EPS,30,0.001}, flags 0).
 * criteria_type/max_count/epsilon are cv::TermCriteria's fields; status is 0/1; err is the min-eigenvalue or the L1
 * patch residual depending on flags.  OPTFLOW_USE_INITIAL_FLOW returns SVO_GPU_E_UNSUPPORTED (the reference never sets it). */
SVO_GPU_API int svo_gpu_klt(svo_gpu_ctx* ctx, int slot_prev, int slot_next, const float* prev_xy, float* next_xy,
                            uint8_t* status, float* err, int n, int win_w, int win_h, int max_level, int max_count,
                            double epsilon, int criteria_type, int flags, double min_eig_threshold);
/* total LK iterations executed by the last svo_gpu_klt / step call (all lanes), for the op model of the roofline */
SVO_GPU_API long long svo_gpu_last_klt_iterations(svo_gpu_ctx* ctx);

/* ---------------------------------------------------------------------------------------------- filters ------- */

/* The loops that follow each LK call in the reference, as an order-preserving compaction on the device:
 *   y_threshold <  0: keep i iff status[i]                                              (src/tracking.cpp:168-175)
 *   y_threshold >= 0: keep i iff status[i] && std::abs(b[i].y - a[i].y) < y_threshold    (src/tracking.cpp:109-115)
 * Survivors keep their input order; index_out[k] is the input position of survivor k (so that the caller can carry
 * Feature::mapPoint over, src/tracking.cpp:112,171).  index_out / a_out / b_out may be NULL; capacity is theirs. */
SVO_GPU_API int svo_gpu_filter_tracks(svo_gpu_ctx* ctx, const float* a_xy, const float* b_xy, const uint8_t* status, int n,
                                      float y_threshold, int* index_out, float* a_out, float* b_out, int capacity,
                                      int* n_out);

/* Tracking::trackFrames (src/tracking.cpp:154-179) in one call on lane 0: LK from slot_prev to slot_next on prev_xy,
 * then the status filter, chained on the device; only the survivors are copied back.
 *   index_out[k] = position in prev_xy of survivor k, curr_out = its tracked position, err_out = its LK err (or NULL). */
SVO_GPU_API int svo_gpu_track_frames(svo_gpu_ctx* ctx, int slot_prev, int slot_next, const float* prev_xy, int n, int win_w,
                                     int win_h, int max_level, int max_count, double epsilon, int criteria_type, int flags,
                                     double min_eig_threshold, int* index_out, float* curr_out, float* err_out,
                                     int capacity, int* n_out);

/* Parameters of Tracking::extractFeatures + findLeftFeaturesInRight (src/tracking.cpp:74-122). */
typedef struct svo_gpu_keyframe_params {
    int fast_threshold;     /* include/config_reader.h:36 (20) */
    int fast_nms;           /* include/config_reader.h:37 (true) */
    float mask_half;        /* src/tracking.cpp:78 (10); <= 0 disables the mask */
    int bucket_size;        /* <= 0: no bucketing (the reference as shipped, "todo" at src/tracking.cpp:88) */
    int per_bucket;
    int win_w, win_h;       /* src/tracking.cpp:104 (11, 11) */
    int max_level;          /* src/tracking.cpp:104 (3) */
    int max_count;          /* src/tracking.cpp:98 (30) */
    double epsilon;         /* src/tracking.cpp:98 (0.001) */
    int criteria_type;      /* COUNT + EPS */
    int flags;              /* 0 */
    double min_eig_threshold;
    float y_threshold;      /* include/config_reader.h:41 (40); < 0 disables the epipolar test */
} svo_gpu_keyframe_params;

/* A key-frame's feature extraction in one call on lane 0, everything chained on the device:
 * mask around prev_xy (:76-79) -> FAST on slot_left (:82) -> optional bucketing (:88) -> LK from slot_left to
 * slot_right on the detections (:101-105) -> status / epipolar filter (:109-115).  Both slots need pyramids built with a
 * window >= (win_w, win_h).  Outputs: left_out / right_out = surviving stereo pairs in detection order,
 * *n_detected = detections LK ran on, *n_out = pairs kept. */
SVO_GPU_API int svo_gpu_keyframe(svo_gpu_ctx* ctx, int slot_left, int slot_right, const float* prev_xy, int n_prev,
                                 const svo_gpu_keyframe_params* params, float* left_out, float* right_out, int capacity,
                                 int* n_detected, int* n_out);

/* ---------------------------------------------------------------------------------------------- per-frame step - */

/* Parameters of one front-end frame, i.e. what Tracking::startStereo does per frame (src/tracking.cpp:240-257):
 * trackFrames(prev, curr) (:154-179) then, on a key-frame, extractFeatures(curr) (:74-92) with bucketing at the "todo" of :88.
 * Two track policies:
 *   carry_tracks == 0  every frame is a key-frame and the lane's tracks are RE-SEEDED from the first max_tracks bucketed
 *                      detections (a fixed, maximal per-frame load: what bench.py times);
 *   carry_tracks != 0  the reference's loop: the LK survivors (status != 0, input order, src/tracking.cpp:168-175) are the
 *                      lane's tracks for the next frame; detection runs only on a key-frame -- the lane's first frame, or
 *                      when the frame before was not a key-frame and ended with fewer than features_to_track tracks
 *                      (src/tracking.cpp:69-70) -- with the mask around the points LK started from (:77), and the bucketed
 *                      detections are appended after the survivors (StereoFrame::insertFeatures, :254) up to max_tracks.
 *                      The decision is taken per lane on the device; lanes that are not key-frames skip FAST and bucketing. */
typedef struct svo_gpu_step_params {
    int w, h;               /* image size of every lane */
    int fast_threshold;     /* include/config_reader.h:36 (20) */
    int fast_nms;           /* include/config_reader.h:37 (true) */
    float mask_half;        /* src/tracking.cpp:78 (10); <= 0 disables the tracked-points mask */
    int bucket_size;        /* FeatureSet::bucketingFeatures argument */
    int per_bucket;         /* FeatureSet::bucketingFeatures argument */
    int max_tracks;         /* tracks carried to the next frame = first max_tracks bucketed features */
    int win_w, win_h;       /* src/tracking.cpp:163 (21, 21) */
    int max_level;          /* src/tracking.cpp:163 (3) */
    int max_count;          /* src/tracking.cpp:157 (50) */
    double epsilon;         /* src/tracking.cpp:157 (0.001) */
    int criteria_type;      /* COUNT + EPS */
    int flags;              /* src/tracking.cpp:164 (OPTFLOW_LK_GET_MIN_EIGENVALS) */
    double min_eig_threshold; /* OpenCV default 1e-4 */
    int carry_tracks;       /* see above */
    int features_to_track;  /* configs/config.yaml:15 (70); used with carry_tracks */
} svo_gpu_step_params;

/* Host-side views of one step's results; any pointer may be NULL (that output is then left on the device).
 * Arrays are [batch][capacity] with the per-lane strides given in elements. */
typedef struct svo_gpu_step_outputs {
    int* n_tracked;         /* [batch] number of points LK was run on in this step (0 on the first step) */
    float* next_xy;         /* [batch][track_capacity][2] */
    uint8_t* status;        /* [batch][track_capacity] */
    float* err;             /* [batch][track_capacity] */
    int track_capacity;
    int* n_keypoints;       /* [batch] FAST keypoints found on the new frame */
    float* kp_xy;           /* [batch][kp_capacity][2] */
    float* kp_response;     /* [batch][kp_capacity] */
    int kp_capacity;
    int* n_bucketed;        /* [batch] features kept by bucketing; the first max_tracks of them are the new tracks */
    float* bucket_xy;       /* [batch][bucket_capacity][2] */
    int bucket_capacity;
    /* carry_tracks mode (zero / NULL otherwise): */
    int* keyframe;          /* [batch] 1 if detection ran for the lane in this step (StereoFrame::isKeyFrame) */
    int* n_survived;        /* [batch] LK survivors = the head of the lane's next track list, in input order */
    int* survivor_index;    /* [batch][track_capacity] position in this step's track list of survivor k (carries
                               Feature::mapPoint over, src/tracking.cpp:171); the next track list is the survivors'
                               next_xy followed by the first min(n_bucketed, max_tracks - n_survived) bucket_xy entries */
} svo_gpu_step_outputs;

/* One frame of `batch` independent streams: ingest images[b] (host or device memory), build its pyramid, run LK from
 * each lane's previous frame on that lane's current tracks, run FAST (+ tracked-points mask) on the new frame, bucket
 * the detections, and make the first max_tracks of them the lane's tracks for the next call.  The context cycles through
 * a ring of three slots internally (SVO_GPU_NUM_SLOTS-3 .. -1: this frame, the previous one, and a free
 * one for the next frame's copy and pyramid build).  Lanes never interact; this is how independent sequences share one
 * GPU (and, with one context per device, several GPUs). */
SVO_GPU_API int svo_gpu_step(svo_gpu_ctx* ctx, const uint8_t* images, int mem_kind, size_t stride_bytes,
                             size_t image_stride_bytes, int batch, const svo_gpu_step_params* params,
                             const svo_gpu_step_outputs* outputs_or_null);
/* The same step with the frame ingest pipelined (the reference's AsyncImageLoader,
 * include/async_image_loader.h:27-47, has the next frames decoded before they are needed): `next_images` (host memory,
 * same geometry, or NULL) are the frames of the FOLLOWING call; their host-to-device copy is started now on a second
 * stream and overlaps this step's kernels.  The following call must pass that same pointer as `images` (otherwise the
 * prefetch is discarded and the frames are copied normally).  The caller keeps `next_images` unchanged until then. */
SVO_GPU_API int svo_gpu_step_pipelined(svo_gpu_ctx* ctx, const uint8_t* images, const uint8_t* next_images, int mem_kind,
                                       size_t stride_bytes, size_t image_stride_bytes, int batch,
                                       const svo_gpu_step_params* params, const svo_gpu_step_outputs* outputs_or_null);
/* The same step split in two, for callers that keep the GPU busy across frames (the reference's loop body, src/tracking.cpp:
 * 240-257, consumes frame t's results on the host while the loader thread already holds frame t+1): _submit enqueues the whole
 * step and returns at once with a ticket; _wait(ticket) blocks until that step is through, fills the n_* counts and reports
 * capacity errors.  Up to SVO_GPU_STEP_DEPTH steps may be outstanding, waited for in submission order.  Frame t+1 never
 * needs frame t's results on the HOST (tracks are re-seeded on the device), so submit(t+1) before wait(t) is legal and keeps
 * the device from idling while the host handles frame t.  The caller keeps `images` (device memory), `next_images` and the
 * `outputs` arrays (pinned host memory for true overlap; a distinct set per outstanding step) alive until the wait.
 * svo_gpu_step_pipelined(...) == submit + wait. */
SVO_GPU_API int svo_gpu_step_submit(svo_gpu_ctx* ctx, const uint8_t* images, const uint8_t* next_images, int mem_kind,
                                    size_t stride_bytes, size_t image_stride_bytes, int batch,
                                    const svo_gpu_step_params* params, const svo_gpu_step_outputs* outputs_or_null,
                                    int* ticket_out);
SVO_GPU_API int svo_gpu_step_wait(svo_gpu_ctx* ctx, int ticket);
/* Forgets the previous frame and tracks of every lane (start of a new set of sequences). */
SVO_GPU_API int svo_gpu_step_reset(svo_gpu_ctx* ctx);
/* Number of kernel launches (graph kernel nodes included) issued by this context so far. */
SVO_GPU_API long long svo_gpu_launch_count(const svo_gpu_ctx* ctx);
/* Per-stage device time of the last svo_gpu_step in milliseconds (CUDA events on the context stream):
 * ms[0] ingest+pyramid, ms[1] LK, ms[2] FAST, ms[3] bucketing.  Only recorded when enabled (adds event overhead). */
SVO_GPU_API int svo_gpu_step_profile(svo_gpu_ctx* ctx, int enable, float ms_out[4]);

#ifdef __cplusplus
}
#endif
#endif /* SVO_GPU_H */
