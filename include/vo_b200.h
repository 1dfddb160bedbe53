/*
 * vo_b200.h - C-ABI of libvo_b200.so: the B200-native (sm_100a) per-frame core of the
 * "locally consistent visual odometry pipeline".
 *
 * The reference (pure Python over OpenCV) has no FFI; its boundary for this path is the
 * class surface of src/vo_pipeline.py consumed by src/main.py:7 (`from vo_pipeline import *`).
 * Each entry point below names the reference call it replaces (file:line under
 * /root/reference/src).  The Python drop-in module (package file vo_pipeline.py) binds these
 * with ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *   - plain C types only; every function returns VO_OK (0) or a negative vo_status and never
 *     throws across the ABI; vo_last_error() gives the message of the last failure;
 *   - the caller owns every host buffer; the context owns all device memory (grows by doubling);
 *   - one vo_ctx per stream / GPU; a ctx is not thread-safe, different ctxs may be used from
 *     different threads; work is enqueued on the ctx's CUDA stream and only the functions
 *     documented as "synchronises" wait for it;
 *   - pointer arguments documented "host or device" are classified with the CUDA unified
 *     address space, so pinned host, pageable host and device buffers are all accepted;
 *   - there is NO CPU fallback: without a usable CUDA device vo_create fails with VO_ERR_CUDA.
 */
#ifndef VO_B200_H
#define VO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden */
#endif

#define VO_B200_ABI_VERSION 1

typedef enum vo_status {
    VO_OK = 0,
    VO_ERR_INVALID = -1,   /* bad argument */
    VO_ERR_CUDA = -2,      /* CUDA runtime error (message has the CUDA string) */
    VO_ERR_NOMEM = -3,     /* allocation failed */
    VO_ERR_NOFRAME = -4,   /* frame key not resident in the pyramid cache */
    VO_ERR_CAPACITY = -5,  /* a fixed table (e.g. RNG draws) was exhausted */
    VO_ERR_STATE = -6      /* call sequence invalid (e.g. no landmarks set) */
} vo_status;

typedef struct vo_ctx vo_ctx;

/* cv2.calcOpticalFlowPyrLK parameters; vo_klt_params_default() = what the reference uses
 * (vo_pipeline.py:215, :501 pass no parameters => OpenCV defaults). */
typedef struct vo_klt_params {
    int win;              /* window side, odd, <= 21 (reference: 21) */
    int max_level;        /* pyramid levels - 1 (reference: 3) */
    int max_iters;        /* Gauss-Newton iterations per level (reference: 30) */
    double epsilon;       /* |delta| stop threshold, squared internally (reference: 0.01) */
    double min_eig_thr;   /* reference: 1e-4 */
    int fb_check;         /* 0 = off (reference). 1 = extra forward-backward check */
    double fb_thresh;     /* max forward-backward distance in px when fb_check=1 */
} vo_klt_params;

/* cv2.solvePnPRansac parameters (vo_pipeline.py:258-264 and :238-243). */
typedef struct vo_pnp_params {
    int max_iters;          /* iterationsCount (reference: OpenCV default 100) */
    double confidence;      /* reference: 0.99999 / config.yaml value */
    double reproj_thresh;   /* reference: REPOJ_THRESH = 2 px */
    int refine_iters;       /* LM iterations of the final refit on the inlier set (>=1) */
    int min_inliers;        /* reference re-bootstraps below 4 (vo_pipeline.py:274) */
} vo_pnp_params;

void vo_klt_params_default(vo_klt_params* p);
void vo_pnp_params_default(vo_pnp_params* p);

int vo_abi_version(void);
/* Number of CUDA devices visible; <0 on CUDA error (no driver / no GPU). */
int vo_device_count(void);

/* ---- context ---------------------------------------------------------------------------- */
/* stream_handle: an existing cudaStream_t to enqueue on (e.g. torch's current stream), or NULL
 * to let the ctx create its own non-blocking stream. */
int vo_create(vo_ctx** out, int device, int max_w, int max_h, int track_capacity, int hyp_capacity,
              void* stream_handle);
int vo_destroy(vo_ctx* ctx);
const char* vo_last_error(vo_ctx* ctx); /* ctx may be NULL: last creation error */
int vo_synchronize(vo_ctx* ctx);
/* number of kernels this library has launched on the ctx since creation (bench "gpu_launches") */
int64_t vo_launch_count(vo_ctx* ctx);

/* pinned host allocations for frame staging (bench / drop-in use them for zero-staging H2D) */
int vo_alloc_pinned(void** out, uint64_t bytes);
int vo_free_pinned(void* p);

/* ---- frames: replaces the implicit buildOpticalFlowPyramid of both cv2.calcOpticalFlowPyrLK
 * calls (vo_pipeline.py:215, :501); one upload + one pyramid per frame, cached by key -------- */
/* img: host or device, 8-bit gray, `pitch` bytes per row.  No-op if frame_key is resident. */
int vo_upload_frame(vo_ctx* ctx, const uint8_t* img, int w, int h, int pitch, int64_t frame_key);
int vo_has_frame(vo_ctx* ctx, int64_t frame_key);
int vo_drop_frames(vo_ctx* ctx);
/* test hook: copy pyramid level `level` of a resident frame (unpadded, w*h bytes) to the host;
 * deriv (optional) receives the int16 (Ix,Iy) plane, 2*w*h values.  Synchronises. */
int vo_get_pyramid_level(vo_ctx* ctx, int64_t frame_key, int level, uint8_t* img, int16_t* deriv,
                         int* w, int* h);

/* ---- unit-level entry point mirroring cv2.calcOpticalFlowPyrLK(prev, next, pts, None)
 * (vo_pipeline.py:215, :501).  pts/next_pts: n x 2 float32, status: n uint8, err: n float32
 * (err, iters may be NULL; iters = Gauss-Newton iterations executed per track).  Host buffers.
 * Synchronises. */
int vo_calc_optical_flow_pyr_lk(vo_ctx* ctx, int64_t prev_key, int64_t next_key, const float* pts,
                                int n, float* next_pts, uint8_t* status, float* err, int32_t* iters,
                                const vo_klt_params* params);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* VO_B200_H */
