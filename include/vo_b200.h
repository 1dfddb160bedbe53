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
/* SURVEY.md 8f-2, frame ingest (main.py:216-223 reads frame k+1 only after frame k is done; here the
 * NEXT frame can be handed over early): same as vo_upload_frame but enqueued on the ctx's ingest
 * stream, so the H2D copy and the pyramid kernels run under the tracking of the current frame.  The
 * first stage that uses frame_key waits for it on the device; nothing synchronises the host.  Call it
 * BEFORE enqueueing the current frame's stages: the ingest is ordered behind the work enqueued so far
 * (its slot may still be read by it) and overlaps whatever is enqueued afterwards.  A pinned or
 * device `img` must stay untouched until a stage using frame_key has been synchronised. */
int vo_prefetch_frame(vo_ctx* ctx, const uint8_t* img, int w, int h, int pitch, int64_t frame_key);
/* The same, announced now and put on the ingest stream by the next pose call (vo_pnp_ransac[_promote])
 * between its last launch and its wait - the ~20 us the enqueue costs on the host then hide under the
 * GPU's PnP - or by the first call that asks for frame_key.  img must stay valid until then. */
int vo_prefetch_frame_deferred(vo_ctx* ctx, const uint8_t* img, int w, int h, int pitch, int64_t frame_key);
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

/* ---- device-resident Markov state: BestVision.state {'P','X'} and .extended_state {'C','F','T'}
 * (vo_pipeline.py:36-37).  P: n x 2 float32, X: n x 3 float64, C: n x 2 float32 (candidate positions
 * are float32-exact in the reference: KLT outputs / detector outputs), F: n x 2 float64,
 * T: n x 16 float64 (4x4 row-major world->camera pose of the first observation). -------------- */
typedef struct vo_counters {
    int32_t n_landmarks, n_candidates;   /* current |P|, |C| */
    int32_t n_tracked, n_lost;           /* last landmark KLT: kept / lost */
    int32_t n_inliers, pnp_ok;           /* last PnP */
    int32_t ransac_iters, best_hypothesis, lm_iters;
    int32_t n_promoted, n_candidates_lost, cf_alias;
    int32_t klt_iters;                   /* sum of LK iterations over the last landmark KLT */
    int32_t klt_iters_candidates;        /* same over the last candidate KLT (set by the promotion step) */
    double reproj_cost;                  /* sum of squared reprojection errors of the refit */
    int32_t n_appended;                  /* rows the last vo_append_landmarks / queued batch added */
    int32_t reserved_;                   /* vo_process_frame_pipelined: bit 0 = the next frame's tracking step is enqueued ahead */
} vo_counters;

/* Upload landmarks after bootstrap (main.py:198-200 vision.state = state).  Synchronises. */
int vo_set_state(vo_ctx* ctx, const float* P, const double* X, int n);
/* Host mirror of state['P'], state['X'] (main.py:248 reads state['X'] every frame).  P/X may be
 * NULL to query n only.  Synchronises. */
int vo_get_state(vo_ctx* ctx, float* P, double* X, int cap, int* n);
/* The landmarks as they were before the last vo_klt_track_landmarks / vo_process_frame; a third
 * landmark buffer keeps them on the device until the next tracking step (main.py:248 reads the
 * previous frame's state['X'] after the current frame's associate + estimatePose).  Synchronises. */
int vo_get_prev_state(vo_ctx* ctx, float* P, double* X, int cap, int* n);
/* ... and as they were before the tracking step before that one (a fourth buffer of the rotation): what main.py:248
 * reads when the next frame's tracking step has already been enqueued inside the pose call
 * (vo_process_frame_pipelined).  Synchronises. */
int vo_get_prev2_state(vo_ctx* ctx, float* P, double* X, int cap, int* n);
/* Asynchronous pinned-host snapshot of P/X as they are now, readable after later calls have
 * replaced them on the device (main.py:248 reads the previous frame's state['X'] after the next
 * frame's associate/estimatePose).  vo_get_snapshot synchronises. */
int vo_snapshot_state(vo_ctx* ctx);
int vo_get_snapshot(vo_ctx* ctx, float* P, double* X, int cap, int* n);
/* Load / read the candidate set (tests, checkpoint/resume).  cf_alias: reference quirk App. B-5. */
int vo_set_candidates(vo_ctx* ctx, const float* C, const double* F, const double* T, int n, int cf_alias);
int vo_get_candidates(vo_ctx* ctx, float* C, double* F, double* T, int cap, int* n);
/* points_tracking_lost of associateKeypointsToLandmarks (vo_pipeline.py:220-221). */
int vo_get_lost(vo_ctx* ctx, float* lost, int cap, int* n);
int vo_get_counters(vo_ctx* ctx, vo_counters* out);   /* synchronises */
int vo_peek_counters(vo_ctx* ctx, vo_counters* out);  /* as of the last synchronising call; no sync */
/* Diagnostics: %globaltimer stamps (ns) taken by the CTA that finished the last RANSAC run - {kernel entry,
 * its P3P done, its hypotheses scored, chunk bookkeeping done, inlier mask done, LM refit done, pruning
 * done, 0}.  Synchronises. */
int vo_debug_pnp_stamps(vo_ctx* ctx, uint64_t out[8]);

/* a1  KeypointsToLandmarksAssociator.associateKeypointsToLandmarks (vo_pipeline.py:188-227):
 * KLT of P from prev to next frame, keep status==1 rows of P and X.  The candidates C are tracked
 * in the same launch (the reference does it later, at vo_pipeline.py:501, on the same frame pair).
 * n_kept may be NULL: then nothing synchronises. */
int vo_klt_track_landmarks(vo_ctx* ctx, int64_t prev_key, int64_t next_key, const vo_klt_params* params,
                           int* n_kept);
/* a1 with the PREVIOUS frame's append of new candidates (vo_pipeline.py:574-591) folded in: same result
 * as vo_add_candidates(new_candidates, n_new, T44) followed by vo_klt_track_landmarks, but the rows
 * travel on the side stream in front of the candidates' KLT, so neither their copy nor the append
 * kernel sits between the caller and the landmark launch (the head of the chain the pose waits for). */
int vo_klt_track_landmarks_add(vo_ctx* ctx, int64_t prev_key, int64_t next_key, const vo_klt_params* params,
                               const double* new_candidates, int n_new, const double T44[16], int* n_kept);
/* The same, plus the pose-independent half of the coming promotion step on the side stream behind the
 * candidates' KLT: lost rows out and positions updated (vo_pipeline.py:503-545), bearing test with
 * threshold early_alpha_deg (:621-642, raw pixels only), compaction of C/F/T.  The promotion call that
 * follows (vo_pnp_ransac_promote / vo_track_triangulate_promote) with the same alpha then only
 * triangulates and appends the promoted rows; with another alpha, without a pose, or if something looks
 * at the candidates first, the early half is taken back and redone the long way.  Results are identical.
 * early_alpha_deg < 0: none (= vo_klt_track_landmarks_add). */
int vo_klt_track_landmarks_ex(vo_ctx* ctx, int64_t prev_key, int64_t next_key, const vo_klt_params* params,
                              const double* new_candidates, int n_new, const double T44[16], double early_alpha_deg,
                              int* n_kept);
/* Undo the last vo_klt_track_landmarks[_add], provided nothing but that tracking step has run since:
 * the drop-in enqueues the tracking of a frame it was handed early (vo_prefetch_frame) right behind the
 * previous frame's promotion, BEFORE associateKeypointsToLandmarks is called for it, and takes it back
 * if that call then comes with other arguments.  P/X are again what the step started from; a folded
 * append stays; the state vo_get_prev_state served before is gone. */
int vo_track_rollback(vo_ctx* ctx);
/* a3  PoseEstimator.estimatePose (vo_pipeline.py:245-335): PnP-RANSAC + LM refit, prunes P/X to
 * the inliers (vo_pipeline.py:327-331), T44 = 4x4 row-major world->camera (vo_pipeline.py:333).
 * ok=0 (no model / n<4) leaves the state untouched: the caller takes the re-bootstrap branch of
 * vo_pipeline.py:274-324.  Synchronises. */
int vo_pnp_ransac(vo_ctx* ctx, const double K[9], const vo_pnp_params* params, double T44[16], int* n_inliers,
                  int* ok);
/* a3 + a6 with ONE synchronisation: vo_pnp_ransac, then - speculatively, before the pose is back on
 * the host - the promotion step of vo_track_triangulate_promote with the pose taken from the device
 * (the class API forces a sync inside estimatePose; with the promotion enqueued ahead of it the GPU
 * does not idle while the host walks from estimatePose to triangulate_landmark).  Needs the
 * candidates tracked with the landmarks (vo_klt_track_landmarks on this frame pair).  Afterwards the
 * caller confirms (vo_promote_commit: triangulate_landmark got the pose/alpha assumed here) or
 * undoes it (vo_promote_rollback: state as after vo_pnp_ransac, candidate tracking still pending).
 * ok=0 is rolled back internally.  Until then the promoted rows sit behind the n_inliers rows of P/X.
 * n_promoted == NULL: the host waits only for the pose (the scalar block is read back between the PnP
 * and the promotion kernels), so the promotion step keeps the GPU busy while the host walks on to
 * triangulate_landmark; vo_peek_counters then shows the landmarks as they were BEFORE the promoted rows
 * are appended (n_landmarks = n_inliers), and the candidates before the promotion step too - unless the
 * tracking call already classified them (vo_klt_track_landmarks_ex), then n_candidates / n_promoted /
 * n_candidates_lost are the final ones. */
int vo_pnp_ransac_promote(vo_ctx* ctx, const double K[9], const vo_pnp_params* params, double alpha_deg,
                          double T44[16], int* n_inliers, int* ok, int* n_promoted);
int vo_promote_commit(vo_ctx* ctx);   /* (a vo_klt_track_landmarks* call behind the speculation confirms it too) */
int vo_promote_rollback(vo_ctx* ctx);
int vo_promote_pending(vo_ctx* ctx);
/* The caller adopts the tracking step vo_process_frame_pipelined enqueued ahead as the next frame's tracking
 * step (the class surface does, when associateKeypointsToLandmarks is called with the arguments assumed):
 * state queries then no longer take it back. */
int vo_ahead_adopt(vo_ctx* ctx);
/* 1 if the candidates were tracked over (prev, next) together with the landmarks and not consumed yet */
int vo_candidates_tracked(vo_ctx* ctx, int64_t prev_key, int64_t next_key);
/* a6  LandmarkTriangulator.triangulate_landmark, steps 1 and 3 (vo_pipeline.py:489-545, 608-721):
 * track C, drop lost rows of C/F/T, alpha(C,F) > alpha_deg => DLT with (T_first, T44) and append to
 * P/X, remove from C/F/T.  prev_key = next_key = -1 runs the promotion step without tracking.
 * n_promoted may be NULL (no sync). */
int vo_track_triangulate_promote(vo_ctx* ctx, int64_t prev_key, int64_t next_key, const double K[9],
                                 const double T44[16], double alpha_deg, const vo_klt_params* params,
                                 int* n_promoted);
/* Track C over (prev, next) now (vo_pipeline.py:501) - a no-op if vo_klt_track_landmarks already
 * did for this frame pair - and read the pending result: previous / tracked positions and status
 * (vo_pipeline.py:503-508; the 'shi-mask' detector masks the well-tracked previous positions,
 * vo_pipeline.py:376-379, :568).  vo_get_candidate_tracks synchronises. */
int vo_track_candidates(vo_ctx* ctx, int64_t prev_key, int64_t next_key, const vo_klt_params* params);
int vo_get_candidate_tracks(vo_ctx* ctx, float* prev_pts, float* next_pts, uint8_t* status, int cap, int* n);
/* a6 step 2 (vo_pipeline.py:564-591): append candidates C := F := pts (n x 2 float64), T := T44.
 * Call after vo_track_triangulate_promote (new rows have alpha = 0 and are never promoted in the
 * frame they arrive, so the order does not change results). */
int vo_add_candidates(vo_ctx* ctx, const double* pts, int n, const double T44[16]);

/* Landmarks from outside the tracking loop, appended behind P/X: what the reference does with the fresh
 * landmarks of its re-bootstrap (vo_pipeline.py:318-320, np.vstack onto associations['P'] / ['X']) and with
 * the bootstrap result (main.py:198-200).  P: n x 2 float32 pixel positions in the frame tracked INTO last,
 * X: n x 3 float64; host or device.  target_total > 0: top-up - only the first
 * min(n, max(0, target_total - |P|)) rows are appended; the count is decided on the device (|P| is not known
 * to the host without a round trip) and reported in vo_counters.n_appended.  Enqueued on the ctx stream, no
 * synchronisation; pageable rows are copied before the call returns, device / pinned rows must stay valid
 * until the next synchronising call.  Valid between frames (after the promotion step of one frame, before the
 * tracking step of the next); a tracking step enqueued ahead (vo_process_frame_pipelined) is taken back. */
int vo_append_landmarks(vo_ctx* ctx, const float* P, const double* X, int n, int target_total);
/* The same, deferred: the batch is appended by the next vo_process_frame[_pipelined] right behind that
 * frame's promotion step and candidate append - in front of a tracking step it enqueues ahead, so the frames
 * stay overlapped.  Up to 4 batches per frame, appended in call order.  Rows must stay valid until the call
 * AFTER that one returns (pageable rows: until that call returns). */
int vo_queue_landmarks(vo_ctx* ctx, const float* P, const double* X, int n, int target_total);

/* ---- the bootstrap on the device.  Replaces, in VOInitializer.getPoseEstimate
 * (vo_pipeline.py:139-162) and in the "< 4 inliers" re-bootstrap of estimatePose (vo_pipeline.py:274-324),
 *     E, mask = cv2.findEssentialMat(kps1, kps2, K, cv2.RANSAC, 0.999, 1.0, None)
 *     _, R, t, _ = cv2.recoverPose(E, kps1, kps2, K, mask=mask)
 * five-point essential-matrix RANSAC (cv::RNG(2^64-1) subsets, Sampson error in float32, first strictly better
 * model wins, adaptive budget, no refit), then the cheirality vote over the four (R, t) decompositions with
 * OpenCV's distance threshold 50, whose mask is ANDed into the RANSAC mask - as the in-place `mask=mask` of the
 * reference's call does. */
typedef struct vo_boot_params {
    double prob;        /* RANSAC confidence (reference: 0.999) */
    double threshold;   /* Sampson threshold in pixels (reference: 1.0) */
    int max_iters;      /* OpenCV's default maxIters: 1000 */
} vo_boot_params;
void vo_boot_params_default(vo_boot_params* p);
/* kps1, kps2: n x 2 float64 pixel positions (host or device).  Outputs (all optional): T34 = [R | t] 3x4 row-major
 * (|t| = 1), mask n bytes (host or device; 1 = RANSAC inlier that passes the winner's cheirality test), E 3x3 (unit
 * Frobenius norm), info = {ok, iterations run, winning iteration, winning model of it, RANSAC inliers, inliers after the
 * cheirality test}.  ok = 0 (fewer than 5 matches, no model): everything zero.  Synchronises. */
int vo_bootstrap_pose(vo_ctx* ctx, const double* kps1, const double* kps2, int n, const double K[9],
                      const vo_boot_params* params, double T34[12], uint8_t* mask, double E[9], int32_t info[6]);

/* ---- SURVEY.md 8f-1 (first "next" row): the Shi-Tomasi candidate detector on the device.
 * Replaces cv2.goodFeaturesToTrack(new_frame, maxCorners=500, qualityLevel=0.03, minDistance=10,
 * mask=mask, blockSize=4, gradientSize=3, useHarrisDetector=False) at vo_pipeline.py:381
 * (LandmarkTriangulator.find_new_candidates_shi, vo_pipeline.py:357-390). */
typedef struct vo_gftt_params {
    int max_corners;       /* 500 (<= 0: no limit) */
    double quality_level;  /* 0.03 */
    double min_distance;   /* 10 */
    int block_size;        /* 4 (1..8); gradientSize is 3 */
    int vec_cols;          /* columns the OpenCV build evaluates with its fused SIMD row filter;
                            * -1 = width & ~31, what the pinned build does (oracle/gftt.c) */
    int use_cluster;       /* 1: corner selection on a thread-block cluster of 8 CTAs (default);
                            * 0: single-CTA kernel only (same result; tests) */
} vo_gftt_params;
void vo_gftt_params_default(vo_gftt_params* p);
/* Unit-level mirror of cv2.goodFeaturesToTrack on a resident frame.  mask: w*h u8 (host or device)
 * or NULL; alternatively exclude_pts (n x 2 float32): the reference's mask, 255 everywhere and 0 at
 * int(x), int(y) of every point (vo_pipeline.py:371-378).  corners: cap x 2 float32 (x, y) in cv2's
 * order.  eig_out: optional w*h float32 = cv2.cornerMinEigenVal(frame, blockSize, ksize=3).
 * info = {candidates (local maxima above the threshold), candidates sorted (+ 2^30 if the single-CTA
 * kernel produced the result), selection rounds, selection-kernel kilocycles packed as
 * sort + 1000 * bucketing + 1000000 * rounds}.  Synchronises. */
int vo_good_features_to_track(vo_ctx* ctx, int64_t frame_key, const uint8_t* mask, const float* exclude_pts,
                              int n_exclude, const vo_gftt_params* params, float* corners, int cap, int* n_corners,
                              float* eig_out, int32_t info[4]);
/* find_new_candidates_shi + the append of vo_pipeline.py:564-591 without leaving the device: corners
 * of frame_key, masked at the previous-frame positions of the candidates the last KLT kept, become
 * candidates C := F := corners, T := T44 (NULL = the pose the last PnP left on the device).
 * No synchronisation.  Call after the promotion step of the same frame. */
int vo_detect_add_candidates_shi(vo_ctx* ctx, int64_t frame_key, const vo_gftt_params* params, const double T44[16]);

/* ---- trajectory / map bookkeeping of main.py:239-255 and a headless top-down plot
 * (main.py:262-271, :296-305) on the device.
 * vo_map_accumulate: pos = (inv(T) @ [0,0,0,1])[0:3] is appended to the path (poses_plotting), and the
 * landmark rows X with r_min < ||X - pos|| < r_max (reference: 3 and 50) to the map (X_plotting), in
 * order.  T44 NULL = the pose the last PnP left on the device.  use_prev_state 1 = the landmarks the
 * current frame STARTED from (main.py:248 reads vision.state, which is the previous iteration's), 0 =
 * the current ones.  n_added NULL and T44 NULL: nothing synchronises. */
int vo_map_reset(vo_ctx* ctx);
int vo_map_accumulate(vo_ctx* ctx, const double T44[16], double r_min, double r_max, int use_prev_state, int* n_added);
/* traj: n_traj x 3, pts: n_pts x 3 float64 (either may be NULL to query the counts), pos: last camera
 * position.  Synchronises. */
int vo_map_get(vo_ctx* ctx, double* traj, int traj_cap, int* n_traj, double* pts, int pts_cap, int* n_pts, double pos[3]);
/* X-Z plot into bgr (h x w x 3 u8, host or device): white canvas, black box `margin` px inside the
 * border, landmarks as blue discs of `radius` px, path as a red 3 px polyline on top; axis limits =
 * data range + 5 % (matplotlib's default margins), returned in bounds = {xlo, xhi, zlo, zhi}.  No text
 * (no font rasteriser).  Synchronises. */
int vo_map_render(vo_ctx* ctx, uint8_t* bgr, int w, int h, int margin, int radius, double bounds[4]);

/* One frame of main.py:216-284 with a single host synchronisation: upload (img may be NULL if
 * next_key is resident), KLT of P and C, PnP-RANSAC + LM + pruning, promotion, append of
 * new_candidates (n_new x 2 float64; the reference's detectors stay on the CPU). */
int vo_process_frame(vo_ctx* ctx, const uint8_t* img, int w, int h, int pitch, int64_t prev_key, int64_t next_key,
                     const double K[9], const vo_klt_params* kparams, const vo_pnp_params* pparams, double alpha_deg,
                     const double* new_candidates, int n_new, double T44[16], vo_counters* counters);

/* The same with the frames overlapped: the host waits for the pose only, so `counters` holds the counts
 * as of the PnP: n_landmarks = n_inliers (the n_promoted rows are appended behind the host's back) and the
 * candidates after the classification but before the append of new_candidates.  If a frame has been handed
 * over with vo_prefetch_frame since the last call, the tracking step (next_key -> that frame) is enqueued
 * right behind the append - and so is its PnP (compaction, RANSAC kernel, refit), launched with this call's
 * K and pnp parameters and gated on this frame's pose on the device.  The next call for that frame pair finds
 * both done (a call with other K / pnp parameters gets its own PnP: the one enqueued ahead is taken back);
 * any other call takes the step back (vo_track_rollback).  While the step is in flight vo_get_state /
 * vo_get_counters show the landmarks its tracking step kept, not yet pruned.  With the stage timers on it
 * behaves like vo_process_frame. */
int vo_process_frame_pipelined(vo_ctx* ctx, const uint8_t* img, int w, int h, int pitch, int64_t prev_key,
                               int64_t next_key, const double K[9], const vo_klt_params* kparams,
                               const vo_pnp_params* pparams, double alpha_deg, const double* new_candidates, int n_new,
                               double T44[16], vo_counters* counters);

/* Stage timers of vo_process_frame (CUDA events on the ctx stream; bench.py's roofline leg).
 * ms[6] = {upload+pyramid, landmark KLT, landmark compaction, PnP-RANSAC+LM, promotion+append (incl.
 * the wait for the candidates' KLT), candidates' KLT on the side stream (overlaps stages 1-3)} summed
 * over `frames` frames since vo_set_profiling(ctx, 1). */
int vo_set_profiling(vo_ctx* ctx, int enable);
int vo_get_stage_times(vo_ctx* ctx, double ms[6], int* frames);

/* ---- unit-level entry point mirroring cv2.solvePnPRansac (vo_pipeline.py:258-264) with P3P
 * hypotheses.  minimal_sets: max_iters x 4 int32 or NULL (cv::RNG(2^64-1) emulation = what cv2
 * draws).  All outputs optional.  info = {iterations run, winning iteration, inliers, LM iters}.
 * Does not touch the Markov state.  Synchronises. */
int vo_solve_pnp_ransac(vo_ctx* ctx, const double* X, const float* P, int n, const double K[9],
                        const vo_pnp_params* params, const int32_t* minimal_sets, double T44[16], double Th12[12],
                        double rvec[3], double tvec[3], uint8_t* inlier_mask, int32_t info[4], int* ok);
/* ---- vo_solve_pnp_ransac for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that
 * forbid cudaMalloc.  Same arguments, kernels and results as vo_solve_pnp_ransac, but the call allocates
 * nothing (no cudaMalloc / cudaFree / allocating library routine): every device buffer of the run - staged
 * points, float32 pack, minimal sets, hypotheses, counts, inlier mask, inlier pack, cv::RNG draw table - is a
 * slice of one file-scope `static __device__` array of < 1 MiB that every GPU loading the library holds until
 * exit.  Limits: n <= VO_NOALLOC_MAX_POINTS and params->max_iters <= VO_NOALLOC_MAX_ITERS, beyond them
 * VO_ERR_CAPACITY (vo_last_error names the limit).  The array is shared by ALL calls on a device (all
 * contexts) and keeps the last call's contents; the call clears what must start at zero itself, but calls
 * that overlap on the GPU (two contexts / host threads at once) race - serialise them.  Synchronises. */
#define VO_NOALLOC_MAX_POINTS 4096
#define VO_NOALLOC_MAX_ITERS 1024
int vo_solve_pnp_ransac_noalloc(vo_ctx* ctx, const double* X, const float* P, int n, const double K[9],
                                const vo_pnp_params* params, const int32_t* minimal_sets, double T44[16],
                                double Th12[12], double rvec[3], double tvec[3], uint8_t* inlier_mask,
                                int32_t info[4], int* ok);
/* ---- unit-level entry point mirroring cv2.triangulatePoints (vo_pipeline.py:666, :172).
 * M1: 3x4 (per_point_m1=0) or n x 12; X4: n x 4 homogeneous, unit norm.  Synchronises. */
int vo_triangulate_points(vo_ctx* ctx, const double* M1, const double* M2, int per_point_m1, const double* p1,
                          const double* p2, int n, double* X4);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* VO_B200_H */
