// Internal context of libvo_b200.so (not part of the ABI).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/vo_b200.h"

#include <nvtx3/nvToolsExt.h>

namespace vo {

// NVTX range over the host-side enqueue of a stage (header-only NVTX v3: a no-op costing a few ns unless a tool
// such as ncu / nsys injects itself); the stage names are the rows of DESIGN.md's kernel table
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};
#define VO_NVTX(name) ::vo::NvtxRange _nvtx_range_##__LINE__(name)

constexpr int kPad = 32;        // border (px) around every pyramid level, REFLECT_101 filled
constexpr int kMaxLevels = 8;   // hard cap on pyramid levels
constexpr int kFrameSlots = 4;  // resident frame pyramids (LRU)

// One pyramid level in HBM: u8 intensities with a kPad border and an int16x2 Scharr plane
// (zero outside the image).  `img`/`deriv` point at the padded origin (-kPad,-kPad).
struct LevelDesc {
    uint8_t* img;
    short2* deriv;
    int w, h;
    int stride;  // elements per padded row (same for img and deriv)
};

struct PyramidDesc {
    LevelDesc lv[kMaxLevels];
    int nlevels;
    // host side only: TMA descriptors of the u8 planes of levels 0..3 (FrameSlot::tmap_img), or NULL
    const CUtensorMap* tmaps;
};

struct FrameSlot {
    int64_t key = 0;
    bool valid = false;
    uint64_t stamp = 0;      // LRU
    int w = 0, h = 0;
    uint8_t* raw = nullptr;  // w*h contiguous copy of the uploaded frame
    uint8_t* img_pool = nullptr;
    short2* deriv_pool = nullptr;
    size_t raw_bytes = 0, img_bytes = 0, deriv_bytes = 0;
    PyramidDesc pyr{};
    // one TMA descriptor per pyramid level over the padded u8 plane (box: the KLT kernel's 64 B x 32 row window tile)
    alignas(64) CUtensorMap tmap_img[4];
    bool tmap_valid = false;
    // frame ingest (vo_prefetch_frame): the slot is being filled on the ingest stream; `ready` is
    // recorded there behind the last pyramid level and the first reader on the main stream waits on it
    cudaEvent_t ready = nullptr;
    bool pending = false;
    uint64_t ingest_seq = 0;
    uint64_t use_step = 0;   // Ctx::step_id when a stage last looked the slot up
};

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;  // elements
};

// ---- Shi-Tomasi detector (det.cu; reference vo_pipeline.py:357-390) -----------------------------
struct DetScalars {
    unsigned max_key;   // order-preserving key of the masked maximum of the min-eigenvalue map (0: none)
    int32_t n_cand;     // local maxima above the quality threshold
    int32_t n_out;      // corners accepted
    int32_t overflow;   // reserved (0)
    int32_t n_sorted;   // candidates the selection kernel sorted
    int32_t rounds;     // parallel rounds the greedy selection needed
    int32_t cyc[3];     // SM cycles of the selection kernel's phases: sort, bucketing, rounds + compaction
    int32_t fallback;   // the cluster kernel left the selection to the single-CTA kernel
};

struct Detector {
    DevBuf<float> eig;                  // w*h min-eigenvalue map
    DevBuf<uint8_t> mask;               // w*h
    DevBuf<unsigned long long> keys;    // (value key << 32 | pixel index), power-of-two capacity >= w*h
    DevBuf<int> cell_cnt, cell_start;   // candidates per minDistance cell, exclusive prefix sum
    DevBuf<int> items;                  // candidate ranks bucketed by cell
    DevBuf<unsigned> xy;                // (x | y << 16) of the candidates in sorted order
    DevBuf<uint8_t> state;              // 0 undecided, 1 accepted, 2 rejected
    DevBuf<int> ghist;                  // key histogram + round flags of the cluster kernel
    DevBuf<float> out;                  // accepted corners, float2, cv2's order
    DetScalars* d_sc = nullptr;
    DetScalars* h_sc = nullptr;         // pinned mirror
    int out_cap = 0;
};

struct Ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    // side stream: the candidates' KLT runs here, concurrently with the landmark chain
    // (landmark KLT -> compaction -> PnP-RANSAC) on `stream`; joined before the promotion step
    cudaStream_t stream2 = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool side_inflight = false;
    bool side_writes_counts = false;   // a deferred candidate append is part of the side stream's work
    // ingest stream: upload + pyramid of the NEXT frame while the current one is tracked
    cudaStream_t stream3 = nullptr;
    // pose stream: LM refit, pose read-back and the pose-dependent rest of the promotion, while the main stream
    // already tracks the next frame (vo_process_frame_pipelined)
    cudaStream_t stream4 = nullptr;
    cudaEvent_t ev_ransac = nullptr, ev_pose_side = nullptr;
    cudaEvent_t ev_snap = nullptr, ev_pose_done = nullptr;   // counters snapshot taken (main) / pose-side work of the frame done
    bool pose_inflight = false;   // work of the last pipelined frame may still be running on stream4
    // tracking stream: the NEXT frame's landmark KLT, started beside the current frame's RANSAC kernel over the rows
    // the tracking step left (before the pruning) plus the rows the frame is about to append - see Pipeline::src_*
    cudaStream_t stream5 = nullptr;
    cudaEvent_t ev_compact = nullptr, ev_trk = nullptr;   // behind the last landmark compaction (main) / the KLT on stream5
    cudaEvent_t ev_ingest_fork = nullptr;
    uint64_t ingest_seq = 0;
    int64_t last_ingest_key = 0; bool last_ingest_valid = false;   // newest frame on the ingest stream, not used yet
    // A "step" closes when the promotion stage has been enqueued (it joins the side stream): from then
    // on no kernel of that step reads a frame slot, and ev_step_end (main stream) marks the point.
    cudaEvent_t ev_step_end[4] = {};   // ring: closing step S records ev_step_end[S & 3]
    uint64_t step_id = 1;        // id of the open step
    int open_uses = 0;           // slot look-ups in the open step
    uint8_t* pinned_stage2 = nullptr;   // pageable sources of the ingest stream are staged here
    // a frame announced with vo_prefetch_frame_deferred: put on the ingest stream by the next pose call,
    // between its enqueue and its wait (or by whoever asks for the key first)
    struct { bool set = false; const uint8_t* img = nullptr; int w = 0, h = 0, pitch = 0; int64_t key = 0; } hint;
    int max_w = 0, max_h = 0;
    int sm_count = 148;
    std::string err;
    int64_t launches = 0;
    uint64_t clock = 0;
    FrameSlot slots[kFrameSlots];
    uint8_t* pinned_stage = nullptr;
    size_t pinned_stage_bytes = 0;

    // unit-level scratch (KLT)
    DevBuf<float> u_pts, u_out, u_err;
    DevBuf<uint8_t> u_status;
    DevBuf<int32_t> u_iters;
    DevBuf<int32_t> u_n;
    // forward-backward check scratch
    DevBuf<float> fb_pts, fb_pts2;
    DevBuf<uint8_t> fb_status, fb_status2;

    Detector det;

    struct Pipeline* pipe = nullptr;  // device-resident Markov state + PnP scratch
    struct MapState* map = nullptr;   // trajectory / landmark map accumulated on the device (map.cu)
};

extern thread_local std::string g_create_error;

int set_error(Ctx* c, int code, const char* fmt, ...);

#define VO_CUDA(c, expr)                                                                       \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return vo::set_error((c), VO_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,             \
                                 cudaGetErrorString(_e), __FILE__, __LINE__);                  \
    } while (0)

#define VO_TRY(expr)                \
    do {                            \
        int _r = (expr);            \
        if (_r != VO_OK) return _r; \
    } while (0)

// ---- programmatic dependent launch ---------------------------------------------------------------
// The frame is a chain of small dependent kernels, so the ~2 us a kernel boundary costs (drain,
// launch, ramp-up) is paid a dozen times per frame.  Kernels launched through launch_pdl may be
// made resident while their predecessor in the stream is still running; every such kernel starts
// with pdl_prologue(), which lets ITS successor do the same and then blocks until the predecessor
// grid has completed and its writes are visible (griddepcontrol.wait) - before touching any data.
// Without the launch attribute both instructions are no-ops.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_prologue()
{
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

template <typename T>
int ensure(Ctx* c, DevBuf<T>& b, size_t n, bool keep = false)
{
    if (n <= b.cap) return VO_OK;
    size_t ncap = b.cap ? b.cap : 256;
    while (ncap < n) ncap *= 2;
    T* np_ = nullptr;
    cudaError_t e = cudaMalloc((void**)&np_, ncap * sizeof(T));
    if (e != cudaSuccess) return set_error(c, VO_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", ncap * sizeof(T), cudaGetErrorString(e));
    if (b.p) {
        if (keep) {
            e = cudaMemcpyAsync(np_, b.p, b.cap * sizeof(T), cudaMemcpyDeviceToDevice, c->stream);
            if (e != cudaSuccess) return set_error(c, VO_ERR_CUDA, "grow copy failed: %s", cudaGetErrorString(e));
        }
        // stream-ordered w.r.t. earlier kernels still reading the old buffer
        cudaStreamSynchronize(c->stream);
        if (c->stream4) cudaStreamSynchronize(c->stream4);
        if (c->stream5) cudaStreamSynchronize(c->stream5);
        if (c->stream2) cudaStreamSynchronize(c->stream2);
        cudaFree(b.p);
    }
    b.p = np_;
    b.cap = ncap;
    return VO_OK;
}

// ---- device-resident Markov state (reference: BestVision.state / .extended_state,
// /root/reference/src/vo_pipeline.py:36-37) and PnP / promotion scratch ------------------------
constexpr int kMinimal = 4;          // P3P on 3 points + 1 disambiguation point
constexpr int kLandBufs = 5;         // landmark buffers in rotation (state before the last two tracking steps + compaction and
                                     // pruning outputs of the current frame and of the step enqueued ahead)

// small block of scalars living in device memory; mirrored to pinned host memory on demand
struct DevScalars {
    int32_t n_land;        // |P| = |X|
    int32_t n_cand;        // |C| = |F| = |T|
    int32_t n_lost;        // landmarks lost by the last KLT
    int32_t n_tracked;     // landmarks kept by the last KLT (input size of the last PnP)
    int32_t n_promoted;    // candidates promoted by the last promotion step
    int32_t n_cand_lost;   // candidates lost by the last KLT
    int32_t cf_alias;      // reference quirk App. B-5: C and F are the same ndarray
    int32_t cand_was_empty;// C was empty after the tracking/removal step (decides cf_alias on add)
    // RANSAC
    uint64_t rng;          // position in the precomputed cv::RNG(2^64-1) output stream
    int32_t niters;        // current (shrinking) iteration budget
    int32_t best_count;
    int32_t best_hyp;      // -1: none
    int32_t iters_run;     // number of loop iterations the sequential reference would have run
    int32_t done;          // scan passed niters
    int32_t sets_ready;    // rows of the minimal-set table generated so far
    int32_t n_inliers;
    int32_t ok;
    int32_t lm_iters;
    int32_t klt_iters;     // sum of LK iterations (roofline accounting)
    int32_t rng_exhausted; // the RNG draw table ran out (host reports VO_ERR_CAPACITY)
    int32_t ticket;        // multi-CTA kernels: last-block-done counter (always 0 between kernels)
    int32_t prev_n_land;   // |P| before the last tracking step (rows of the `prev` landmark buffer)
    int32_t prev2_n_land;  // ... and before the tracking step before that (`prev2` buffer)
    // values before the last promotion step (rollback of a speculative promotion)
    int32_t spec_n_land, spec_n_cand, spec_cf_alias, spec_cand_was_empty;
    int32_t finalized;     // the refit of the current RANSAC run has been done
    int32_t klt_iters_cand; // sum of LK iterations over the candidates (roofline accounting)
    int32_t ticket2;       // same for kernels enqueued on the side stream (they overlap main-stream kernels)
    int32_t klt_iters_cand2; // side-stream accumulator of klt_iters_cand (early classification), moved over by the finish kernel
    int32_t n_appended;    // rows the last landmark append added (top-up decided on the device)
    int32_t epoch;         // round flag of the persistent RANSAC kernel (run id << 12 | round)
    int32_t prom_base;     // row of P/X where the promoted rows of the frame start (split promotion)
    // counters as of the pose (vo_process_frame_pipelined reads the block back while the main stream goes on)
    int32_t snap_n_land, snap_n_cand, snap_cf_alias, snap_n_appended;
    int32_t n_cand_early;  // |C| as the last early classification left it (before the frame tail appends the new candidates)
    int32_t gate_ok;       // `ok` of the last pipelined frame as its frame tail saw it: what kernels enqueued ahead are gated on
    double T[16];          // refined pose, row-major 4x4 world->camera
    double Th[12];         // winning unrefined hypothesis [R|t]
    double cost;           // final sum of squared reprojection errors over the inliers
    // %globaltimer (ns) stamps of the CTA that finished the last RANSAC run (vo_debug_pnp_stamps):
    // entry, P3P done, hypotheses scored, chunk bookkeeping done, inlier mask done, LM done, pruning done
    unsigned long long stamps[8];
};

struct Pipeline {
    int track_capacity = 0, hyp_capacity = 0;
    // landmarks, kLandBufs buffers in rotation (cur = index of the live ones): the post-KLT compaction
    // writes cur+1, the inlier pruning cur+2, so the state a frame started from (prev) stays intact
    // until the next frame's compaction - main.py:248 reads it after estimatePose
    int cur = 0;
    int prev = -1;          // buffer holding the state before the last tracking step (-1: none)
    DevBuf<float> P[kLandBufs];     // n x 2
    DevBuf<double> X[kLandBufs];    // n x 3
    int n_land_upper = 0;   // host-known upper bound of n_land
    int prev_n_upper = 0;   // same for the rows of the `prev` buffer
    // the state before the tracking step before that: with a tracking step enqueued one frame ahead (inside the pose
    // call) this is the buffer main.py:248 reads - the rotation is long enough to keep it intact
    int prev2 = -1, prev2_n_upper = 0;
    // candidates, ping-pong
    int ccur = 0;
    DevBuf<float> C[2];     // n x 2 (current observation; float32-exact)
    DevBuf<double> F[2];    // n x 2 (first observation)
    DevBuf<int32_t> Ti[2];  // n (index into the pose table)
    int n_cand_upper = 0;
    DevBuf<double> poses;   // pose table, 16 doubles per entry (4x4 row-major)
    int n_poses = 0;
    // KLT outputs for landmarks + candidates (one fused launch)
    DevBuf<float> klt_in, klt_out, klt_err;
    DevBuf<uint8_t> klt_status;
    DevBuf<int32_t> klt_iters;
    // second set: a candidates' KLT that runs ahead of the current frame's pose writes here (the sets are swapped), so
    // the current frame's results survive until its promotion step is settled (no pose: it is redone from them)
    DevBuf<float> klt_out2, klt_err2; DevBuf<uint8_t> klt_status2; DevBuf<int32_t> klt_iters2; bool klt_swapped = false;
    DevBuf<float> lost;     // positions of lost landmark tracks (debug output of a1)
    int64_t cand_tracked_prev = -1, cand_tracked_next = -1;  // frames of the cached candidate KLT
    bool cand_tracked_valid = false;
    // PnP scratch
    DevBuf<float> q4;       // n x 4: float32(X) and u
    DevBuf<float> qv;       // n: v
    DevBuf<int32_t> sets;   // H x 4
    DevBuf<uint32_t> rng_tab; // raw 32-bit outputs of cv::RNG(2^64-1), in draw order
    int rng_tab_len = 0;
    DevBuf<int32_t> blk_counts; // per-CTA counts of the multi-CTA compactions (3 per CTA)
    DevBuf<uint8_t> cand_flag;  // per candidate: 0 lost, 1 stays, 2 promoted
    DevBuf<double> hyp;     // H x 12 [R|t]
    DevBuf<int32_t> counts; // H (-1 = no model)
    DevBuf<uint8_t> mask;   // n
    DevBuf<float> q4i, qvi; // the inliers of the last RANSAC, packed for the refit kernel
    unsigned pnp_run = 0;   // RANSAC launches so far (epoch base of the persistent kernel)
    DevBuf<uint8_t> promote;// n_cand
    DevBuf<double> stage_d; // host->device staging (unit-level calls, candidate feed)
    DevBuf<float> stage_f;
    DevBuf<int32_t> stage_i;
    // deferred append riding on the side stream (vo_klt_track_landmarks_add): double-buffered staging
    double* h_add[2] = { nullptr, nullptr }; size_t h_add_cap = 0; int add_flip = 0;
    DevBuf<double> d_add[2];
    // landmark rows waiting for the next vo_process_frame (vo_queue_landmarks), and the pinned staging ring
    // pageable rows are copied into (the append kernel reads it over the unified address space)
    struct LandSeg { const float* P; const double* X; int n; int target; };
    LandSeg land_q[4]; int n_land_q = 0;
    double* h_land[4] = { nullptr, nullptr, nullptr, nullptr }; size_t h_land_cap[4] = { 0, 0, 0, 0 }; int land_flip = 0;
    cudaEvent_t ev_land_slot[4] = {}; int land_slot_used = -1;
    int klt_b_off = 0;      // offset of the candidate results inside the klt_* buffers
    bool packed_valid = false;  // q4/qv hold the current landmarks (set by the post-KLT compaction)
    // the last promotion step consumed a candidate KLT: C[ccur ^ 1] holds the candidates' previous
    // positions, klt_status + mask_src_off their status, d_sc->spec_n_cand their number (detector mask)
    bool mask_src_valid = false; int mask_src_off = 0;
    // early classification (tri.cu): the candidate side of the promotion step - lost rows, bearing test,
    // compaction of C/F/T - already ran on the side stream behind the candidates' KLT with this alpha;
    // the promoted rows wait in prom_rows / prom_ti for the pose (promote_finish_kernel)
    bool early_valid = false; double early_alpha = 0.; int early_ncand_upper = 0;
    DevBuf<double> prom_rows; DevBuf<int32_t> prom_ti;
    // ---- tracking ahead of the pruning (pipelined frames) -------------------------------------------------------
    // A track depends on nothing but its own start position and the two frames, so the next frame's landmark KLT
    // need not wait for the RANSAC kernel to prune the landmarks: it runs over a SUPERSET of the rows the frame will
    // hold - list 0: the rows the last compaction kept (buffer b1_buf, sc->n_tracked rows), list 1: the rows the early
    // classification parked for promotion (prom_P, sc->n_promoted rows), lists 2..: the queued batches (every offered
    // row) - and every landmark row carries the index of its track in that superset (Lsrc, written by the pruning
    // and by the frame-tail kernel); the next compaction looks status and position up through it.  Same results.
    DevBuf<float> lk_out[2], lk_err[2]; DevBuf<uint8_t> lk_status[2]; DevBuf<int32_t> lk_iters[2]; int lk_flip = 0;
    DevBuf<int32_t> Lsrc[kLandBufs];   // per landmark buffer: row -> superset track index
    DevBuf<float> prom_P;              // float2 positions of the parked rows (list 1)
    int b1_buf = -1, b1_upper = 0;     // buffer the last compaction wrote, and the bound it was launched with
    bool src_valid = false;            // Lsrc[cur] describes the live landmarks under the plan below
    int src_prom_base = 0, src_prom_upper = 0, src_n_segs = 0, src_total = 0;
    int src_seg_base[4] = {}, src_seg_n[4] = {}; const float* src_seg_P[4] = {};
    DevBuf<double> feed_d[2]; int feed_flip = 0;   // the frame's new candidates (pipelined frames: also list 1 of the candidates' KLT)
    cudaEvent_t ev_feed = nullptr;     // behind the feed's H2D copy (main stream)
    // speculative promotion (vo_pnp_ransac_promote): host-side values to restore on rollback
    bool spec_pending = false;
    // host-side values before the last tracking step (vo_track_rollback)
    // vo_process_frame_pipelined: the tracking step (ahead_prev -> ahead_next) is already enqueued
    bool ahead_valid = false; int64_t ahead_prev = 0, ahead_next = 0;
    // ... and so is its PnP (RANSAC kernel + refit, gated like the tracking step): the next call only has to ask for
    // the same camera matrix and parameters; pa_*: what to put back if it asks for something else
    bool pnp_ahead = false; double pa_K[9] = {}; vo_pnp_params pa_prm = {}; int pa_cur = 0, pa_prev = -1, pa_prev2 = -1;
    double keep_T[16] = {}; int keep_ok = 0;   // the last frame's pose as the host got it (restored on the device by an undo)
    bool rb_valid = false; int rb_cur = 0, rb_prev = -1, rb_prev_n_upper = 0, rb_n_land_upper = 0, rb_prev2 = -1, rb_prev2_n_upper = 0;
    double trace_wait_us = 0.; int trace_ready = 0, trace_n = 0;   // VO_TRACE=1 diagnostics
    cudaEvent_t ev_land = nullptr;   // behind the landmark KLT launch (candidates' KLT ordered after it, see track_both)
    cudaEvent_t ev_pose = nullptr;   // behind the scalar read-back between PnP and a speculative promotion
    int64_t spec_prev_key = -1, spec_next_key = -1;
    DevScalars* d_sc = nullptr;
    DevScalars* h_sc = nullptr;  // pinned mirror
    // pinned snapshot of the landmarks before a tracking step (serves main.py:248)
    float* h_snapP = nullptr; double* h_snapX = nullptr; DevScalars* h_snap_sc = nullptr;
    size_t snap_cap = 0; int snap_rows = 0;
    // optional stage timers (vo_set_profiling)
    bool profiling = false; cudaEvent_t ev[6] = {}; cudaEvent_t ev_side[2] = {}; double stage_ms[6] = {}; int stage_frames = 0;
};

// ---- stage launchers implemented in the .cu files -------------------------------------------
int build_pyramid(Ctx* c, FrameSlot& s, int max_level_cap, cudaStream_t st = nullptr);
int launch_klt(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const float* d_pts,
               const int32_t* d_n, int n_upper, float* d_out, uint8_t* d_status, float* d_err,
               int32_t* d_iters, const vo_klt_params& prm);
// two track lists in one launch (landmarks + candidates): list B starts at element offset
// `b_off` of out/status/err/iters
int launch_klt2(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const float* d_ptsA,
                const int32_t* d_nA, int nA_upper, const float* d_ptsB, const int32_t* d_nB, int nB_upper,
                int b_off, float* d_out, uint8_t* d_status, float* d_err, int32_t* d_iters,
                const vo_klt_params& prm, cudaStream_t st = nullptr, const int32_t* gate = nullptr);
// up to kKltSegs track lists in one launch: list k = pts[k][0 .. min(*n_dev[k], n_upper[k])), results at element
// offset out_off[k] (+ *off_dev[k]) of out/status/err/iters (n_dev[k] == nullptr: exactly n_upper[k] rows)
constexpr int kKltSegs = 6;
struct KltSegs {
    int n;
    const float2* pts[kKltSegs];
    const int32_t* n_dev[kKltSegs];
    int n_upper[kKltSegs];
    int out_off[kKltSegs];
    const int32_t* off_dev[kKltSegs];   // optional: added to out_off on the device (a list appended behind a live count)
    uint8_t f64[kKltSegs];              // the list holds double2 rows (converted to float32 like the candidate append does)
};
int launch_klt_segs(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const KltSegs& segs, float* d_out,
                    uint8_t* d_status, float* d_err, int32_t* d_iters, const vo_klt_params& prm,
                    cudaStream_t st = nullptr, const int32_t* gate = nullptr);
// make `stream` wait for the work in flight on the side stream (no-op if there is none)
int join_side_stream(Ctx* c);
// make `stream` wait for the pose-side work of the last pipelined frame (LM refit, X of the promoted rows)
int join_pose_stream(Ctx* c);
// put a deferred frame hint on the ingest stream now (no-op if there is none)
int flush_frame_hint(Ctx* c, bool under_current_stages = false);
// the promotion stage of a frame has been enqueued: close the step (see Ctx::ev_step_end)
int close_step(Ctx* c);

// optional extra (off in parity mode): track `out` back into `prev` and clear status where the
// round trip misses the start point by more than prm.fb_thresh
int klt_fb_check(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const float* d_pts,
                 const int32_t* d_n, int n_upper, const float* d_out, uint8_t* d_status,
                 const vo_klt_params& prm, cudaStream_t st = nullptr, int which = 0, const int32_t* gate = nullptr);
FrameSlot* get_frame(Ctx* c, int64_t key);
int pipeline_init(Ctx* c, int track_capacity, int hyp_capacity);
void pipeline_free(Ctx* c);

// pnp.cu: RANSAC over the packed points (q4,qv; n on device at sc->n_tracked or given),
// then mask + LM refit + (optionally) compaction of the landmark buffers by the inlier mask.
// The landmark compaction of a tracking step done by the RANSAC kernel itself (its first phase, all CTAs; see
// pnp_ransac_kernel): saves the compaction kernel and the kernel boundary behind it on the frame's chain.
constexpr int kFuseMaxRows = 8192;      // 256 scan chunks x 32 rows
struct FusedCompact {
    int on, n_upper;
    const float2* klt_out; const uint8_t* status; const int32_t* iters;   // the tracking step's results (row = track)
    const double* X_in;                                                    // 3-D points of the rows that were tracked
    float2* P_out; double* X_out; float2* lost;                            // kept rows (new position, 3-D point), lost positions
    float4* q4; float* qv;                                                 // the float32 PnP pack of the kept rows
};

struct PnpLaunch {
    const float* q4; const float* qv;   // packed float32 points
    const int32_t* d_n; int n_upper;    // point count (device) and its host upper bound
    const int32_t* d_sets_in;           // explicit H x 4 table or nullptr (cv::RNG emulation)
    double K[9];
    vo_pnp_params prm;
    bool compact;                       // prune landmark buffers to the inliers (estimatePose)
    bool refit_on_pose_stream = false;  // LM refit on Ctx::stream4: the landmark side of the frame goes on without it
    const int32_t* gate = nullptr;      // device flag: the RANSAC kernel does nothing when it is 0 (a run enqueued ahead of a
                                        // pose that may not come); the refit then finds ok == 0 and writes "no pose"
    bool fuse_compact = false;          // the RANSAC kernel compacts the pending tracking step first (process_frame_impl)
    bool small_grid = false;            // kernels of the next frame run beside the RANSAC kernel: leave them most of the SMs
};
int launch_pnp(Ctx* c, const PnpLaunch& L);
// raw cv::RNG output table on the device, large enough for H iterations (grows, never shrinks)
int ensure_rng_table(Ctx* c, int H);
// tri.cu
// after KLT: keep status==1, pack q4/qv.  gate (device flag, may be NULL): the kernel does nothing when it is 0
// - for work enqueued ahead of a pose that may not come (vo_process_frame_pipelined)
int launch_landmark_compact(Ctx* c, int n_upper, const int32_t* gate = nullptr, bool from_superset = false);
// Tnew == nullptr: take the pose the last PnP left on the device (fused per-frame call)
// split (pipelined frames, early classification done): the promoted rows' positions are appended on the main
// stream at once, their 3-D points (DLT with the refined pose) and the pose-table row `pose_idx` (-1: none)
// follow on the pose stream
int launch_promote(Ctx* c, const double K[9], const double* Tnew, double alpha_deg, int ncand_upper,
                   bool use_klt, int b_off, bool split = false, int pose_idx = -1);
// the pose-independent half of the promotion step on the side stream, right behind the candidates' KLT
int launch_promote_early(Ctx* c, double alpha_deg, int ncand_upper, int b_off, const int32_t* gate = nullptr);
// take an early classification back (the candidates are again as the KLT found them); no-op if there is none
int early_undo(Ctx* c);
// d_pts: n_new x 2 doubles on the device; d_Thost: 16 doubles on the device or nullptr (= device pose)
// d_n_new: optional device count (the rows actually appended are min(*d_n_new, n_new))
int launch_cand_add(Ctx* c, const double* d_pts, int n_new, const double* d_Thost, const int32_t* d_n_new = nullptr,
                    bool on_side_stream = false, bool write_pose = true);
// map.cu
void map_free(Ctx* c);
// det.cu
void gftt_defaults(vo_gftt_params* p);
int detector_free(Ctx* c);
int launch_gftt(Ctx* c, const FrameSlot& f, const uint8_t* d_mask, const vo_gftt_params& prm);
int build_point_mask(Ctx* c, int w, int h, const float* d_pts, const uint8_t* d_status, const int32_t* d_n, int n_upper);
int corners_to_stage(Ctx* c, double* d_stage);
int launch_frame_tail(Ctx* c, const double K[9], double alpha_deg, int b_off, const double* d_feed, int n_new,
                      const Pipeline::LandSeg* segs, int n_segs);
int launch_land_append(Ctx* c, const float* d_P, const double* d_X, int n_rows, int target, const int32_t* gate = nullptr);
int launch_triangulate(Ctx* c, const double* d_M1, const double* d_M2, int per_point, const double* d_p1,
                       const double* d_p2, int n, double* d_X4);

}  // namespace vo

struct vo_ctx : public vo::Ctx {};
