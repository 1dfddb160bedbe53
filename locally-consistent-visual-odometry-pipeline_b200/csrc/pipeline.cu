// Pipeline-level C-ABI: the device-resident Markov state (P, X | C, F, T) and the three per-frame
// stages of /root/reference/src/main.py:229-284 (associate -> estimatePose -> triangulate_landmark),
// plus unit-level entry points that mirror the cv2 signatures for differential tests.
#include <math.h>
#include <stdlib.h>
#include <time.h>
#include <utility>
#include "ctx.cuh"
#include "pnp_sets.cuh"

namespace vo {

int pipeline_init(Ctx* c, int track_capacity, int hyp_capacity)
{
    Pipeline* p = new (std::nothrow) Pipeline();
    if (!p) return set_error(c, VO_ERR_NOMEM, "out of host memory");
    c->pipe = p;
    p->track_capacity = track_capacity > 0 ? track_capacity : 4096;
    p->hyp_capacity = hyp_capacity > 0 ? hyp_capacity : 2048;
    VO_CUDA(c, cudaMalloc((void**)&p->d_sc, sizeof(DevScalars)));
    VO_CUDA(c, cudaMemsetAsync(p->d_sc, 0, sizeof(DevScalars), c->stream));
    VO_CUDA(c, cudaMallocHost((void**)&p->h_sc, sizeof(DevScalars)));
    memset(p->h_sc, 0, sizeof(DevScalars));
    const size_t n = p->track_capacity;
    for (int k = 0; k < kLandBufs; k++) { VO_TRY(ensure(c, p->P[k], n * 2)); VO_TRY(ensure(c, p->X[k], n * 3)); }
    for (int k = 0; k < 2; k++) {
        VO_TRY(ensure(c, p->C[k], n * 2)); VO_TRY(ensure(c, p->F[k], n * 2)); VO_TRY(ensure(c, p->Ti[k], n));
    }
    VO_TRY(ensure(c, p->klt_out, n * 4)); VO_TRY(ensure(c, p->klt_err, n * 2));
    VO_TRY(ensure(c, p->klt_status, n * 2)); VO_TRY(ensure(c, p->klt_iters, n * 2));
    VO_TRY(ensure(c, p->q4, n * 4)); VO_TRY(ensure(c, p->qv, n)); VO_TRY(ensure(c, p->lost, n * 2));
    VO_TRY(ensure(c, p->mask, n)); VO_TRY(ensure(c, p->poses, 16 * 1024));
    VO_TRY(ensure(c, p->sets, (size_t)p->hyp_capacity * kMinimal));
    VO_TRY(ensure(c, p->hyp, (size_t)p->hyp_capacity * 12));
    VO_TRY(ensure(c, p->counts, (size_t)p->hyp_capacity));
    return VO_OK;
}

void pipeline_free(Ctx* c)
{
    Pipeline* p = c->pipe;
    if (!p) return;
    for (int k = 0; k < kLandBufs; k++) { cudaFree(p->P[k].p); cudaFree(p->X[k].p); }
    for (int k = 0; k < 2; k++) { cudaFree(p->C[k].p); cudaFree(p->F[k].p); cudaFree(p->Ti[k].p); }
    cudaFree(p->poses.p); cudaFree(p->klt_in.p); cudaFree(p->klt_out.p); cudaFree(p->klt_err.p);
    cudaFree(p->klt_status.p); cudaFree(p->klt_iters.p); cudaFree(p->lost.p); cudaFree(p->q4.p); cudaFree(p->qv.p);
    cudaFree(p->sets.p); cudaFree(p->hyp.p); cudaFree(p->counts.p); cudaFree(p->mask.p); cudaFree(p->promote.p);
    cudaFree(p->q4i.p); cudaFree(p->qvi.p); cudaFree(p->prom_rows.p); cudaFree(p->prom_ti.p); cudaFree(p->prom_P.p);
    for (int k = 0; k < 2; k++) { cudaFree(p->lk_out[k].p); cudaFree(p->lk_err[k].p); cudaFree(p->lk_status[k].p); cudaFree(p->lk_iters[k].p); }
    for (int k = 0; k < kLandBufs; k++) cudaFree(p->Lsrc[k].p);
    cudaFree(p->klt_out2.p); cudaFree(p->klt_err2.p); cudaFree(p->klt_status2.p); cudaFree(p->klt_iters2.p);
    cudaFree(p->feed_d[0].p); cudaFree(p->feed_d[1].p);
    if (p->ev_feed) cudaEventDestroy(p->ev_feed);
    cudaFree(p->stage_d.p); cudaFree(p->stage_f.p); cudaFree(p->stage_i.p);
    cudaFree(p->rng_tab.p); cudaFree(p->blk_counts.p); cudaFree(p->cand_flag.p);
    for (int k = 0; k < 2; k++) { cudaFree(p->d_add[k].p); if (p->h_add[k]) cudaFreeHost(p->h_add[k]); }
    for (int k = 0; k < 4; k++) { if (p->h_land[k]) cudaFreeHost(p->h_land[k]); if (p->ev_land_slot[k]) cudaEventDestroy(p->ev_land_slot[k]); }
    cudaFree(p->d_sc);
    cudaFreeHost(p->h_sc);
    for (int k = 0; k < 6; k++) if (p->ev[k]) cudaEventDestroy(p->ev[k]);
    for (int k = 0; k < 2; k++) if (p->ev_side[k]) cudaEventDestroy(p->ev_side[k]);
    if (p->ev_pose) cudaEventDestroy(p->ev_pose);
    if (p->ev_land) cudaEventDestroy(p->ev_land);
    cudaFreeHost(p->h_snapP); cudaFreeHost(p->h_snapX); cudaFreeHost(p->h_snap_sc);
    delete p;
    c->pipe = nullptr;
}

// D2H of the scalar block + stream sync; tightens the host-side upper bounds
static int read_scalars(Ctx* c)
{
    Pipeline* p = c->pipe;
    VO_TRY(join_pose_stream(c));                               // the refined pose comes from there
    if (c->side_writes_counts) VO_TRY(join_side_stream(c));   // n_cand is being updated over there
    VO_CUDA(c, cudaMemcpyAsync(p->h_sc, p->d_sc, sizeof(DevScalars), cudaMemcpyDeviceToHost, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    p->n_land_upper = p->h_sc->n_land;
    p->n_cand_upper = p->h_sc->n_cand;
    return VO_OK;
}

// the RNG draw table ran out during the last RANSAC (needs > 16 draws per iteration on average)
static int check_rng(Ctx* c)
{
    if (c->pipe->h_sc->rng_exhausted)
        return set_error(c, VO_ERR_CAPACITY, "cv::RNG draw table exhausted during RANSAC");
    return VO_OK;
}

__global__ void set_counts_kernel(DevScalars* sc, int n_land, int n_cand, int alias)
{
    if (n_land >= 0) sc->n_land = n_land;
    if (n_cand >= 0) { sc->n_cand = n_cand; sc->cf_alias = alias; sc->cand_was_empty = n_cand == 0; }
}

// A deferred append of new candidates (vo_pipeline.py:574-591 of the PREVIOUS frame) that rides on the
// side stream in front of the candidates' KLT: host rows, copied through pinned staging
struct DeferredAdd { const double* pts; int n; const double* T44; };

// KLT of landmarks (list A) and candidates (list B, stored behind A at offset n_land_upper).  The
// landmark launch - the head of the chain the pose depends on - is enqueued FIRST; the candidates
// (usually most of the tracks, and with them the slowest stragglers) follow on the side stream, behind
// the optional deferred append, and are joined by the promotion step.
// superset: the landmark launch goes to the tracking stream and covers the rows of the last compaction (before the
// pruning), the rows parked for promotion and every row of the queued batches (Pipeline::src_*) - it starts as soon
// as that compaction and the early classification are done, beside the RANSAC kernel, and needs no gate: it writes
// nothing but its own result buffers (lk_*), which the gated compaction reads through Lsrc.
static int track_both(Ctx* c, FrameSlot* a, FrameSlot* b, const vo_klt_params& prm, bool with_candidates,
                      const DeferredAdd* add = nullptr, double early_alpha = -1., const int32_t* gate = nullptr,
                      bool superset = false, const double* d_feed = nullptr, int n_feed = 0)
{
    Pipeline* p = c->pipe;
    const int nA = p->n_land_upper;
    const int n_add = (add && with_candidates && add->n > 0) ? add->n : 0;
    const int nB = with_candidates ? p->n_cand_upper + n_add : 0;
    const size_t tot = (size_t)nA + nB > 0 ? (size_t)nA + nB : 1;
    VO_TRY(early_undo(c));          // (a classification nobody finished: its tracking is being replaced)
    VO_TRY(join_side_stream(c));
    // superset: the candidates' KLT does not wait for the frame tail either - list 0: the rows the last early
    // classification kept (sc->n_cand_early of them), list 1: the frame's new candidates straight from the staging
    // buffer (the tail appends them behind list 0, so their tracks sit where the classification will look)
    const bool cand_ahead = superset && nB > 0 && n_add == 0 && (n_feed == 0 || d_feed != nullptr);
    p->klt_swapped = false;
    if (cand_ahead) {   // (ungated: into the other set of result buffers)
        std::swap(p->klt_out, p->klt_out2); std::swap(p->klt_err, p->klt_err2);
        std::swap(p->klt_status, p->klt_status2); std::swap(p->klt_iters, p->klt_iters2);
        p->klt_swapped = true;
    }
    VO_TRY(ensure(c, p->klt_out, tot * 2)); VO_TRY(ensure(c, p->klt_err, tot));
    VO_TRY(ensure(c, p->klt_status, tot)); VO_TRY(ensure(c, p->klt_iters, tot));
    if (nB > 0) VO_CUDA(c, cudaEventRecord(c->ev_fork, c->stream));   // the side stream forks BEFORE the landmark launch
    if (superset) {
        const int f = p->lk_flip ^= 1;
        const size_t stot = p->src_total > 0 ? p->src_total : 1;
        VO_TRY(ensure(c, p->lk_out[f], stot * 2)); VO_TRY(ensure(c, p->lk_err[f], stot));
        VO_TRY(ensure(c, p->lk_status[f], stot)); VO_TRY(ensure(c, p->lk_iters[f], stot));
        KltSegs sg{};
        sg.n = 2 + p->src_n_segs;
        sg.pts[0] = (const float2*)p->P[p->b1_buf].p; sg.n_dev[0] = &p->d_sc->n_tracked; sg.n_upper[0] = p->b1_upper; sg.out_off[0] = 0;
        sg.pts[1] = (const float2*)p->prom_P.p; sg.n_dev[1] = &p->d_sc->n_promoted; sg.n_upper[1] = p->src_prom_upper; sg.out_off[1] = p->src_prom_base;
        for (int k = 0; k < p->src_n_segs; k++) {
            sg.pts[2 + k] = (const float2*)p->src_seg_P[k]; sg.n_dev[2 + k] = nullptr; sg.n_upper[2 + k] = p->src_seg_n[k];
            sg.out_off[2 + k] = p->src_seg_base[k];
        }
        VO_CUDA(c, cudaStreamWaitEvent(c->stream5, c->ev_compact, 0));   // list 0 and its count
        VO_CUDA(c, cudaStreamWaitEvent(c->stream5, c->ev_join, 0));      // list 1 and its count (early classification, side stream)
        if (b->ready) VO_CUDA(c, cudaStreamWaitEvent(c->stream5, b->ready, 0));   // the frame, if it came over the ingest stream
        VO_TRY(launch_klt_segs(c, a->pyr, b->pyr, sg, p->lk_out[f].p, p->lk_status[f].p, p->lk_err[f].p, p->lk_iters[f].p, prm,
                               c->stream5, nullptr));
        VO_CUDA(c, cudaEventRecord(c->ev_trk, c->stream5));
    } else
    VO_TRY(launch_klt2(c, a->pyr, b->pyr, p->P[p->cur].p, &p->d_sc->n_land, nA, nullptr, nullptr, 0, 0, p->klt_out.p,
                       p->klt_status.p, p->klt_err.p, p->klt_iters.p, prm, nullptr, gate));
    if (prm.fb_check)
        VO_TRY(klt_fb_check(c, a->pyr, b->pyr, p->P[p->cur].p, &p->d_sc->n_land, nA, p->klt_out.p, p->klt_status.p, prm, c->stream, 0, gate));
    p->klt_b_off = nA;
    if (nB > 0) {
        if (!cand_ahead) VO_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev_fork, 0));
        if (n_add > 0) {
            const size_t need = (size_t)n_add * 2 + 16;
            if (need > p->h_add_cap) {
                VO_CUDA(c, cudaStreamSynchronize(c->stream2));
                size_t cap = 4096;
                while (cap < need) cap *= 2;
                for (int k = 0; k < 2; k++) {
                    if (p->h_add[k]) cudaFreeHost(p->h_add[k]);
                    p->h_add[k] = nullptr;
                    VO_CUDA(c, cudaMallocHost((void**)&p->h_add[k], cap * sizeof(double)));
                }
                p->h_add_cap = cap;
            }
            // two staging buffers in turn: the copy out of the one used two frames ago was joined by that
            // frame's promotion step, which the pose read-back of the frame after it has waited for
            const int f = p->add_flip ^= 1;
            memcpy(p->h_add[f], add->T44, 16 * sizeof(double));
            memcpy(p->h_add[f] + 16, add->pts, (size_t)n_add * 2 * sizeof(double));
            // the append kernel reads the rows straight out of the pinned buffer (unified addressing:
            // cudaMallocHost memory is device-accessible under the same pointer) - a few KB over PCIe on the
            // side stream, and one driver call less on the host
            VO_TRY(launch_cand_add(c, p->h_add[f] + 16, n_add, p->h_add[f], nullptr, true));
            c->side_writes_counts = true;
        }
        static const int cand_order = getenv("VO_CAND_AFTER_LAND") ? atoi(getenv("VO_CAND_AFTER_LAND")) : 0;
        if (cand_order && nA > 0) {
            if (!p->ev_land) VO_CUDA(c, cudaEventCreateWithFlags(&p->ev_land, cudaEventDisableTiming));
            VO_CUDA(c, cudaEventRecord(p->ev_land, c->stream));
            VO_CUDA(c, cudaStreamWaitEvent(c->stream2, p->ev_land, 0));
        }
        if (p->profiling) cudaEventRecord(p->ev_side[0], c->stream2);
        if (cand_ahead) {
            KltSegs sg{};
            sg.n = 2;
            sg.pts[0] = (const float2*)p->C[p->ccur].p; sg.n_dev[0] = &p->d_sc->n_cand_early; sg.n_upper[0] = nB - n_feed; sg.out_off[0] = nA;
            sg.pts[1] = (const float2*)d_feed; sg.n_dev[1] = nullptr; sg.n_upper[1] = n_feed; sg.out_off[1] = nA;
            sg.off_dev[1] = &p->d_sc->n_cand_early; sg.f64[1] = 1;
            if (n_feed > 0) VO_CUDA(c, cudaStreamWaitEvent(c->stream2, p->ev_feed, 0));
            if (b->ready) VO_CUDA(c, cudaStreamWaitEvent(c->stream2, b->ready, 0));
            VO_TRY(launch_klt_segs(c, a->pyr, b->pyr, sg, p->klt_out.p, p->klt_status.p, p->klt_err.p, p->klt_iters.p, prm,
                                   c->stream2, nullptr));
            VO_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev_fork, 0));   // the classification needs the tail's append
        } else
        VO_TRY(launch_klt2(c, a->pyr, b->pyr, nullptr, nullptr, 0, p->C[p->ccur].p, &p->d_sc->n_cand, nB, nA, p->klt_out.p,
                           p->klt_status.p, p->klt_err.p, p->klt_iters.p, prm, c->stream2, gate));
        if (prm.fb_check)
            VO_TRY(klt_fb_check(c, a->pyr, b->pyr, p->C[p->ccur].p, &p->d_sc->n_cand, nB, p->klt_out.p + 2 * (size_t)nA,
                                p->klt_status.p + nA, prm, c->stream2, 1, gate));
        if (p->profiling) cudaEventRecord(p->ev_side[1], c->stream2);
        // the pose-independent half of the promotion step follows the candidates' KLT on the side stream
        if (early_alpha >= 0.) {
            // The pose stream may still owe the previous pipelined frame its scalar read-back and the DLT of the rows it
            // parked for promotion (prom_rows / prom_ti / n_promoted), all of which this classification overwrites: wait
            // for the END of that work (ev_pose_done), not just for the read-back.  Usually long over - but when a
            // tracking launch larger than the GPU (landmarks + an appended map batch + candidates > 148 x 16 warps)
            // fills every SM, the pose stream's small kernels wait for a free slot and come AFTER the candidates' KLT.
            if (c->pose_inflight) VO_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev_pose_done, 0));
            // (the tracking stream's launch reads the parked rows and their count, which this classification rewrites)
            if (superset) VO_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev_trk, 0));
            VO_TRY(launch_promote_early(c, early_alpha, nB, nA, gate));
        }
        VO_CUDA(c, cudaEventRecord(c->ev_join, c->stream2));
        c->side_inflight = true;
    }
    return VO_OK;
}

}  // namespace vo

using namespace vo;

__global__ void expand_poses_kernel(const int32_t* __restrict__ Ti, const double* __restrict__ poses, int n,
                                    double* __restrict__ T)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * 16) return;
    T[i] = poses[16 * (size_t)Ti[i >> 4] + (i & 15)];
}

__global__ void pack_points_kernel(DevScalars* sc, const double* __restrict__ X, const float2* __restrict__ P, int n,
                                   int from_state, float4* __restrict__ q4, float* __restrict__ qv,
                                   int32_t* __restrict__ sets, vo::RngTab rt)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int m = from_state ? min(sc->n_land, n) : n;
    if (i == 0) sc->n_tracked = m;
    if (i < m) {
        const float2 p = P[i];
        q4[i] = make_float4((float)X[3 * (size_t)i], (float)X[3 * (size_t)i + 1], (float)X[3 * (size_t)i + 2], p.x);
        qv[i] = p.y;
    }
    // the point count of the coming PnP-RANSAC is known: draw its first minimal sets here
    if (blockIdx.x == 0 && threadIdx.x < 32) {
        vo::RngPreload pre;
        vo::pnp_preload_draws(rt, pre);
        vo::pnp_prepare_sets(sc, sets, rt, m, pre);
    }
}

// ---- landmark append (reference vo_pipeline.py:318-320, main.py:198-200) ---------------------------------
// Rows the kernel can read in place (device memory, pinned or registered host memory) are used where they
// are; pageable rows are copied into a slot of the pinned staging ring first, so the caller's arrays may
// go away as soon as the call returns.
static int land_rows_device_view(Ctx* c, const float* P, const double* X, int n, const float** dP, const double** dX)
{
    Pipeline* p = c->pipe;
    cudaPointerAttributes aP{}, aX{};
    const bool okP = cudaPointerGetAttributes(&aP, P) == cudaSuccess && aP.type != cudaMemoryTypeUnregistered;
    const bool okX = cudaPointerGetAttributes(&aX, X) == cudaSuccess && aX.type != cudaMemoryTypeUnregistered;
    cudaGetLastError();   // (older drivers report unregistered host pointers as an error)
    if (okP && okX) { *dP = P; *dX = X; return VO_OK; }
    const int slot = p->land_flip = (p->land_flip + 1) & 3;
    const size_t need = (size_t)n * 4;   // doubles: 3 per X row + 1 (= 2 floats) per P row
    if (p->ev_land_slot[slot]) VO_CUDA(c, cudaEventSynchronize(p->ev_land_slot[slot]));   // its last reader is done
    else VO_CUDA(c, cudaEventCreateWithFlags(&p->ev_land_slot[slot], cudaEventDisableTiming));
    if (need > p->h_land_cap[slot]) {
        size_t cap = 4096;
        while (cap < need) cap *= 2;
        if (p->h_land[slot]) cudaFreeHost(p->h_land[slot]);
        p->h_land[slot] = nullptr; p->h_land_cap[slot] = 0;
        VO_CUDA(c, cudaMallocHost((void**)&p->h_land[slot], cap * sizeof(double)));
        p->h_land_cap[slot] = cap;
    }
    double* hx = p->h_land[slot];
    float* hp = reinterpret_cast<float*>(hx + 3 * (size_t)n);
    memcpy(hx, X, sizeof(double) * 3 * n);
    memcpy(hp, P, sizeof(float) * 2 * n);
    *dP = hp; *dX = hx;
    p->land_slot_used = slot;
    return VO_OK;
}

static int append_landmarks_now(Ctx* c, const float* P, const double* X, int n, int target)
{
    Pipeline* p = c->pipe;
    const float* dP; const double* dX;
    p->land_slot_used = -1;
    VO_TRY(land_rows_device_view(c, P, X, n, &dP, &dX));
    VO_TRY(launch_land_append(c, dP, dX, n, target));
    if (p->land_slot_used >= 0) VO_CUDA(c, cudaEventRecord(p->ev_land_slot[p->land_slot_used], c->stream));
    return VO_OK;
}

#define VO_NEED(c) do { if (!(c) || !(c)->pipe) return VO_ERR_INVALID; } while (0)

extern "C" {

// ---- state I/O --------------------------------------------------------------------------------
int vo_track_rollback(vo_ctx* c);

// A tracking step that vo_process_frame_pipelined enqueued ahead belongs to the NEXT frame: whoever looks at
// the state or replaces it in between gets (and leaves) the state as the last frame left it.
static int drop_ahead(vo_ctx* c)
{
    Pipeline* p = c->pipe;
    if (p->ahead_valid && p->rb_valid && p->packed_valid) VO_TRY(vo_track_rollback(c));
    p->ahead_valid = false;
    return VO_OK;
}

int vo_set_state(vo_ctx* c, const float* P, const double* X, int n)
{
    VO_NEED(c);
    if (n < 0 || (n > 0 && (!P || !X))) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    VO_TRY(drop_ahead(c));
    VO_TRY(join_pose_stream(c));
    VO_CUDA(c, cudaStreamSynchronize(c->stream5));   // (a landmark KLT that ran ahead reads the rotation's buffers)
    p->b1_buf = -1; p->src_valid = false;
    const size_t m = n > 0 ? n : 1;
    VO_TRY(ensure(c, p->P[p->cur], m * 2)); VO_TRY(ensure(c, p->X[p->cur], m * 3));
    if (n > 0) {
        VO_CUDA(c, cudaMemcpyAsync(p->P[p->cur].p, P, sizeof(float) * 2 * n, cudaMemcpyDefault, c->stream));
        VO_CUDA(c, cudaMemcpyAsync(p->X[p->cur].p, X, sizeof(double) * 3 * n, cudaMemcpyDefault, c->stream));
    }
    set_counts_kernel<<<1, 1, 0, c->stream>>>(p->d_sc, n, -1, 0);
    c->launches++;
    p->n_land_upper = n;
    p->cand_tracked_valid = false;
    p->packed_valid = false;
    p->spec_pending = false;
    // the caller's buffers may be pageable and go away: finish the copies before returning
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

int vo_get_state(vo_ctx* c, float* P, double* X, int cap, int* n)
{
    VO_NEED(c);
    Pipeline* p = c->pipe;
    VO_TRY(drop_ahead(c));
    VO_TRY(read_scalars(c));   // (joins the pose stream: X of rows promoted by a pipelined frame)
    // (a tracking step the caller has taken over, with its PnP already on the device: the state between the two is the
    // rows the tracking step kept - the buffer the pruning read, not the one it wrote)
    const int buf = p->pnp_ahead ? p->pa_cur : p->cur;
    const int m = p->pnp_ahead ? p->h_sc->n_tracked : p->h_sc->n_land;
    if (n) *n = m;
    if (m > cap) return (P || X) ? set_error(c, VO_ERR_CAPACITY, "state has %d rows, buffer %d", m, cap) : VO_OK;
    if (m > 0 && P) VO_CUDA(c, cudaMemcpyAsync(P, p->P[buf].p, sizeof(float) * 2 * m, cudaMemcpyDefault, c->stream));
    if (m > 0 && X) VO_CUDA(c, cudaMemcpyAsync(X, p->X[buf].p, sizeof(double) * 3 * m, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}


// The landmarks as they were before the last vo_klt_track_landmarks / vo_process_frame: the third
// landmark buffer keeps them until the next tracking step (main.py:248 reads the previous frame's
// state['X'] after the current frame's associate + estimatePose).  Synchronises.
static int get_state_before(vo_ctx* c, int steps_back, float* P, double* X, int cap, int* n)
{
    VO_NEED(c);
    Pipeline* p = c->pipe;
    const int buf = steps_back == 2 ? p->prev2 : p->prev;
    if (buf < 0) return set_error(c, VO_ERR_STATE, "no state from %d tracking step(s) back is held", steps_back);
    VO_TRY(read_scalars(c));
    const int m = steps_back == 2 ? p->h_sc->prev2_n_land : p->h_sc->prev_n_land;
    if (n) *n = m;
    if (m > cap) return (P || X) ? set_error(c, VO_ERR_CAPACITY, "previous state has %d rows, buffer %d", m, cap) : VO_OK;
    if (m > 0 && P) VO_CUDA(c, cudaMemcpyAsync(P, p->P[buf].p, sizeof(float) * 2 * m, cudaMemcpyDefault, c->stream));
    if (m > 0 && X) VO_CUDA(c, cudaMemcpyAsync(X, p->X[buf].p, sizeof(double) * 3 * m, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

int vo_get_prev_state(vo_ctx* c, float* P, double* X, int cap, int* n) { return get_state_before(c, 1, P, X, cap, n); }
// ... and the state before the tracking step before that one: what main.py:248 reads when the tracking of the next
// frame has been enqueued one frame ahead (inside the pose call, see vo_process_frame_pipelined)
int vo_get_prev2_state(vo_ctx* c, float* P, double* X, int cap, int* n) { return get_state_before(c, 2, P, X, cap, n); }

// main.py:248 reads the PREVIOUS frame's state['X'] after the next frame's associate/estimatePose
// have already replaced it on the device: keep an asynchronous pinned-host snapshot (64 KB at 2k
// landmarks; the copy engine is idle at that point) that the stale dict can still be served from.
int vo_snapshot_state(vo_ctx* c)
{
    VO_NEED(c);
    Pipeline* p = c->pipe;
    VO_TRY(join_pose_stream(c));
    const size_t n = p->n_land_upper > 0 ? p->n_land_upper : 1;
    if (n > p->snap_cap) {
        VO_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFreeHost(p->h_snapP); cudaFreeHost(p->h_snapX);
        p->h_snapP = nullptr; p->h_snapX = nullptr; p->snap_cap = 0;
        size_t cap = 4096;
        while (cap < n) cap *= 2;
        VO_CUDA(c, cudaMallocHost((void**)&p->h_snapP, cap * 2 * sizeof(float)));
        VO_CUDA(c, cudaMallocHost((void**)&p->h_snapX, cap * 3 * sizeof(double)));
        p->snap_cap = cap;
    }
    if (!p->h_snap_sc) VO_CUDA(c, cudaMallocHost((void**)&p->h_snap_sc, sizeof(DevScalars)));
    VO_CUDA(c, cudaMemcpyAsync(p->h_snap_sc, p->d_sc, sizeof(DevScalars), cudaMemcpyDeviceToHost, c->stream));
    if (p->n_land_upper > 0) {
        VO_CUDA(c, cudaMemcpyAsync(p->h_snapP, p->P[p->cur].p, sizeof(float) * 2 * p->n_land_upper, cudaMemcpyDeviceToHost, c->stream));
        VO_CUDA(c, cudaMemcpyAsync(p->h_snapX, p->X[p->cur].p, sizeof(double) * 3 * p->n_land_upper, cudaMemcpyDeviceToHost, c->stream));
    }
    p->snap_rows = p->n_land_upper;
    return VO_OK;
}

int vo_get_snapshot(vo_ctx* c, float* P, double* X, int cap, int* n)
{
    VO_NEED(c);
    Pipeline* p = c->pipe;
    if (!p->h_snap_sc) return set_error(c, VO_ERR_STATE, "no snapshot taken");
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    int m = p->h_snap_sc->n_land;
    if (m > p->snap_rows) m = p->snap_rows;
    if (n) *n = m;
    if (m > cap) return (P || X) ? set_error(c, VO_ERR_CAPACITY, "snapshot has %d rows, buffer %d", m, cap) : VO_OK;
    if (P && m > 0) memcpy(P, p->h_snapP, sizeof(float) * 2 * m);
    if (X && m > 0) memcpy(X, p->h_snapX, sizeof(double) * 3 * m);
    return VO_OK;
}

// test / bootstrap hook: load a candidate set; T is n x 16 (4x4 row-major per candidate)
int vo_set_candidates(vo_ctx* c, const float* C, const double* F, const double* T, int n, int cf_alias)
{
    VO_NEED(c);
    VO_TRY(drop_ahead(c));
    VO_TRY(join_pose_stream(c));
    VO_TRY(early_undo(c));
    VO_TRY(join_side_stream(c));
    if (n < 0 || (n > 0 && (!C || !F || !T))) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    const size_t m = n > 0 ? n : 1;
    VO_TRY(ensure(c, p->C[p->ccur], m * 2)); VO_TRY(ensure(c, p->F[p->ccur], m * 2)); VO_TRY(ensure(c, p->Ti[p->ccur], m));
    VO_TRY(ensure(c, p->poses, m * 16));
    if (n > 0) {
        std::vector<int32_t> idx(n);
        for (int i = 0; i < n; i++) idx[i] = i;
        VO_CUDA(c, cudaMemcpyAsync(p->C[p->ccur].p, C, sizeof(float) * 2 * n, cudaMemcpyDefault, c->stream));
        VO_CUDA(c, cudaMemcpyAsync(p->F[p->ccur].p, F, sizeof(double) * 2 * n, cudaMemcpyDefault, c->stream));
        VO_CUDA(c, cudaMemcpyAsync(p->poses.p, T, sizeof(double) * 16 * n, cudaMemcpyDefault, c->stream));
        VO_CUDA(c, cudaMemcpyAsync(p->Ti[p->ccur].p, idx.data(), sizeof(int32_t) * n, cudaMemcpyDefault, c->stream));
        VO_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    p->n_poses = n;
    set_counts_kernel<<<1, 1, 0, c->stream>>>(p->d_sc, -1, n, cf_alias ? 1 : 0);
    c->launches++;
    p->n_cand_upper = n;
    p->cand_tracked_valid = false;
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

int vo_get_candidates(vo_ctx* c, float* C, double* F, double* T, int cap, int* n)
{
    VO_NEED(c);
    VO_TRY(drop_ahead(c));
    VO_TRY(early_undo(c));
    VO_TRY(join_side_stream(c));
    Pipeline* p = c->pipe;
    VO_TRY(read_scalars(c));
    const int m = p->h_sc->n_cand;
    if (n) *n = m;
    if (m > cap) return (C || F || T) ? set_error(c, VO_ERR_CAPACITY, "%d candidates, buffer %d", m, cap) : VO_OK;
    if (m > 0 && C) VO_CUDA(c, cudaMemcpyAsync(C, p->C[p->ccur].p, sizeof(float) * 2 * m, cudaMemcpyDefault, c->stream));
    if (m > 0 && F) VO_CUDA(c, cudaMemcpyAsync(F, p->F[p->ccur].p, sizeof(double) * 2 * m, cudaMemcpyDefault, c->stream));
    if (m > 0 && T) {
        VO_TRY(ensure(c, p->stage_d, (size_t)m * 16));
        expand_poses_kernel<<<(m * 16 + 255) / 256, 256, 0, c->stream>>>(p->Ti[p->ccur].p, p->poses.p, m, p->stage_d.p);
        c->launches++;
        VO_CUDA(c, cudaMemcpyAsync(T, p->stage_d.p, sizeof(double) * 16 * m, cudaMemcpyDefault, c->stream));
    }
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

int vo_get_lost(vo_ctx* c, float* lost, int cap, int* n)
{
    VO_NEED(c);
    Pipeline* p = c->pipe;
    VO_TRY(read_scalars(c));
    const int m = p->h_sc->n_lost;
    if (n) *n = m;
    if (m > cap) return lost ? set_error(c, VO_ERR_CAPACITY, "%d lost tracks, buffer %d", m, cap) : VO_OK;
    if (m > 0 && lost) VO_CUDA(c, cudaMemcpyAsync(lost, p->lost.p, sizeof(float) * 2 * m, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

static int fill_counters(vo_ctx* c, vo_counters* out, bool sync)
{
    VO_NEED(c);
    if (!out) return VO_ERR_INVALID;
    if (sync) VO_TRY(read_scalars(c));
    const DevScalars* s = c->pipe->h_sc;
    out->n_landmarks = c->pipe->pnp_ahead ? s->n_tracked : s->n_land;   // (a PnP enqueued ahead has not happened yet for the caller)
    out->n_candidates = s->n_cand; out->n_tracked = s->n_tracked; out->n_lost = s->n_lost;
    out->n_inliers = s->n_inliers; out->pnp_ok = s->ok; out->ransac_iters = s->iters_run; out->best_hypothesis = s->best_hyp;
    out->lm_iters = s->lm_iters; out->n_promoted = s->n_promoted; out->n_candidates_lost = s->n_cand_lost;
    out->cf_alias = s->cf_alias; out->klt_iters = s->klt_iters; out->klt_iters_candidates = s->klt_iters_cand;
    out->reproj_cost = s->cost; out->n_appended = s->n_appended; out->reserved_ = 0;
    return VO_OK;
}

// (counters are read as they are: with a tracking step enqueued ahead they already show its effect)
int vo_get_counters(vo_ctx* c, vo_counters* out) { return fill_counters(c, out, true); }
// diagnostics: %globaltimer stamps (ns) of the last RANSAC run, see DevScalars::stamps.  Synchronises.
int vo_debug_pnp_stamps(vo_ctx* c, uint64_t out[8])
{
    VO_NEED(c);
    if (!out) return VO_ERR_INVALID;
    VO_TRY(read_scalars(c));
    for (int k = 0; k < 8; k++) out[k] = c->pipe->h_sc->stamps[k];
    return VO_OK;
}
// values as of the last synchronising call (no device access)
int vo_peek_counters(vo_ctx* c, vo_counters* out) { return fill_counters(c, out, false); }

// ---- a1: KeypointsToLandmarksAssociator.associateKeypointsToLandmarks (vo_pipeline.py:188-227)
static int track_landmarks_impl(vo_ctx* c, int64_t prev_key, int64_t next_key, const vo_klt_params* params,
                                const DeferredAdd* add, int* n_kept, double early_alpha = -1.)
{
    VO_NEED(c);
    Pipeline* p = c->pipe;
    vo_klt_params prm;
    if (params) prm = *params; else vo_klt_params_default(&prm);
    FrameSlot* a = get_frame(c, prev_key);
    FrameSlot* b = get_frame(c, next_key);
    if (!a || !b) return set_error(c, VO_ERR_NOFRAME, "frame %lld not resident", (long long)(!a ? prev_key : next_key));
    p->spec_pending = false;   // a tracking step behind a speculative promotion confirms it (= vo_promote_commit)
    // candidates ride along: triangulate_landmark (vo:501) tracks them over the same frame pair, and
    // their tracking does not depend on the pose
    VO_TRY(track_both(c, a, b, prm, true, add, early_alpha));
    p->cand_tracked_prev = prev_key; p->cand_tracked_next = next_key; p->cand_tracked_valid = true;
    p->rb_cur = p->cur; p->rb_prev = p->prev; p->rb_prev_n_upper = p->prev_n_upper; p->rb_n_land_upper = p->n_land_upper;
    p->rb_prev2 = p->prev2; p->rb_prev2_n_upper = p->prev2_n_upper;
    VO_TRY(launch_landmark_compact(c, p->n_land_upper));
    p->rb_valid = true;
    p->packed_valid = true;
    if (n_kept) { VO_TRY(join_side_stream(c)); VO_TRY(read_scalars(c)); *n_kept = p->h_sc->n_land; }
    return VO_OK;
}

int vo_klt_track_landmarks(vo_ctx* c, int64_t prev_key, int64_t next_key, const vo_klt_params* params, int* n_kept)
{
    return track_landmarks_impl(c, prev_key, next_key, params, nullptr, n_kept);
}

// a1 with the previous frame's append of new candidates (vo_pipeline.py:574-591) folded in: the rows go
// to the side stream in front of the candidates' KLT, so neither the copy nor the append kernel sits
// between the host and the landmark launch.  Same result as vo_add_candidates + vo_klt_track_landmarks.
int vo_klt_track_landmarks_add(vo_ctx* c, int64_t prev_key, int64_t next_key, const vo_klt_params* params,
                               const double* new_candidates, int n_new, const double T44[16], int* n_kept)
{
    if (n_new < 0 || (n_new > 0 && (!new_candidates || !T44))) return VO_ERR_INVALID;
    DeferredAdd add{ new_candidates, n_new, T44 };
    return track_landmarks_impl(c, prev_key, next_key, params, n_new > 0 ? &add : nullptr, n_kept);
}

// A PnP that vo_process_frame_pipelined enqueued ahead (PnpLaunch::gate) is not the one the caller wants after all:
// the landmarks are again the rows its tracking step kept (the pruning wrote into another buffer of the rotation),
// and the device-side pose is the last frame's again.  Rare path: waits for the pose stream.
__global__ void pnp_unahead_kernel(DevScalars* sc, int ok)
{
    sc->n_land = sc->n_tracked; sc->ok = ok; sc->n_inliers = 0;
}

static int pnp_ahead_undo(vo_ctx* c)
{
    Pipeline* p = c->pipe;
    if (!p->pnp_ahead) return VO_OK;
    p->pnp_ahead = false;
    VO_CUDA(c, cudaStreamSynchronize(c->stream4));   // its refit
    pnp_unahead_kernel<<<1, 1, 0, c->stream>>>(p->d_sc, p->keep_ok);
    c->launches++;
    VO_CUDA(c, cudaGetLastError());
    VO_CUDA(c, cudaMemcpyAsync(p->d_sc->T, p->keep_T, sizeof(p->keep_T), cudaMemcpyHostToDevice, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));    // (keep_T is pageable)
    p->cur = p->pa_cur; p->prev = p->pa_prev; p->prev2 = p->pa_prev2;
    p->src_valid = false;
    return VO_OK;
}

__global__ void track_rollback_kernel(DevScalars* sc)
{
    sc->n_land = sc->prev_n_land; sc->prev_n_land = sc->prev2_n_land; sc->n_tracked = 0; sc->n_lost = 0;
}

// The same, and the pose-independent half of the coming promotion step (vo_pipeline.py:503-545: lost rows
// out, positions updated; :621-642: bearing test with threshold early_alpha_deg; compaction of C/F/T) goes
// to the side stream behind the candidates' KLT.  The promotion call that follows with the same alpha
// only triangulates and appends the promoted rows; with another alpha, or if something looks at the
// candidates first, the early half is taken back.  early_alpha_deg < 0: none.
int vo_klt_track_landmarks_ex(vo_ctx* c, int64_t prev_key, int64_t next_key, const vo_klt_params* params,
                              const double* new_candidates, int n_new, const double T44[16], double early_alpha_deg,
                              int* n_kept)
{
    if (n_new < 0 || (n_new > 0 && (!new_candidates || !T44))) return VO_ERR_INVALID;
    DeferredAdd add{ new_candidates, n_new, T44 };
    return track_landmarks_impl(c, prev_key, next_key, params, n_new > 0 ? &add : nullptr, n_kept, early_alpha_deg);
}

// Undo the last vo_klt_track_landmarks[_add] (nothing but the tracking step may have run since): the
// landmarks are again the ones it started from - the compaction wrote into another buffer of the
// rotation - and the candidates' pending KLT result is dropped.  A folded append stays (it is not part
// of the tracking).  The state before the PREVIOUS tracking step (vo_get_prev_state) is gone.
int vo_track_rollback(vo_ctx* c)
{
    VO_NEED(c);
    Pipeline* p = c->pipe;
    if (!p->rb_valid || !p->packed_valid) return set_error(c, VO_ERR_STATE, "no tracking step to undo");
    VO_TRY(pnp_ahead_undo(c));
    VO_TRY(early_undo(c));
    VO_TRY(join_side_stream(c));
    track_rollback_kernel<<<1, 1, 0, c->stream>>>(p->d_sc);
    c->launches++;
    VO_CUDA(c, cudaGetLastError());
    // (the rotation keeps the state before the tracking step before this one, too)
    p->cur = p->rb_cur; p->prev = p->rb_prev; p->prev_n_upper = p->rb_prev_n_upper; p->prev2 = -1; p->prev2_n_upper = 0;
    p->n_land_upper = p->rb_n_land_upper;
    p->rb_valid = false; p->packed_valid = false; p->cand_tracked_valid = false;
    p->b1_buf = -1; p->src_valid = false;
    return VO_OK;
}

// ---- a3: PoseEstimator.estimatePose (vo_pipeline.py:245-335); synchronises
int vo_pnp_ransac(vo_ctx* c, const double K[9], const vo_pnp_params* params, double T44[16], int* n_inliers, int* ok)
{
    VO_NEED(c);
    if (!K) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    VO_TRY(pnp_ahead_undo(c));
    PnpLaunch L{};
    if (params) L.prm = *params; else vo_pnp_params_default(&L.prm);
    if (!p->packed_valid) {   // no tracking step preceded (state set directly): pack P/X now
        const size_t m = p->n_land_upper > 0 ? p->n_land_upper : 1;
        VO_TRY(ensure(c, p->q4, m * 4)); VO_TRY(ensure(c, p->qv, m));
        VO_TRY(ensure_rng_table(c, p->hyp_capacity));
        pack_points_kernel<<<(int)((m + 255) / 256), 256, 0, c->stream>>>(p->d_sc, p->X[p->cur].p, (const float2*)p->P[p->cur].p,
                                                                          p->n_land_upper, 1, (float4*)p->q4.p, p->qv.p, p->sets.p,
                                                                          RngTab{ p->rng_tab.p, p->rng_tab_len });
        c->launches++;
    }
    L.q4 = p->q4.p; L.qv = p->qv.p; L.d_n = &p->d_sc->n_tracked; L.n_upper = p->n_land_upper;
    L.d_sets_in = nullptr; L.compact = true;
    memcpy(L.K, K, sizeof(L.K));
    VO_TRY(launch_pnp(c, L));
    p->packed_valid = false;
    VO_TRY(flush_frame_hint(c, true));   // the next frame's ingest: enqueued while the GPU runs the PnP
    VO_TRY(read_scalars(c));
    VO_TRY(check_rng(c));
    if (T44) memcpy(T44, p->h_sc->T, sizeof(double) * 16);
    if (n_inliers) *n_inliers = p->h_sc->n_inliers;
    if (ok) *ok = p->h_sc->ok;
    return VO_OK;
}

__global__ void promote_rollback_kernel(DevScalars* sc)
{
    sc->n_land = sc->spec_n_land; sc->n_cand = sc->spec_n_cand; sc->cf_alias = sc->spec_cf_alias;
    sc->cand_was_empty = sc->spec_cand_was_empty; sc->n_promoted = 0; sc->n_cand_lost = 0;
}

static int promote_rollback(Ctx* c)
{
    Pipeline* p = c->pipe;
    promote_rollback_kernel<<<1, 1, 0, c->stream>>>(p->d_sc);
    c->launches++;
    p->ccur ^= 1;                                  // the pre-promotion candidate buffers are intact
    p->mask_src_valid = false;
    p->n_land_upper = p->h_sc->spec_n_land;        // h_sc is current: every speculation ends in a sync
    p->n_cand_upper = p->h_sc->spec_n_cand;
    p->h_sc->n_land = p->h_sc->spec_n_land; p->h_sc->n_cand = p->h_sc->spec_n_cand;
    p->h_sc->n_promoted = 0; p->h_sc->n_cand_lost = 0;
    p->cand_tracked_prev = p->spec_prev_key; p->cand_tracked_next = p->spec_next_key;
    p->cand_tracked_valid = true;                  // the KLT results of the candidates are still there
    p->spec_pending = false;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

// a3 + a6 in one enqueue, ONE synchronisation: estimatePose, then - speculatively, before the pose
// is back on the host - the promotion step of triangulate_landmark with the pose taken from the
// device.  The class API forces a sync in estimatePose (it returns T); with the promotion already
// behind it, the GPU has no idle gap while the host walks from estimatePose to
// triangulate_landmark.  Needs the candidates tracked together with the landmarks
// (vo_klt_track_landmarks).  The caller then either confirms (vo_promote_commit: the arguments of
// triangulate_landmark were the ones assumed) or undoes it (vo_promote_rollback).  No model (ok=0):
// rolled back here, the state is as after vo_pnp_ransac.  The promoted rows sit behind the
// n_inliers rows of P/X.
int vo_pnp_ransac_promote(vo_ctx* c, const double K[9], const vo_pnp_params* params, double alpha_deg, double T44[16],
                          int* n_inliers, int* ok, int* n_promoted)
{
    VO_NEED(c);
    if (!K) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    if (!p->cand_tracked_valid || !p->packed_valid)
        return set_error(c, VO_ERR_STATE, "speculative promotion needs vo_klt_track_landmarks on this frame pair first");
    VO_TRY(pnp_ahead_undo(c));
    PnpLaunch L{};
    if (params) L.prm = *params; else vo_pnp_params_default(&L.prm);
    L.q4 = p->q4.p; L.qv = p->qv.p; L.d_n = &p->d_sc->n_tracked; L.n_upper = p->n_land_upper;
    L.d_sets_in = nullptr; L.compact = true;
    memcpy(L.K, K, sizeof(L.K));
    // n_promoted == NULL: the caller only needs the pose now; the host waits for the scalar block only, so the
    // promotion step (and whatever is enqueued next) runs while the host walks on to triangulate_landmark.
    const bool early = n_promoted == nullptr;
    const int ncand_before = p->n_cand_upper;
    const bool cls_done = p->early_valid && alpha_deg == p->early_alpha;   // candidate side already classified
    // ... and with the candidates classified by the tracking call (vo_klt_track_landmarks_ex) the landmark side is
    // final when the RANSAC kernel ends: refit, read-back and the promoted rows' DLT go to the pose stream
    const bool split = early && cls_done;
    L.refit_on_pose_stream = split;
    VO_TRY(launch_pnp(c, L));
    p->packed_valid = false;
    p->spec_prev_key = p->cand_tracked_prev; p->spec_next_key = p->cand_tracked_next;
    if (early && !p->ev_pose) VO_CUDA(c, cudaEventCreateWithFlags(&p->ev_pose, cudaEventDisableTiming));
    if (early && !split) {
        VO_TRY(join_side_stream(c));   // a deferred append on the side stream writes n_cand (long done by now)
        VO_CUDA(c, cudaMemcpyAsync(p->h_sc, p->d_sc, sizeof(DevScalars), cudaMemcpyDeviceToHost, c->stream));
        VO_CUDA(c, cudaEventRecord(p->ev_pose, c->stream));
    }
    VO_TRY(launch_promote(c, K, nullptr, alpha_deg, p->n_cand_upper, true, p->klt_b_off, split, -1));
    VO_TRY(close_step(c));
    p->cand_tracked_valid = false;
    p->spec_pending = true;
    VO_TRY(flush_frame_hint(c, true));   // the next frame's ingest: enqueued while the GPU runs the PnP
    if (early) {
        static const bool trace = getenv("VO_TRACE") != nullptr;
        if (trace) {   // host time blocked on the pose (tools/profile_e2e.py)
            timespec t0, t1;
            clock_gettime(CLOCK_MONOTONIC, &t0);
            const bool ready = cudaEventQuery(p->ev_pose) == cudaSuccess;
            cudaEventSynchronize(p->ev_pose);
            clock_gettime(CLOCK_MONOTONIC, &t1);
            p->trace_wait_us += (t1.tv_sec - t0.tv_sec) * 1e6 + (t1.tv_nsec - t0.tv_nsec) * 1e-3;
            p->trace_ready += ready; p->trace_n++;
            if (p->trace_n % 100 == 0)
                fprintf(stderr, "[vo trace] pose wait %.1f us/frame, pose already there %d/%d\n", p->trace_wait_us / p->trace_n,
                        p->trace_ready, p->trace_n);
        }
        VO_CUDA(c, cudaEventSynchronize(p->ev_pose));
        DevScalars* s = p->h_sc;   // as of the PnP: the promoted rows are not appended yet
        if (split) {   // (read back while the main stream went on: the counters come from the snapshot taken at the pose)
            s->n_land = s->snap_n_land; s->n_cand = s->snap_n_cand; s->cf_alias = s->snap_cf_alias; s->n_appended = s->snap_n_appended;
        }
        s->spec_n_land = s->n_land;   // what a rollback restores on the host side
        if (cls_done) {   // the side stream has classified the candidates: their counts are final already
            p->n_land_upper = s->n_land + s->n_promoted;
            p->n_cand_upper = s->n_cand;
        } else {          // values BEFORE the promotion step: bounds stay upper bounds
            p->n_land_upper = s->n_land + ncand_before;
            p->n_cand_upper = s->n_cand;
            s->spec_n_cand = s->n_cand;
            s->n_promoted = 0; s->n_cand_lost = 0;
        }
    } else {
        VO_TRY(read_scalars(c));
    }
    VO_TRY(check_rng(c));
    if (T44) memcpy(T44, p->h_sc->T, sizeof(double) * 16);
    if (n_inliers) *n_inliers = p->h_sc->n_inliers;
    if (ok) *ok = p->h_sc->ok;
    if (n_promoted) *n_promoted = p->h_sc->n_promoted;
    if (!p->h_sc->ok) VO_TRY(promote_rollback(c));
    return VO_OK;
}

int vo_promote_commit(vo_ctx* c)
{
    VO_NEED(c);
    if (!c->pipe->spec_pending) return set_error(c, VO_ERR_STATE, "no speculative promotion pending");
    c->pipe->spec_pending = false;
    return VO_OK;
}

int vo_promote_rollback(vo_ctx* c)
{
    VO_NEED(c);
    if (!c->pipe->spec_pending) return set_error(c, VO_ERR_STATE, "no speculative promotion pending");
    return promote_rollback(c);
}

int vo_promote_pending(vo_ctx* c) { return c && c->pipe && c->pipe->spec_pending ? 1 : 0; }

// The caller takes over a tracking step vo_process_frame_pipelined enqueued ahead (it is about to use it as the
// tracking step of the next frame): state queries no longer take it back.
int vo_ahead_adopt(vo_ctx* c)
{
    VO_NEED(c);
    c->pipe->ahead_valid = false;
    return VO_OK;
}

int vo_candidates_tracked(vo_ctx* c, int64_t prev_key, int64_t next_key)
{
    if (!c || !c->pipe) return 0;
    const Pipeline* p = c->pipe;
    return p->cand_tracked_valid && p->packed_valid && p->cand_tracked_prev == prev_key && p->cand_tracked_next == next_key;
}

// ---- externally supplied landmarks (vo_pipeline.py:318-320: np.vstack of the re-bootstrap's fresh landmarks
// onto associations['P'], ['X']; main.py:198-200) ---------------------------------------------------------------
int vo_append_landmarks(vo_ctx* c, const float* P, const double* X, int n, int target_total)
{
    VO_NEED(c);
    if (n < 0 || (n > 0 && (!P || !X))) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    if (p->ahead_valid && p->rb_valid && p->packed_valid) {   // a tracking step enqueued ahead did not see these rows
        VO_TRY(vo_track_rollback(c));
        p->ahead_valid = false;
    }
    if (p->packed_valid)
        return set_error(c, VO_ERR_STATE, "landmarks are appended between frames, not between a tracking step and its PnP");
    p->spec_pending = false;   // (an append behind a speculative promotion confirms it, as a tracking step does)
    if (n == 0) return VO_OK;
    return append_landmarks_now(c, P, X, n, target_total);
}

int vo_queue_landmarks(vo_ctx* c, const float* P, const double* X, int n, int target_total)
{
    VO_NEED(c);
    if (n < 0 || (n > 0 && (!P || !X))) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    if (n == 0) return VO_OK;
    if (p->n_land_q >= 4) return set_error(c, VO_ERR_CAPACITY, "at most 4 landmark batches can wait for a frame");
    p->land_q[p->n_land_q++] = Pipeline::LandSeg{ P, X, n, target_total };
    return VO_OK;
}

// ---- a6 (second half): append new candidates, F := C, T := pose (vo_pipeline.py:574-591)
int vo_add_candidates(vo_ctx* c, const double* pts, int n, const double T44[16])
{
    VO_NEED(c);
    VO_TRY(early_undo(c));
    if (n < 0 || (n > 0 && !pts) || !T44) return VO_ERR_INVALID;
    if (n == 0) return VO_OK;   // the reference appends nothing (vo:574)
    Pipeline* p = c->pipe;
    VO_TRY(ensure(c, p->stage_d, (size_t)n * 2 + 16));
    VO_CUDA(c, cudaMemcpyAsync(p->stage_d.p + 16, pts, sizeof(double) * 2 * n, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaMemcpyAsync(p->stage_d.p, T44, sizeof(double) * 16, cudaMemcpyDefault, c->stream));
    VO_TRY(launch_cand_add(c, p->stage_d.p + 16, n, p->stage_d.p));
    return VO_OK;
}

// ---- 8f-1: find_new_candidates_shi on the device (vo_pipeline.py:357-390 + :564-591) ---------------
// Shi-Tomasi corners of the resident frame, masked at the truncated PREVIOUS-frame positions of the
// candidates the last KLT kept (the reference's mask, :371-378), appended as candidates C := F :=
// corners, T := T44 (NULL: the pose the last PnP left on the device).  Nothing comes back to the
// host: no synchronisation.  Call after the promotion step of the same frame.
int vo_detect_add_candidates_shi(vo_ctx* c, int64_t frame_key, const vo_gftt_params* params, const double T44[16])
{
    VO_NEED(c);
    VO_TRY(join_pose_stream(c));
    VO_TRY(early_undo(c));
    Pipeline* p = c->pipe;
    vo_gftt_params prm;
    if (params) prm = *params; else vo_gftt_params_default(&prm);
    if (prm.max_corners <= 0) return set_error(c, VO_ERR_INVALID, "vo_detect_add_candidates_shi needs maxCorners > 0");
    FrameSlot* f = get_frame(c, frame_key);
    if (!f) return set_error(c, VO_ERR_NOFRAME, "frame %lld not resident", (long long)frame_key);
    VO_TRY(join_side_stream(c));
    const uint8_t* d_mask = nullptr;
    if (p->mask_src_valid) {
        // upper bound of the pre-promotion candidate count: what the promotion step was sized with
        const int n_upper = (int)(p->C[p->ccur ^ 1].cap / 2);
        VO_TRY(build_point_mask(c, f->w, f->h, p->C[p->ccur ^ 1].p, p->klt_status.p + p->mask_src_off,
                                &p->d_sc->spec_n_cand, n_upper));
        d_mask = c->det.mask.p;
    }
    VO_TRY(launch_gftt(c, *f, d_mask, prm));
    VO_TRY(ensure(c, p->stage_d, (size_t)prm.max_corners * 2 + 16));
    if (T44) VO_CUDA(c, cudaMemcpyAsync(p->stage_d.p, T44, sizeof(double) * 16, cudaMemcpyDefault, c->stream));
    VO_TRY(corners_to_stage(c, p->stage_d.p + 16));
    VO_TRY(launch_cand_add(c, p->stage_d.p + 16, prm.max_corners, T44 ? p->stage_d.p : nullptr, &c->det.d_sc->n_out));
    return VO_OK;
}

// Tracks the candidates C over (prev, next) unless vo_klt_track_landmarks already did for this
// frame pair; the result is consumed by vo_track_triangulate_promote.  (vo_pipeline.py:501)
int vo_track_candidates(vo_ctx* c, int64_t prev_key, int64_t next_key, const vo_klt_params* params)
{
    VO_NEED(c);
    VO_TRY(join_side_stream(c));
    Pipeline* p = c->pipe;
    if (p->cand_tracked_valid && p->cand_tracked_prev == prev_key && p->cand_tracked_next == next_key) return VO_OK;
    vo_klt_params prm;
    if (params) prm = *params; else vo_klt_params_default(&prm);
    FrameSlot* a = get_frame(c, prev_key);
    FrameSlot* b = get_frame(c, next_key);
    if (!a || !b) return set_error(c, VO_ERR_NOFRAME, "frame %lld not resident", (long long)(!a ? prev_key : next_key));
    const int nB = p->n_cand_upper;
    const size_t tot = nB > 0 ? nB : 1;
    VO_TRY(ensure(c, p->klt_out, tot * 2)); VO_TRY(ensure(c, p->klt_err, tot));
    VO_TRY(ensure(c, p->klt_status, tot)); VO_TRY(ensure(c, p->klt_iters, tot));
    VO_TRY(launch_klt(c, a->pyr, b->pyr, p->C[p->ccur].p, &p->d_sc->n_cand, nB, p->klt_out.p, p->klt_status.p,
                      p->klt_err.p, p->klt_iters.p, prm));
    if (prm.fb_check)
        VO_TRY(klt_fb_check(c, a->pyr, b->pyr, p->C[p->ccur].p, &p->d_sc->n_cand, nB, p->klt_out.p, p->klt_status.p, prm));
    p->klt_b_off = 0;
    p->cand_tracked_prev = prev_key; p->cand_tracked_next = next_key; p->cand_tracked_valid = true;
    return VO_OK;
}

// Host view of the pending candidate tracking (before vo_track_triangulate_promote consumes it):
// previous positions, tracked positions, status - what vo_pipeline.py:503-508 computes; the
// 'shi-mask' detector needs the well-tracked previous positions (vo_pipeline.py:568).  Synchronises.
int vo_get_candidate_tracks(vo_ctx* c, float* prev_pts, float* next_pts, uint8_t* status, int cap, int* n)
{
    VO_NEED(c);
    VO_TRY(early_undo(c));
    VO_TRY(join_side_stream(c));
    Pipeline* p = c->pipe;
    if (!p->cand_tracked_valid) return set_error(c, VO_ERR_STATE, "no pending candidate tracking");
    const int off = p->klt_b_off;
    VO_TRY(read_scalars(c));
    const int m = p->h_sc->n_cand;
    if (n) *n = m;
    if (m > cap) return (prev_pts || next_pts || status) ? set_error(c, VO_ERR_CAPACITY, "%d candidates, buffer %d", m, cap) : VO_OK;
    if (m > 0 && prev_pts) VO_CUDA(c, cudaMemcpyAsync(prev_pts, p->C[p->ccur].p, sizeof(float) * 2 * m, cudaMemcpyDefault, c->stream));
    if (m > 0 && next_pts) VO_CUDA(c, cudaMemcpyAsync(next_pts, p->klt_out.p + 2 * (size_t)off, sizeof(float) * 2 * m, cudaMemcpyDefault, c->stream));
    if (m > 0 && status) VO_CUDA(c, cudaMemcpyAsync(status, p->klt_status.p + off, m, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

// ---- a6 (first half): track C, drop lost, bearing test, DLT, promote (vo_pipeline.py:489-545, 608-721)
int vo_track_triangulate_promote(vo_ctx* c, int64_t prev_key, int64_t next_key, const double K[9], const double T44[16],
                                 double alpha_deg, const vo_klt_params* params, int* n_promoted)
{
    VO_NEED(c);
    if (!K || !T44) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    bool use_klt = true;
    if (prev_key == next_key && prev_key < 0) use_klt = false;   // promotion step only (tests)
    if (use_klt) VO_TRY(vo_track_candidates(c, prev_key, next_key, params));
    p->cand_tracked_valid = false;
    VO_TRY(launch_promote(c, K, T44, alpha_deg, p->n_cand_upper, use_klt, p->klt_b_off));
    VO_TRY(close_step(c));
    if (n_promoted) { VO_TRY(read_scalars(c)); *n_promoted = p->h_sc->n_promoted; }
    return VO_OK;
}

// ---- fused per-frame call: main.py:216-284 for one frame with ONE host sync ----------------------
// upload + pyramid of the new frame, KLT of P and C (one launch), landmark compaction, PnP-RANSAC +
// LM + pruning, candidate update + bearing test + DLT + promotion, append of the new candidates
// (pose taken from the device), D2H of the pose and counters.
//
// pipelined: the host waits for the pose only (scalars read back between the PnP and the promotion, so
// the counters are those BEFORE the promotion and the append), and if a frame has been handed over with
// vo_prefetch_frame since the last call, the tracking step into it is enqueued right behind the append
// - the next call for (next_key, that frame) finds it done, any other call takes it back - so the GPU
// runs the chain PnP -> promotion -> append -> KLT -> compaction without a gap across frames.
static void fill_counters_from(const DevScalars* s, vo_counters* counters)
{
    counters->n_landmarks = s->n_land; counters->n_candidates = s->n_cand; counters->n_tracked = s->n_tracked;
    counters->n_lost = s->n_lost; counters->n_inliers = s->n_inliers; counters->pnp_ok = s->ok;
    counters->ransac_iters = s->iters_run; counters->best_hypothesis = s->best_hyp; counters->lm_iters = s->lm_iters;
    counters->n_promoted = s->n_promoted; counters->n_candidates_lost = s->n_cand_lost; counters->cf_alias = s->cf_alias;
    counters->klt_iters = s->klt_iters; counters->klt_iters_candidates = s->klt_iters_cand; counters->reproj_cost = s->cost;
    counters->n_appended = s->n_appended; counters->reserved_ = 0;
}

static int process_frame_impl(vo_ctx* c, const uint8_t* img, int w, int h, int pitch, int64_t prev_key, int64_t next_key,
                              const double K[9], const vo_klt_params* kparams, const vo_pnp_params* pparams, double alpha_deg,
                              const double* new_candidates, int n_new, double T44[16], vo_counters* counters, bool pipelined)
{
    VO_NVTX(pipelined ? "vo:process_frame_pipelined" : "vo:process_frame");
    VO_NEED(c);
    if (!K || n_new < 0 || (n_new > 0 && !new_candidates)) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    vo_klt_params kp;
    if (kparams) kp = *kparams; else vo_klt_params_default(&kp);
    if (p->profiling) pipelined = false;   // the stage timers bracket the stages of ONE frame
    if (p->profiling) cudaEventRecord(p->ev[0], c->stream);
    // a tracking step enqueued ahead by the previous pipelined call: ours, or to be taken back
    // ... or one the caller made itself (vo_klt_track_landmarks* for this frame pair, the class surface's associate call)
    const bool tracked = p->packed_valid && p->cand_tracked_valid && p->cand_tracked_prev == prev_key &&
                         p->cand_tracked_next == next_key;
    if (p->ahead_valid && !tracked && p->rb_valid && p->packed_valid) VO_TRY(vo_track_rollback(c));
    p->ahead_valid = false;
    if (img) VO_TRY(vo_upload_frame(c, img, w, h, pitch, next_key));
    FrameSlot *a = nullptr, *b = nullptr;
    if (!tracked) {
        a = get_frame(c, prev_key);
        b = get_frame(c, next_key);
        if (!a || !b) return set_error(c, VO_ERR_NOFRAME, "frame %lld not resident", (long long)(!a ? prev_key : next_key));
    }
    // the frame's new candidates: two staging buffers in turn (the candidates' KLT of the NEXT frame, which a pipelined
    // call may enqueue ahead, reads the rows straight from here - and may still be running when the next feed arrives)
    double* feed_ptr = nullptr;
    if (n_new > 0) {
        DevBuf<double>& fbuf = p->feed_d[p->feed_flip ^= 1];
        VO_TRY(ensure(c, fbuf, (size_t)n_new * 2));
        feed_ptr = fbuf.p;
        if (!p->ev_feed) VO_CUDA(c, cudaEventCreateWithFlags(&p->ev_feed, cudaEventDisableTiming));
    }
    if (n_new > 0 && !tracked) {   // stage the candidate feed early so the copy overlaps the tracking
        VO_CUDA(c, cudaMemcpyAsync(feed_ptr, new_candidates, sizeof(double) * 2 * n_new, cudaMemcpyDefault, c->stream));
        VO_CUDA(c, cudaEventRecord(p->ev_feed, c->stream));
    }
    const bool prof = p->profiling;
    if (prof) cudaEventRecord(p->ev[1], c->stream);
    // Tracking ahead of the pruning (track_both, superset): the next frame's KLT launches run beside this frame's
    // RANSAC kernel and refit instead of behind them.  Same results.  OFF by default (VO_TRACK_AHEAD=1 turns it on):
    // measured on the B200 it shortens the main stream's chain (device-resident leg +4 %) but delays the pose the host
    // is waiting for and costs about ten more driver calls per frame (class-surface leg -9 %).
    static const int ahead_mode = getenv("VO_TRACK_AHEAD") ? atoi(getenv("VO_TRACK_AHEAD")) : 0;
    const bool want_superset = pipelined && ahead_mode != 0;
    if (!tracked) VO_TRY(track_both(c, a, b, kp, true, nullptr, alpha_deg));
    if (prof) cudaEventRecord(p->ev[2], c->stream);
    if (!tracked) VO_TRY(launch_landmark_compact(c, p->n_land_upper));
    p->packed_valid = true;
    PnpLaunch L{};
    if (pparams) L.prm = *pparams; else vo_pnp_params_default(&L.prm);
    L.q4 = p->q4.p; L.qv = p->qv.p; L.d_n = &p->d_sc->n_tracked; L.n_upper = p->n_land_upper;
    L.d_sets_in = nullptr; L.compact = true;
    memcpy(L.K, K, sizeof(L.K));
    if (prof) cudaEventRecord(p->ev[3], c->stream);
    // Pipelined frames with the candidates already classified (the steady state): the landmark side of the frame is
    // final when the RANSAC kernel ends (pruned P/X, promoted rows' positions, appended batches), so the main stream goes
    // straight on to the next frame's tracking, while the pose stream runs the LM refit, hands the pose to the host and
    // triangulates the promoted rows - needed again only by the next frame's compaction, which waits for it.
    const bool split = pipelined && p->early_valid && alpha_deg == p->early_alpha;
    L.refit_on_pose_stream = split;
    L.small_grid = split && want_superset;
    if (pipelined && !p->ev_pose) VO_CUDA(c, cudaEventCreateWithFlags(&p->ev_pose, cudaEventDisableTiming));
    // this frame's PnP may be on the device already (enqueued by the previous call behind the tracking step): it stands if
    // it was launched for the same camera matrix and parameters and this call splits the frame the same way
    bool pnp_done = false;
    if (p->pnp_ahead) {
        pnp_done = tracked && split && memcmp(p->pa_K, L.K, sizeof(L.K)) == 0 && memcmp(&p->pa_prm, &L.prm, sizeof(L.prm)) == 0;
        if (pnp_done) p->pnp_ahead = false;
        else VO_TRY(pnp_ahead_undo(c));
    }
    if (!pnp_done) VO_TRY(launch_pnp(c, L));
    if (n_new > 0 && tracked) {   // (tracking done ahead: the PnP launch is what the GPU is waiting for - the feed goes behind it)
        VO_CUDA(c, cudaMemcpyAsync(feed_ptr, new_candidates, sizeof(double) * 2 * n_new, cudaMemcpyDefault, c->stream));
        VO_CUDA(c, cudaEventRecord(p->ev_feed, c->stream));
    }
    if (prof) cudaEventRecord(p->ev[4], c->stream);
    const int ncand_before = p->n_cand_upper;
    const bool early_used = p->early_valid;   // the promotion below only finishes an early classification
    if (pipelined && !split) {
        VO_TRY(join_side_stream(c));
        VO_CUDA(c, cudaMemcpyAsync(p->h_sc, p->d_sc, sizeof(DevScalars), cudaMemcpyDeviceToHost, c->stream));
        VO_CUDA(c, cudaEventRecord(p->ev_pose, c->stream));
    }
    // landmark batches waiting for this frame (vo_queue_landmarks): behind the promoted rows, in call order
    Pipeline::LandSeg segs[4];
    const int n_segs = p->n_land_q;
    for (int k = 0; k < n_segs; k++) segs[k] = p->land_q[k];
    p->n_land_q = 0;
    if (split) {
        // one launch for the promoted rows' positions, the new candidates and the batches
        Pipeline::LandSeg dsegs[4];
        int slots[4];
        for (int k = 0; k < n_segs; k++) {
            const float* dP; const double* dX;
            p->land_slot_used = -1;
            VO_TRY(land_rows_device_view(c, segs[k].P, segs[k].X, segs[k].n, &dP, &dX));
            slots[k] = p->land_slot_used;
            dsegs[k] = Pipeline::LandSeg{ dP, dX, segs[k].n, segs[k].target };
        }
        VO_TRY(launch_frame_tail(c, K, alpha_deg, p->klt_b_off, feed_ptr, n_new, dsegs, n_segs));
        for (int k = 0; k < n_segs; k++) if (slots[k] >= 0) VO_CUDA(c, cudaEventRecord(p->ev_land_slot[slots[k]], c->stream));
        VO_TRY(close_step(c));
    } else {
        if (n_new > 0) VO_TRY(ensure(c, p->poses, (size_t)(p->n_poses + 1) * 16, true));   // (before a kernel gets the pointer)
        VO_TRY(launch_promote(c, K, nullptr, alpha_deg, p->n_cand_upper, true, p->klt_b_off, false, -1));
        VO_TRY(close_step(c));
        if (n_new > 0) VO_TRY(launch_cand_add(c, feed_ptr, n_new, nullptr));
        for (int k = 0; k < n_segs; k++) VO_TRY(append_landmarks_now(c, segs[k].P, segs[k].X, segs[k].n, segs[k].target));
    }
    if (prof) cudaEventRecord(p->ev[5], c->stream);
    p->cand_tracked_valid = false;
    p->packed_valid = false;
    p->rb_valid = false;
    const int b_off_k = p->klt_b_off, mask_off_k = p->mask_src_off;   // this frame's, should the step ahead be taken back
    const bool mask_valid_k = p->mask_src_valid;
    if (pipelined) {
        VO_TRY(flush_frame_hint(c, true));   // a frame announced for later goes to the ingest stream now: it may be the next one
        // the next frame is on the ingest stream already: its tracking goes right behind the append
        if (c->last_ingest_valid && c->last_ingest_key != next_key && vo_has_frame(c, c->last_ingest_key)) {
            const int64_t nk = c->last_ingest_key;
            FrameSlot* a2 = get_frame(c, next_key);
            FrameSlot* b2 = get_frame(c, nk);
            if (a2 && b2) {
                // every kernel of the step ahead looks at this frame's `ok` first and leaves all buffers alone
                // when no pose came: the failure path below then finds this frame's tracking results, candidate
                // buffers and scalars exactly as they were
                const int32_t* gate = split ? &p->d_sc->gate_ok : &p->d_sc->ok;
                const bool superset = split && p->src_valid && p->b1_buf >= 0 && !kp.fb_check && want_superset;
                VO_TRY(track_both(c, a2, b2, kp, true, nullptr, alpha_deg, gate, superset, feed_ptr, n_new));
                p->cand_tracked_prev = next_key; p->cand_tracked_next = nk; p->cand_tracked_valid = true;
                p->rb_cur = p->cur; p->rb_prev = p->prev; p->rb_prev_n_upper = p->prev_n_upper; p->rb_n_land_upper = p->n_land_upper;
    p->rb_prev2 = p->prev2; p->rb_prev2_n_upper = p->prev2_n_upper;
                // Its compaction and its PnP: the RANSAC kernel needs nothing the next call brings (new candidates and
                // landmark batches come in behind it), so the GPU goes from one frame to the next without waiting for
                // the host (VO_PNP_AHEAD=0: A/B switch, same results) - and then the kernel does the compaction itself,
                // as its first phase (VO_FUSE_COMPACT=0: A/B switch, same results), unless the tracks came from the
                // superset launch or there are more rows than its scan handles.
                static const int pnp_ahead_mode = getenv("VO_PNP_AHEAD") ? atoi(getenv("VO_PNP_AHEAD")) : 1;
                static const int fuse_mode = getenv("VO_FUSE_COMPACT") ? atoi(getenv("VO_FUSE_COMPACT")) : 1;
                const bool pnp_ahead = split && pnp_ahead_mode != 0;
                const bool fuse = pnp_ahead && fuse_mode != 0 && !superset && !kp.fb_check && p->n_land_upper <= kFuseMaxRows &&
                                  p->n_land_upper > 0;
                if (!fuse) VO_TRY(launch_landmark_compact(c, p->n_land_upper, gate, superset));
                p->rb_valid = true; p->packed_valid = true;
                p->ahead_valid = true; p->ahead_prev = next_key; p->ahead_next = nk;
                if (pnp_ahead) {
                    PnpLaunch L2 = L;
                    L2.q4 = p->q4.p; L2.qv = p->qv.p;   // (the compaction above may have grown them)
                    L2.n_upper = p->n_land_upper; L2.gate = gate; L2.refit_on_pose_stream = true; L2.small_grid = superset;
                    L2.fuse_compact = fuse;
                    if (fuse) {   // (what an undo of the PnP alone goes back to: the compacted buffer the kernel is about to write)
                        p->pa_cur = (p->cur + 1) % kLandBufs; p->pa_prev = p->cur; p->pa_prev2 = p->prev == p->pa_cur ? -1 : p->prev;
                    } else {
                        p->pa_cur = p->cur; p->pa_prev = p->prev; p->pa_prev2 = p->prev2;
                    }
                    memcpy(p->pa_K, L2.K, sizeof(p->pa_K)); p->pa_prm = L2.prm;
                    VO_TRY(launch_pnp(c, L2));
                    p->pnp_ahead = true;
                }
            }
        }
        VO_TRY(flush_frame_hint(c, true));
        {
            static const bool trace = getenv("VO_TRACE") != nullptr;
            if (trace) {   // host time blocked on the pose, and how often it was there already (tools/profile_e2e.py)
                timespec t0, t1;
                clock_gettime(CLOCK_MONOTONIC, &t0);
                const bool ready = cudaEventQuery(p->ev_pose) == cudaSuccess;
                cudaEventSynchronize(p->ev_pose);
                clock_gettime(CLOCK_MONOTONIC, &t1);
                p->trace_wait_us += (t1.tv_sec - t0.tv_sec) * 1e6 + (t1.tv_nsec - t0.tv_nsec) * 1e-3;
                p->trace_ready += ready; p->trace_n++;
                if (p->trace_n % 300 == 0) {
                    fprintf(stderr, "[vo trace] pipelined frames: pose wait %.1f us/frame, pose already there %d/%d\n",
                            p->trace_wait_us / p->trace_n, p->trace_ready, p->trace_n);
                    p->trace_wait_us = 0.; p->trace_ready = 0; p->trace_n = 0;
                }
            }
        }
        VO_CUDA(c, cudaEventSynchronize(p->ev_pose));
        DevScalars* s = p->h_sc;   // as of the PnP: before the append of the promoted rows and of the new candidates
        if (split) {   // (read back while the main stream went on: the counters come from the snapshot taken at the pose)
            s->n_land = s->snap_n_land; s->n_cand = s->snap_n_cand; s->cf_alias = s->snap_cf_alias; s->n_appended = s->snap_n_appended;
        }
        // (still upper bounds for what the promotion, the append and a tracking step enqueued above leave)
        if (early_used) { p->n_land_upper = s->n_land + s->n_promoted; p->n_cand_upper = s->n_cand + n_new; }
        else {
            p->n_land_upper = s->n_land + ncand_before; p->n_cand_upper = s->n_cand + n_new;
            s->n_promoted = 0; s->n_cand_lost = 0;
        }
        for (int k = 0; k < n_segs; k++) {   // the appended batches: same rule as the kernel, on the bounds
            int up = p->n_land_upper + segs[k].n;
            if (segs[k].target > 0 && up > segs[k].target) up = segs[k].target > p->n_land_upper ? segs[k].target : p->n_land_upper;
            p->n_land_upper = up;
        }
        if (!s->ok && p->ahead_valid) {
            // the step ahead did nothing on the device (gated): take back its host-side bookkeeping only
            VO_TRY(join_side_stream(c));
            p->cur = p->rb_cur; p->prev = p->rb_prev; p->prev_n_upper = p->rb_prev_n_upper; p->n_land_upper = p->rb_n_land_upper;
            p->prev2 = p->rb_prev2; p->prev2_n_upper = p->rb_prev2_n_upper;
            if (p->early_valid) {
                p->early_valid = false; p->ccur ^= 1;
                if (p->n_cand_upper < p->early_ncand_upper) p->n_cand_upper = p->early_ncand_upper;
            }
            p->klt_b_off = b_off_k; p->mask_src_valid = mask_valid_k; p->mask_src_off = mask_off_k;
            if (p->klt_swapped) {   // this frame's candidate tracks are in the other set
                VO_CUDA(c, cudaStreamSynchronize(c->stream2));
                std::swap(p->klt_out, p->klt_out2); std::swap(p->klt_err, p->klt_err2);
                std::swap(p->klt_status, p->klt_status2); std::swap(p->klt_iters, p->klt_iters2);
                p->klt_swapped = false;
            }
            p->b1_buf = -1; p->src_valid = false;
            p->rb_valid = false; p->packed_valid = false; p->cand_tracked_valid = false; p->ahead_valid = false;
            if (p->pnp_ahead) VO_CUDA(c, cudaStreamSynchronize(c->stream4));
            p->pnp_ahead = false;   // (gated: it did nothing; its refit found ok == 0 and wrote "no pose", as this frame's did)
        }
    } else {
        VO_TRY(read_scalars(c));
    }
    if (!p->h_sc->ok && early_used) {
        // no pose, but the early classification assumed one and took the would-be promoted rows out of the
        // candidates: back to the tracked candidates, then the step as it goes without a pose (lost rows out,
        // positions updated, nothing promoted)
        VO_TRY(read_scalars(c));
        VO_TRY(promote_rollback(c));
        VO_TRY(launch_promote(c, K, nullptr, alpha_deg, p->n_cand_upper, true, p->klt_b_off));
        p->cand_tracked_valid = false;
        // (the rollback also dropped the batches appended behind the would-be promoted rows: once more)
        for (int k = 0; k < n_segs; k++) VO_TRY(append_landmarks_now(c, segs[k].P, segs[k].X, segs[k].n, segs[k].target));
        VO_TRY(read_scalars(c));
    }
    VO_TRY(check_rng(c));
    if (prof) {   // accumulate stage times (ms): upload+pyramid, KLT, compaction, PnP, promotion
        for (int k = 0; k < 5; k++) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, p->ev[k], p->ev[k + 1]) == cudaSuccess) p->stage_ms[k] += ms;
        }
        float ms = 0.f;   // the candidates' KLT on the side stream (overlaps stages 1-3)
        if (cudaEventElapsedTime(&ms, p->ev_side[0], p->ev_side[1]) == cudaSuccess) p->stage_ms[5] += ms;
        p->stage_frames++;
    }
    if (pipelined) { memcpy(p->keep_T, p->h_sc->T, sizeof(p->keep_T)); p->keep_ok = p->h_sc->ok; }
    if (T44) memcpy(T44, p->h_sc->T, sizeof(double) * 16);
    if (counters) {
        fill_counters_from(p->h_sc, counters);
        counters->reserved_ = p->ahead_valid ? 1 : 0;   // bit 0: the next frame's tracking step has been enqueued ahead
    }
    return VO_OK;
}

int vo_process_frame(vo_ctx* c, const uint8_t* img, int w, int h, int pitch, int64_t prev_key, int64_t next_key,
                     const double K[9], const vo_klt_params* kparams, const vo_pnp_params* pparams, double alpha_deg,
                     const double* new_candidates, int n_new, double T44[16], vo_counters* counters)
{
    return process_frame_impl(c, img, w, h, pitch, prev_key, next_key, K, kparams, pparams, alpha_deg, new_candidates, n_new,
                              T44, counters, false);
}

int vo_process_frame_pipelined(vo_ctx* c, const uint8_t* img, int w, int h, int pitch, int64_t prev_key, int64_t next_key,
                               const double K[9], const vo_klt_params* kparams, const vo_pnp_params* pparams, double alpha_deg,
                               const double* new_candidates, int n_new, double T44[16], vo_counters* counters)
{
    return process_frame_impl(c, img, w, h, pitch, prev_key, next_key, K, kparams, pparams, alpha_deg, new_candidates, n_new,
                              T44, counters, true);
}

// Stage timers of vo_process_frame (CUDA events on the ctx stream).  enable: 1 = on (resets the
// accumulators), 0 = off.  vo_get_stage_times: ms summed over `frames` frames for the 5 stages
// {upload+pyramid, KLT, landmark compaction, PnP-RANSAC+LM, promotion+append}.
int vo_set_profiling(vo_ctx* c, int enable)
{
    VO_NEED(c);
    Pipeline* p = c->pipe;
    if (enable && !p->ev[0]) {
        for (int k = 0; k < 6; k++) VO_CUDA(c, cudaEventCreate(&p->ev[k]));
        for (int k = 0; k < 2; k++) VO_CUDA(c, cudaEventCreate(&p->ev_side[k]));
    }
    p->profiling = enable != 0;
    for (int k = 0; k < 6; k++) p->stage_ms[k] = 0.;
    p->stage_frames = 0;
    return VO_OK;
}

int vo_get_stage_times(vo_ctx* c, double ms[6], int* frames)
{
    VO_NEED(c);
    if (!ms) return VO_ERR_INVALID;
    for (int k = 0; k < 6; k++) ms[k] = c->pipe->stage_ms[k];
    if (frames) *frames = c->pipe->stage_frames;
    return VO_OK;
}

// ---- unit level: cv2.solvePnPRansac (vo_pipeline.py:258-264) ------------------------------------

static void rotation_to_rvec(const double* T, double* rvec)
{
    const double R[9] = { T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10] };
    double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
    const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
    double cs = (R[0] + R[4] + R[8] - 1.) * 0.5;
    cs = cs > 1. ? 1. : (cs < -1. ? -1. : cs);
    const double theta = acos(cs);
    if (s < 1e-5) {
        if (cs > 0.) { rvec[0] = rvec[1] = rvec[2] = 0.; return; }
        double t;
        t = (R[0] + 1.) * 0.5; rx = sqrt(t > 0. ? t : 0.);
        t = (R[4] + 1.) * 0.5; ry = sqrt(t > 0. ? t : 0.) * (R[1] < 0. ? -1. : 1.);
        t = (R[8] + 1.) * 0.5; rz = sqrt(t > 0. ? t : 0.) * (R[2] < 0. ? -1. : 1.);
        if (fabs(rx) < fabs(ry) && fabs(rx) < fabs(rz) && (R[5] > 0.) != (ry * rz > 0.)) rz = -rz;
        const double nn = theta / sqrt(rx * rx + ry * ry + rz * rz);
        rvec[0] = rx * nn; rvec[1] = ry * nn; rvec[2] = rz * nn;
    } else {
        const double vth = 0.5 / s * theta;
        rvec[0] = rx * vth; rvec[1] = ry * vth; rvec[2] = rz * vth;
    }
}

// X: n x 3 float64, P: n x 2 float32 (host or device).  minimal_sets: H x 4 int32 (H >= max_iters)
// or NULL = cv::RNG(2^64-1) emulation.  Outputs (all optional): T44 refined pose, Th12 the winning
// unrefined hypothesis [R|t], rvec/tvec as cv2 returns them, inlier_mask n bytes, info[4] =
// {iterations run, winning iteration, inlier count, LM iterations}.  Does not touch the state.
int vo_solve_pnp_ransac(vo_ctx* c, const double* X, const float* P, int n, const double K[9],
                        const vo_pnp_params* params, const int32_t* minimal_sets, double T44[16], double Th12[12],
                        double rvec[3], double tvec[3], uint8_t* inlier_mask, int32_t info[4], int* ok)
{
    VO_NEED(c);
    if (n < 0 || !K || (n > 0 && (!X || !P))) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    VO_TRY(drop_ahead(c));   // (a step enqueued ahead keeps its point count in the scalar block this run uses)
    PnpLaunch L{};
    if (params) L.prm = *params; else vo_pnp_params_default(&L.prm);
    const size_t m = n > 0 ? n : 1;
    VO_TRY(ensure(c, p->stage_d, m * 3)); VO_TRY(ensure(c, p->stage_f, m * 2));
    VO_TRY(ensure(c, p->q4, m * 4)); VO_TRY(ensure(c, p->qv, m));
    if (n > 0) {
        VO_CUDA(c, cudaMemcpyAsync(p->stage_d.p, X, sizeof(double) * 3 * n, cudaMemcpyDefault, c->stream));
        VO_CUDA(c, cudaMemcpyAsync(p->stage_f.p, P, sizeof(float) * 2 * n, cudaMemcpyDefault, c->stream));
    }
    VO_TRY(ensure_rng_table(c, p->hyp_capacity));
    pack_points_kernel<<<(int)((m + 255) / 256), 256, 0, c->stream>>>(p->d_sc, p->stage_d.p, (const float2*)p->stage_f.p, n, 0,
                                                                      (float4*)p->q4.p, p->qv.p, p->sets.p,
                                                                      RngTab{ p->rng_tab.p, p->rng_tab_len });
    c->launches++;
    if (minimal_sets) {
        const size_t H = L.prm.max_iters > 0 ? L.prm.max_iters : 1;
        VO_TRY(ensure(c, p->stage_i, H * kMinimal));
        VO_CUDA(c, cudaMemcpyAsync(p->stage_i.p, minimal_sets, sizeof(int32_t) * kMinimal * H, cudaMemcpyDefault, c->stream));
        L.d_sets_in = p->stage_i.p;
    }
    L.q4 = p->q4.p; L.qv = p->qv.p; L.d_n = nullptr; L.n_upper = n; L.compact = false;
    memcpy(L.K, K, sizeof(L.K));
    p->packed_valid = false;
    VO_TRY(launch_pnp(c, L));
    if (inlier_mask && n > 0) VO_CUDA(c, cudaMemcpyAsync(inlier_mask, p->mask.p, n, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaMemcpyAsync(p->h_sc, p->d_sc, sizeof(DevScalars), cudaMemcpyDeviceToHost, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    VO_TRY(check_rng(c));
    const DevScalars* s = p->h_sc;
    if (ok) *ok = s->ok;
    if (T44) memcpy(T44, s->T, sizeof(double) * 16);
    if (Th12) memcpy(Th12, s->Th, sizeof(double) * 12);
    if (rvec) { if (s->ok) rotation_to_rvec(s->T, rvec); else rvec[0] = rvec[1] = rvec[2] = 0.; }
    if (tvec) { tvec[0] = s->T[3]; tvec[1] = s->T[7]; tvec[2] = s->T[11]; }
    if (info) { info[0] = s->iters_run; info[1] = s->best_hyp; info[2] = s->n_inliers; info[3] = s->lm_iters; }
    return VO_OK;
}

// ---- unit level: cv2.triangulatePoints (vo_pipeline.py:666, :172) --------------------------------
// M1: 3x4 row-major (per_point_m1 = 0) or n x 12 (one first-view matrix per point, as the
// reference's loop uses K @ T[idx]); M2: 3x4; p1, p2: n x 2 float64; X4: n x 4 homogeneous, unit
// norm, sign arbitrary (as OpenCV).  Host or device pointers.  Synchronises.
int vo_triangulate_points(vo_ctx* c, const double* M1, const double* M2, int per_point_m1, const double* p1,
                          const double* p2, int n, double* X4)
{
    VO_NEED(c);
    if (n < 0 || (n > 0 && (!M1 || !M2 || !p1 || !p2 || !X4))) return VO_ERR_INVALID;
    if (n == 0) return VO_OK;
    Pipeline* p = c->pipe;
    const size_t nm1 = per_point_m1 ? (size_t)n * 12 : 12;
    VO_TRY(ensure(c, p->stage_d, nm1 + 12 + (size_t)n * 8));
    double* dM1 = p->stage_d.p; double* dM2 = dM1 + nm1; double* dp1 = dM2 + 12; double* dp2 = dp1 + 2 * (size_t)n;
    double* dX = dp2 + 2 * (size_t)n;
    VO_CUDA(c, cudaMemcpyAsync(dM1, M1, sizeof(double) * nm1, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaMemcpyAsync(dM2, M2, sizeof(double) * 12, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaMemcpyAsync(dp1, p1, sizeof(double) * 2 * n, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaMemcpyAsync(dp2, p2, sizeof(double) * 2 * n, cudaMemcpyDefault, c->stream));
    VO_TRY(launch_triangulate(c, dM1, dM2, per_point_m1, dp1, dp2, n, dX));
    VO_CUDA(c, cudaMemcpyAsync(X4, dX, sizeof(double) * 4 * n, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

}  // extern "C"

// ---- vo_solve_pnp_ransac without a single allocation -------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and
// same results as vo_solve_pnp_ransac; every device buffer the run needs - the staged points, the float32 pack,
// minimal sets, hypotheses, counts, inlier mask, inlier pack, the cv::RNG draw table - lives in ONE file-scope
// device array of kNoAllocScratchBytes, which every GPU that loads the library holds until exit.
//   limits: n <= VO_NOALLOC_MAX_POINTS correspondences, params->max_iters <= VO_NOALLOC_MAX_ITERS.
//   All calls (of all contexts on a device) share the array and it keeps the last call's contents: the regions a
//   run reads before writing are cleared per call; calls that overlap on the GPU race - serialise them.
constexpr int kNoAllocMaxPoints = VO_NOALLOC_MAX_POINTS;
constexpr int kNoAllocMaxIters = VO_NOALLOC_MAX_ITERS;
constexpr int kNoAllocRngLen = 32768;   // >= kNoAllocMaxIters * 16 + 1024 draws (ensure_rng_table's rule)
static_assert(kNoAllocRngLen >= kNoAllocMaxIters * 16 + 1024, "draw table too short for the iteration limit");

constexpr size_t al(size_t v) { return (v + 255) & ~(size_t)255; }
struct NoAllocLayout {   // byte offsets into the scratch array, each a multiple of 256
    static constexpr size_t kSetRows = kNoAllocMaxIters > kPnpChunk0 ? kNoAllocMaxIters : kPnpChunk0;
    static constexpr size_t X = 0;                                                         // n x 3 f64
    static constexpr size_t P = X + al(sizeof(double) * 3 * kNoAllocMaxPoints);            // n x 2 f32
    static constexpr size_t q4 = P + al(sizeof(float) * 2 * kNoAllocMaxPoints);            // n x 4 f32
    static constexpr size_t qv = q4 + al(sizeof(float) * 4 * kNoAllocMaxPoints);           // n f32
    static constexpr size_t q4i = qv + al(sizeof(float) * kNoAllocMaxPoints);
    static constexpr size_t qvi = q4i + al(sizeof(float) * 4 * kNoAllocMaxPoints);
    static constexpr size_t sets = qvi + al(sizeof(float) * kNoAllocMaxPoints);            // rows x 4 i32
    static constexpr size_t sets_in = sets + al(sizeof(int32_t) * kMinimal * kSetRows);    // caller's table, staged
    static constexpr size_t rng = sets_in + al(sizeof(int32_t) * kMinimal * kNoAllocMaxIters);
    // --- cleared per call from here on ---
    static constexpr size_t hyp = rng + al(sizeof(uint32_t) * kNoAllocRngLen);             // H x 12 f64
    static constexpr size_t counts = hyp + al(sizeof(double) * 12 * kNoAllocMaxIters);     // H i32
    static constexpr size_t mask = counts + al(sizeof(int32_t) * kNoAllocMaxIters);        // n u8
    static constexpr size_t end = mask + al(kNoAllocMaxPoints);
};
constexpr size_t kNoAllocScratchBytes = NoAllocLayout::end;
static_assert(kNoAllocScratchBytes <= (1u << 20), "the no-alloc scratch is meant to stay modest (<= 1 MiB)");

// In host code this name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

__global__ void noalloc_rng_fill_kernel(uint32_t* tab, int len)
{
    // cv::RNG(2^64-1): raw 32-bit outputs in draw order (as ensure_rng_table draws them on the host)
    uint64_t s = 0xFFFFFFFFFFFFFFFFULL;
    for (int i = 0; i < len; i++) {
        s = (uint64_t)(uint32_t)s * 4164903690ULL + (uint32_t)(s >> 32);
        tab[i] = (uint32_t)s;
    }
}

constexpr int kNoAllocMaxDevices = 64;
static unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {};   // per device: address of ITS copy of the array

// once per device, with that device current: the array's device address, and the draw table inside it
static int noalloc_setup(Ctx* c, unsigned char** out)
{
    const int dev = c->device;
    if (dev < 0 || dev >= kNoAllocMaxDevices) return set_error(c, VO_ERR_INVALID, "device %d outside the no-alloc table", dev);
    if (!g_noalloc_ptr[dev]) {
        VO_CUDA(c, cudaSetDevice(dev));
        unsigned char* ptr = nullptr;
        VO_CUDA(c, cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
        if (!ptr) return set_error(c, VO_ERR_CUDA, "cudaGetSymbolAddress returned no address for the no-alloc scratch");
        noalloc_rng_fill_kernel<<<1, 1, 0, c->stream>>>(reinterpret_cast<uint32_t*>(ptr + NoAllocLayout::rng), kNoAllocRngLen);
        VO_CUDA(c, cudaGetLastError());
        VO_CUDA(c, cudaStreamSynchronize(c->stream));
        c->launches++;
        g_noalloc_ptr[dev] = ptr;
    }
    *out = g_noalloc_ptr[dev];
    return VO_OK;
}

// The launchers take their buffers from the Pipeline: for the duration of one call they are handed views into the
// scratch array (capacity = the stated maximum, so no ensure() in there can reach cudaMalloc), then the real ones back.
struct NoAllocViews {
    Pipeline* p;
    DevBuf<double> stage_d, hyp; DevBuf<float> stage_f, q4, qv, q4i, qvi; DevBuf<int32_t> stage_i, sets, counts;
    DevBuf<uint8_t> mask; DevBuf<uint32_t> rng_tab; int rng_tab_len;
    template <typename T> static void view(DevBuf<T>& b, unsigned char* base, size_t off, size_t cap)
    {
        b.p = reinterpret_cast<T*>(base + off); b.cap = cap;
    }
    NoAllocViews(Pipeline* pp, unsigned char* s) : p(pp), stage_d(pp->stage_d), hyp(pp->hyp), stage_f(pp->stage_f), q4(pp->q4),
        qv(pp->qv), q4i(pp->q4i), qvi(pp->qvi), stage_i(pp->stage_i), sets(pp->sets), counts(pp->counts), mask(pp->mask),
        rng_tab(pp->rng_tab), rng_tab_len(pp->rng_tab_len)
    {
        typedef NoAllocLayout Lay;
        view(p->stage_d, s, Lay::X, (size_t)3 * kNoAllocMaxPoints); view(p->stage_f, s, Lay::P, (size_t)2 * kNoAllocMaxPoints);
        view(p->q4, s, Lay::q4, (size_t)4 * kNoAllocMaxPoints); view(p->qv, s, Lay::qv, kNoAllocMaxPoints);
        view(p->q4i, s, Lay::q4i, (size_t)4 * kNoAllocMaxPoints); view(p->qvi, s, Lay::qvi, kNoAllocMaxPoints);
        view(p->sets, s, Lay::sets, Lay::kSetRows * kMinimal); view(p->stage_i, s, Lay::sets_in, (size_t)kNoAllocMaxIters * kMinimal);
        view(p->rng_tab, s, Lay::rng, kNoAllocRngLen); p->rng_tab_len = kNoAllocRngLen;
        view(p->hyp, s, Lay::hyp, (size_t)12 * kNoAllocMaxIters); view(p->counts, s, Lay::counts, kNoAllocMaxIters);
        view(p->mask, s, Lay::mask, kNoAllocMaxPoints);
    }
    ~NoAllocViews()
    {
        p->stage_d = stage_d; p->hyp = hyp; p->stage_f = stage_f; p->q4 = q4; p->qv = qv; p->q4i = q4i; p->qvi = qvi;
        p->stage_i = stage_i; p->sets = sets; p->counts = counts; p->mask = mask; p->rng_tab = rng_tab; p->rng_tab_len = rng_tab_len;
    }
};

extern "C" {

int vo_solve_pnp_ransac_noalloc(vo_ctx* c, const double* X, const float* P, int n, const double K[9],
                                const vo_pnp_params* params, const int32_t* minimal_sets, double T44[16], double Th12[12],
                                double rvec[3], double tvec[3], uint8_t* inlier_mask, int32_t info[4], int* ok)
{
    VO_NEED(c);
    if (n < 0 || !K || (n > 0 && (!X || !P))) return VO_ERR_INVALID;
    Pipeline* p = c->pipe;
    PnpLaunch L{};
    if (params) L.prm = *params; else vo_pnp_params_default(&L.prm);
    if (n > kNoAllocMaxPoints)
        return set_error(c, VO_ERR_CAPACITY, "vo_solve_pnp_ransac_noalloc: %d correspondences, the static scratch holds at most %d "
                         "(use vo_solve_pnp_ransac)", n, kNoAllocMaxPoints);
    if (L.prm.max_iters > kNoAllocMaxIters)
        return set_error(c, VO_ERR_CAPACITY, "vo_solve_pnp_ransac_noalloc: max_iters %d, the static scratch holds at most %d "
                         "hypotheses (use vo_solve_pnp_ransac)", L.prm.max_iters, kNoAllocMaxIters);
    VO_TRY(drop_ahead(c));   // (a step enqueued ahead keeps its point count in the scalar block this run uses)
    unsigned char* scratch = nullptr;
    VO_TRY(noalloc_setup(c, &scratch));
    VO_TRY(join_side_stream(c));
    NoAllocViews views(p, scratch);   // (restored on every way out)
    // what a run reads before it writes must not be the last call's: hypotheses, counts, mask
    VO_CUDA(c, cudaMemsetAsync(scratch + NoAllocLayout::hyp, 0, NoAllocLayout::end - NoAllocLayout::hyp, c->stream));
    if (n > 0) {
        VO_CUDA(c, cudaMemcpyAsync(p->stage_d.p, X, sizeof(double) * 3 * n, cudaMemcpyDefault, c->stream));
        VO_CUDA(c, cudaMemcpyAsync(p->stage_f.p, P, sizeof(float) * 2 * n, cudaMemcpyDefault, c->stream));
    }
    const size_t m = n > 0 ? n : 1;
    pack_points_kernel<<<(int)((m + 255) / 256), 256, 0, c->stream>>>(p->d_sc, p->stage_d.p, (const float2*)p->stage_f.p, n, 0,
                                                                      (float4*)p->q4.p, p->qv.p, p->sets.p,
                                                                      RngTab{ p->rng_tab.p, p->rng_tab_len });
    c->launches++;
    if (minimal_sets) {
        const size_t H = L.prm.max_iters > 0 ? L.prm.max_iters : 1;
        VO_CUDA(c, cudaMemcpyAsync(p->stage_i.p, minimal_sets, sizeof(int32_t) * kMinimal * H, cudaMemcpyDefault, c->stream));
        L.d_sets_in = p->stage_i.p;
    }
    L.q4 = p->q4.p; L.qv = p->qv.p; L.d_n = nullptr; L.n_upper = n; L.compact = false;
    memcpy(L.K, K, sizeof(L.K));
    p->packed_valid = false;
    VO_TRY(launch_pnp(c, L));
    if (inlier_mask && n > 0) VO_CUDA(c, cudaMemcpyAsync(inlier_mask, p->mask.p, n, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaMemcpyAsync(p->h_sc, p->d_sc, sizeof(DevScalars), cudaMemcpyDeviceToHost, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    VO_TRY(check_rng(c));
    const DevScalars* s = p->h_sc;
    if (ok) *ok = s->ok;
    if (T44) memcpy(T44, s->T, sizeof(double) * 16);
    if (Th12) memcpy(Th12, s->Th, sizeof(double) * 12);
    if (rvec) { if (s->ok) rotation_to_rvec(s->T, rvec); else rvec[0] = rvec[1] = rvec[2] = 0.; }
    if (tvec) { tvec[0] = s->T[3]; tvec[1] = s->T[7]; tvec[2] = s->T[11]; }
    if (info) { info[0] = s->iters_run; info[1] = s->best_hyp; info[2] = s->n_inliers; info[3] = s->lm_iters; }
    return VO_OK;
}

}  // extern "C"
