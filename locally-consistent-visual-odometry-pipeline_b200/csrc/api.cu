// C-ABI entry points of libvo_b200.so: context, frame cache, unit-level KLT.
// (Pipeline-level entry points live in pipeline.cu.)
#include <stdarg.h>
#include "ctx.cuh"

namespace vo {

thread_local std::string g_create_error;

int set_error(Ctx* c, int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    else g_create_error = buf;
    return code;
}

static FrameSlot* find_slot(Ctx* c, int64_t key)
{
    if (c->hint.set && c->hint.key == key) flush_frame_hint(c);
    for (auto& s : c->slots)
        if (s.valid && s.key == key) {
            s.stamp = ++c->clock;
            s.use_step = c->step_id; c->open_uses++;
            if (s.pending) {   // filled by the ingest stream: everything enqueued from here on waits for it
                s.pending = false;
                if (c->last_ingest_valid && c->last_ingest_key == key) c->last_ingest_valid = false;
                if (cudaStreamWaitEvent(c->stream, s.ready, 0) != cudaSuccess) return nullptr;
            }
            return &s;
        }
    return nullptr;
}

FrameSlot* get_frame(Ctx* c, int64_t key) { return find_slot(c, key); }

int join_side_stream(Ctx* c)
{
    if (!c->side_inflight) return VO_OK;
    c->side_inflight = false; c->side_writes_counts = false;
    VO_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    return VO_OK;
}

int join_pose_stream(Ctx* c)
{
    if (!c->pose_inflight) return VO_OK;
    c->pose_inflight = false;
    VO_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_pose_done, 0));
    return VO_OK;
}

int close_step(Ctx* c)
{
    VO_CUDA(c, cudaEventRecord(c->ev_step_end[c->step_id & 3], c->stream));
    c->step_id++; c->open_uses = 0;
    return VO_OK;
}

}  // namespace vo

using namespace vo;

extern "C" {

void vo_klt_params_default(vo_klt_params* p)
{
    p->win = 21; p->max_level = 3; p->max_iters = 30; p->epsilon = 0.01; p->min_eig_thr = 1e-4;
    p->fb_check = 0; p->fb_thresh = 1.0;
}

void vo_pnp_params_default(vo_pnp_params* p)
{
    p->max_iters = 100; p->confidence = 0.99999; p->reproj_thresh = 2.0; p->refine_iters = 20;
    p->min_inliers = 4;
}

int vo_abi_version(void) { return VO_B200_ABI_VERSION; }

int vo_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { cudaGetLastError(); return -(int)e; }
    return n;
}

int vo_create(vo_ctx** out, int device, int max_w, int max_h, int track_capacity, int hyp_capacity,
              void* stream_handle)
{
    if (!out) return VO_ERR_INVALID;
    *out = nullptr;
    if (max_w <= 0 || max_h <= 0) return set_error(nullptr, VO_ERR_INVALID, "bad frame size %dx%d", max_w, max_h);
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess)
        return set_error(nullptr, VO_ERR_CUDA, "cudaSetDevice(%d) failed: %s - libvo_b200 has no CPU fallback", device, cudaGetErrorString(e));
    vo_ctx* c = new (std::nothrow) vo_ctx();
    if (!c) return set_error(nullptr, VO_ERR_NOMEM, "out of host memory");
    c->device = device; c->max_w = max_w; c->max_h = max_h;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) { delete c; return set_error(nullptr, VO_ERR_CUDA, "cudaGetDeviceProperties failed: %s", cudaGetErrorString(e)); }
    c->sm_count = prop.multiProcessorCount;
    if (stream_handle) { c->stream = (cudaStream_t)stream_handle; c->own_stream = false; }
    else {
        e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) { delete c; return set_error(nullptr, VO_ERR_CUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(e)); }
        c->own_stream = true;
    }
    if (cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream3, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream4, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream5, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_compact, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_trk, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_ransac, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_pose_side, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_snap, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_pose_done, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_ingest_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_step_end[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_step_end[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_step_end[2], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_step_end[3], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess) {
        delete c;
        return set_error(nullptr, VO_ERR_CUDA, "side stream creation failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    c->pinned_stage_bytes = (size_t)max_w * max_h;
    e = cudaMallocHost((void**)&c->pinned_stage, c->pinned_stage_bytes);
    if (e == cudaSuccess) e = cudaMallocHost((void**)&c->pinned_stage2, c->pinned_stage_bytes);
    if (e != cudaSuccess) { delete c; return set_error(nullptr, VO_ERR_NOMEM, "cudaMallocHost failed: %s", cudaGetErrorString(e)); }
    int r = pipeline_init(c, track_capacity, hyp_capacity);
    if (r != VO_OK) { g_create_error = c->err; vo_destroy(c); return r; }
    *out = c;
    return VO_OK;
}

int vo_destroy(vo_ctx* c)
{
    if (!c) return VO_OK;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->stream2) cudaStreamSynchronize(c->stream2);
    if (c->stream3) cudaStreamSynchronize(c->stream3);
    if (c->stream4) cudaStreamSynchronize(c->stream4);
    if (c->stream5) cudaStreamSynchronize(c->stream5);
    map_free(c);
    pipeline_free(c);
    detector_free(c);
    for (auto& s : c->slots) { cudaFree(s.raw); cudaFree(s.img_pool); cudaFree(s.deriv_pool); if (s.ready) cudaEventDestroy(s.ready); }
    cudaFree(c->u_pts.p); cudaFree(c->u_out.p); cudaFree(c->u_err.p); cudaFree(c->u_status.p);
    cudaFree(c->u_iters.p); cudaFree(c->u_n.p);
    cudaFree(c->fb_pts.p); cudaFree(c->fb_pts2.p); cudaFree(c->fb_status.p); cudaFree(c->fb_status2.p);
    cudaFreeHost(c->pinned_stage);
    cudaFreeHost(c->pinned_stage2);
    if (c->stream3) cudaStreamDestroy(c->stream3);
    if (c->stream4) cudaStreamDestroy(c->stream4);
    if (c->stream5) cudaStreamDestroy(c->stream5);
    if (c->ev_compact) cudaEventDestroy(c->ev_compact);
    if (c->ev_trk) cudaEventDestroy(c->ev_trk);
    if (c->ev_ransac) cudaEventDestroy(c->ev_ransac);
    if (c->ev_pose_side) cudaEventDestroy(c->ev_pose_side);
    if (c->ev_snap) cudaEventDestroy(c->ev_snap);
    if (c->ev_pose_done) cudaEventDestroy(c->ev_pose_done);
    if (c->ev_ingest_fork) cudaEventDestroy(c->ev_ingest_fork);
    for (auto& e : c->ev_step_end) if (e) cudaEventDestroy(e);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    if (c->stream2) cudaStreamDestroy(c->stream2);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    delete c;
    return VO_OK;
}

const char* vo_last_error(vo_ctx* c) { return c ? c->err.c_str() : g_create_error.c_str(); }

int vo_synchronize(vo_ctx* c)
{
    if (!c) return VO_ERR_INVALID;
    VO_TRY(flush_frame_hint(c));
    VO_TRY(join_side_stream(c));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream3));
    return VO_OK;
}

int64_t vo_launch_count(vo_ctx* c) { return c ? c->launches : 0; }

int vo_alloc_pinned(void** out, uint64_t bytes)
{
    if (!out) return VO_ERR_INVALID;
    cudaError_t e = cudaMallocHost(out, bytes);
    if (e != cudaSuccess) { *out = nullptr; return set_error(nullptr, VO_ERR_NOMEM, "cudaMallocHost(%llu) failed: %s", (unsigned long long)bytes, cudaGetErrorString(e)); }
    return VO_OK;
}

int vo_free_pinned(void* p)
{
    if (p) cudaFreeHost(p);
    return VO_OK;
}

int vo_has_frame(vo_ctx* c, int64_t key)
{
    if (!c) return 0;
    if (c->hint.set && c->hint.key == key) flush_frame_hint(c);
    for (auto& s : c->slots) if (s.valid && s.key == key) return 1;
    return 0;
}

int vo_drop_frames(vo_ctx* c)
{
    if (!c) return VO_ERR_INVALID;
    c->hint.set = false;
    VO_TRY(join_side_stream(c));
    for (auto& s : c->slots) s.valid = false;
    return VO_OK;
}

}  // extern "C"

// Upload + pyramid of one frame on `st` (the main stream, or the ingest stream for a prefetch).
static int upload_impl(vo_ctx* c, const uint8_t* img, int w, int h, int pitch, int64_t key, bool ingest, bool overlap = false)
{
    VO_NVTX(ingest ? "vo:frame_ingest(H2D+pyramid)" : "vo:frame_upload(H2D+pyramid)");
    if (!c || !img) return VO_ERR_INVALID;
    if (w <= 0 || h <= 0 || w > c->max_w || h > c->max_h || pitch < w)
        return set_error(c, VO_ERR_INVALID, "frame %dx%d (pitch %d) outside ctx limits %dx%d", w, h, pitch, c->max_w, c->max_h);
    if (ingest ? vo_has_frame(c, key) : (find_slot(c, key) != nullptr)) return VO_OK;
    cudaStream_t st = ingest ? c->stream3 : c->stream;
    if (!ingest) VO_TRY(join_side_stream(c));   // a pyramid slot may be recycled under the side stream's reads
    // victim: a free slot, else the least recently used one - but not a frame that was handed over early
    // and has not been used yet (its stamp is older than those of the frames tracked since), unless
    // that hint is stale (two newer ingests) or nothing else is left
    FrameSlot* s = nullptr;
    for (auto& t : c->slots) if (!t.valid) { s = &t; break; }
    if (!s) {
        FrameSlot* any = &c->slots[0];
        for (auto& t : c->slots) {
            if (t.stamp < any->stamp) any = &t;
            const bool fresh_hint = t.pending && c->ingest_seq - t.ingest_seq < 2;
            if (!fresh_hint && (!s || t.stamp < s->stamp)) s = &t;
        }
        if (!s) s = any;
    }
    if (ingest) {
        // The recycled slot may still be read by work that is enqueued but not finished.  If the stages
        // of the current frame are already on the stream and none of them looked this slot up, its last
        // readers belong to a closed step: wait for that step's end only, so the ingest runs UNDER the
        // current frame's stages.  Otherwise order it behind everything enqueued so far (stages enqueued
        // after this call then overlap it).
        // (`overlap`: called from inside a pose call, whose kernels - this frame's last - are on the stream.)
        // The event of the step that last used the slot; a ring entry reused since belongs to a later
        // step, which is only more conservative.
        if ((c->open_uses > 0 || overlap) && s->use_step < c->step_id) {
            VO_CUDA(c, cudaStreamWaitEvent(st, c->ev_step_end[s->use_step & 3], 0));
        } else {
            VO_CUDA(c, cudaEventRecord(c->ev_ingest_fork, c->stream));
            VO_CUDA(c, cudaStreamWaitEvent(st, c->ev_ingest_fork, 0));
            if (c->side_inflight) VO_CUDA(c, cudaStreamWaitEvent(st, c->ev_join, 0));
        }
    }
    s->valid = false;
    if (ingest) s->ingest_seq = ++c->ingest_seq;
    if (s->pending) {   // an ingest into this very slot is still in flight: write behind it
        VO_CUDA(c, cudaStreamWaitEvent(st, s->ready, 0));
        s->pending = false;
    }
    const size_t bytes = (size_t)w * h;
    if (bytes > s->raw_bytes) {
        if (s->raw) { cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->stream3); cudaFree(s->raw); s->raw = nullptr; s->raw_bytes = 0; }
        VO_CUDA(c, cudaMalloc((void**)&s->raw, bytes));
        s->raw_bytes = bytes;
    }
    // Pageable host memory is staged through a pinned buffer of the ctx (one per stream) so the copy
    // is a true async DMA; pinned host and device sources are copied directly.
    cudaPointerAttributes at;
    cudaError_t e = cudaPointerGetAttributes(&at, img);
    bool pageable = (e != cudaSuccess) || at.type == cudaMemoryTypeUnregistered;
    if (e != cudaSuccess) cudaGetLastError();
    if (pageable) {
        // the previous async copy out of the staging buffer must have completed
        VO_CUDA(c, cudaStreamSynchronize(st));
        uint8_t* stage = ingest ? c->pinned_stage2 : c->pinned_stage;
        for (int y = 0; y < h; y++) memcpy(stage + (size_t)y * w, img + (size_t)y * pitch, w);
        VO_CUDA(c, cudaMemcpyAsync(s->raw, stage, bytes, cudaMemcpyHostToDevice, st));
    } else {
        if (pitch == w) VO_CUDA(c, cudaMemcpyAsync(s->raw, img, bytes, cudaMemcpyDefault, st));
        else VO_CUDA(c, cudaMemcpy2DAsync(s->raw, w, img, pitch, w, h, cudaMemcpyDefault, st));
    }
    s->w = w; s->h = h;
    if (!ingest) { s->use_step = c->step_id; c->open_uses++; }   // written on the main stream in the open step
    VO_TRY(build_pyramid(c, *s, 3, st));
    if (ingest) {
        if (!s->ready) VO_CUDA(c, cudaEventCreateWithFlags(&s->ready, cudaEventDisableTiming));
        VO_CUDA(c, cudaEventRecord(s->ready, st));
        s->pending = true;
        c->last_ingest_key = key; c->last_ingest_valid = true;
    }
    s->key = key; s->valid = true; s->stamp = ++c->clock;
    return VO_OK;
}

namespace vo {
int flush_frame_hint(Ctx* c, bool under_current_stages)
{
    if (!c->hint.set) return VO_OK;
    c->hint.set = false;
    return upload_impl(static_cast<vo_ctx*>(c), c->hint.img, c->hint.w, c->hint.h, c->hint.pitch, c->hint.key, true,
                       under_current_stages);
}
}  // namespace vo

extern "C" {

int vo_upload_frame(vo_ctx* c, const uint8_t* img, int w, int h, int pitch, int64_t key)
{
    return upload_impl(c, img, w, h, pitch, key, false);
}

// SURVEY.md 8f-2 (frame ingest): the same upload + pyramid, enqueued on the ctx's ingest stream so
// that it runs under the tracking of the current frame; the first stage that uses `key` waits for it
// on the device (no host synchronisation anywhere).
int vo_prefetch_frame(vo_ctx* c, const uint8_t* img, int w, int h, int pitch, int64_t key)
{
    if (c && c->hint.set) VO_TRY(flush_frame_hint(c));
    return upload_impl(c, img, w, h, pitch, key, true);
}

// The same, announced now and enqueued by the next pose call (vo_pnp_ransac / vo_pnp_ransac_promote)
// between its last kernel launch and its wait for the pose - the host-side cost of the enqueue (~20 us)
// then hides under the GPU's PnP - or earlier, by the first call that asks for frame_key.  img must
// stay valid and unchanged until then (and, as for vo_prefetch_frame, until the copy has run).
int vo_prefetch_frame_deferred(vo_ctx* c, const uint8_t* img, int w, int h, int pitch, int64_t key)
{
    if (!c || !img) return VO_ERR_INVALID;
    if (w <= 0 || h <= 0 || w > c->max_w || h > c->max_h || pitch < w)
        return set_error(c, VO_ERR_INVALID, "frame %dx%d (pitch %d) outside ctx limits %dx%d", w, h, pitch, c->max_w, c->max_h);
    if (c->hint.set) VO_TRY(flush_frame_hint(c));
    c->hint.set = true; c->hint.img = img; c->hint.w = w; c->hint.h = h; c->hint.pitch = pitch; c->hint.key = key;
    return VO_OK;
}

int vo_get_pyramid_level(vo_ctx* c, int64_t key, int level, uint8_t* img, int16_t* deriv, int* w, int* h)
{
    if (!c) return VO_ERR_INVALID;
    FrameSlot* s = find_slot(c, key);
    if (!s) return set_error(c, VO_ERR_NOFRAME, "frame %lld not resident", (long long)key);
    if (level < 0 || level >= s->pyr.nlevels) return set_error(c, VO_ERR_INVALID, "level %d of %d", level, s->pyr.nlevels);
    const LevelDesc& d = s->pyr.lv[level];
    if (w) *w = d.w;
    if (h) *h = d.h;
    if (img)
        VO_CUDA(c, cudaMemcpy2DAsync(img, d.w, d.img + (size_t)kPad * d.stride + kPad, d.stride, d.w, d.h, cudaMemcpyDeviceToHost, c->stream));
    if (deriv)
        VO_CUDA(c, cudaMemcpy2DAsync(deriv, (size_t)d.w * 4, d.deriv + (size_t)kPad * d.stride + kPad, (size_t)d.stride * 4, (size_t)d.w * 4, d.h, cudaMemcpyDeviceToHost, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

int vo_calc_optical_flow_pyr_lk(vo_ctx* c, int64_t prev_key, int64_t next_key, const float* pts, int n,
                                float* next_pts, uint8_t* status, float* err, int32_t* iters,
                                const vo_klt_params* params)
{
    if (!c || n < 0 || (n > 0 && (!pts || !next_pts || !status))) return VO_ERR_INVALID;
    vo_klt_params prm;
    if (params) prm = *params; else vo_klt_params_default(&prm);
    FrameSlot* a = find_slot(c, prev_key);
    FrameSlot* b = find_slot(c, next_key);
    if (!a || !b) return set_error(c, VO_ERR_NOFRAME, "frame %lld not resident", (long long)(!a ? prev_key : next_key));
    if (n == 0) return VO_OK;
    VO_TRY(ensure(c, c->u_pts, (size_t)n * 2));
    VO_TRY(ensure(c, c->u_out, (size_t)n * 2));
    VO_TRY(ensure(c, c->u_err, (size_t)n));
    VO_TRY(ensure(c, c->u_status, (size_t)n));
    VO_TRY(ensure(c, c->u_iters, (size_t)n));
    VO_CUDA(c, cudaMemcpyAsync(c->u_pts.p, pts, sizeof(float) * 2 * n, cudaMemcpyDefault, c->stream));
    VO_TRY(launch_klt(c, a->pyr, b->pyr, c->u_pts.p, nullptr, n, c->u_out.p, c->u_status.p, c->u_err.p, c->u_iters.p, prm));
    if (prm.fb_check) VO_TRY(klt_fb_check(c, a->pyr, b->pyr, c->u_pts.p, nullptr, n, c->u_out.p, c->u_status.p, prm));
    VO_CUDA(c, cudaMemcpyAsync(next_pts, c->u_out.p, sizeof(float) * 2 * n, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaMemcpyAsync(status, c->u_status.p, n, cudaMemcpyDefault, c->stream));
    if (err) VO_CUDA(c, cudaMemcpyAsync(err, c->u_err.p, sizeof(float) * n, cudaMemcpyDefault, c->stream));
    if (iters) VO_CUDA(c, cudaMemcpyAsync(iters, c->u_iters.p, sizeof(int32_t) * n, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

}  // extern "C"
