// Internal context of libvo_b200.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/vo_b200.h"

namespace vo {

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
};

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;  // elements
};

struct Ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
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
    DevBuf<float> fb_pts;
    DevBuf<uint8_t> fb_status;

    struct Pipeline* pipe = nullptr;  // device-resident Markov state + PnP scratch (pipeline.cu)
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
        cudaFree(b.p);
    }
    b.p = np_;
    b.cap = ncap;
    return VO_OK;
}

// ---- stage launchers implemented in the .cu files -------------------------------------------
int build_pyramid(Ctx* c, FrameSlot& s, int max_level_cap);
int launch_klt(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const float* d_pts,
               const int32_t* d_n, int n_upper, float* d_out, uint8_t* d_status, float* d_err,
               int32_t* d_iters, const vo_klt_params& prm);

// optional extra (off in parity mode): track `out` back into `prev` and clear status where the
// round trip misses the start point by more than prm.fb_thresh
int klt_fb_check(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const float* d_pts,
                 const int32_t* d_n, int n_upper, const float* d_out, uint8_t* d_status,
                 const vo_klt_params& prm);
FrameSlot* get_frame(Ctx* c, int64_t key);
int pipeline_init(Ctx* c, int track_capacity, int hyp_capacity);
void pipeline_free(Ctx* c);

}  // namespace vo

struct vo_ctx : public vo::Ctx {};
