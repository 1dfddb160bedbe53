// K9: Shi-Tomasi corner detector on the device (the first "next" row).
//
// Replaces cv2.goodFeaturesToTrack(new_frame, maxCorners=500, qualityLevel=0.03, minDistance=10,
// mask=mask, blockSize=4, gradientSize=3, useHarrisDetector=False) at
//   /root/reference/src/vo_pipeline.py:381 (LandmarkTriangulator.find_new_candidates_shi, :357-390)
// including the reference's mask (single pixels of the well-tracked candidates, :371-378).
// The arithmetic follows the probed behaviour of the OpenCV build restated in oracle/gftt.c
// (float32 Sobel with the scale folded into the smoothing taps and its fused/unfused evaluation,
// float64 box sums, unfused min-eigenvalue formula), so the corner list is the one cv2 returns.
//
//   shi_eig_kernel     32x16 px tiles: u8 tile -> (Dx, Dy) tile in shared memory -> 4x4 box sums of the
//                      products in float64 -> min eigenvalue; masked maximum by atomicMax on an
//                      order-preserving key.  HBM: 1 B/px in (+1 B/px mask), 4 B/px out.
//   shi_nms_kernel     threshold (quality x max), 3x3 non-maximum test on the thresholded map, interior
//                      and mask rules; survivors appended as 64-bit keys (value | pixel index).
//   shi_select_kernel  one CTA: bitonic sort of the keys (value descending, address descending - the
//                      order cv2 sorts in), then the greedy minimum-distance acceptance in that order:
//                      a warp decides 32 candidates per step (accepted-corner grid of minDistance cells
//                      + an in-batch dependency resolution with ballots), identical to the sequential scan.
#include "ctx.cuh"

namespace vo {

namespace {

constexpr int TX = 32, TY = 16;
constexpr int kMaxBlock = 8;
constexpr int SW = TX + kMaxBlock + 1, SH = TY + kMaxBlock + 1;   // source tile (virtual coordinates)
constexpr int DW = TX + kMaxBlock - 1, DH = TY + kMaxBlock - 1;   // derivative tile
constexpr int kSortSmem = 16384;                                  // keys sorted in shared memory (128 KB)
constexpr int kCellCap = 4;

__device__ __forceinline__ int reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

// float -> unsigned, monotonic (negative values included); 0 is below every float
__device__ __forceinline__ unsigned ordered_key(float v)
{
    const unsigned u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_value(unsigned k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void __launch_bounds__(TX * TY)
shi_eig_kernel(const uint8_t* __restrict__ img, int w, int h, int pitch, int block, int vec_cols, float k1, float k0,
               const uint8_t* __restrict__ mask, float* __restrict__ eig, DetScalars* sc)
{
    __shared__ uint8_t src[SH][SW + 3];
    __shared__ float dxs[DH][DW + 1], dys[DH][DW + 1];
    __shared__ unsigned wmax[TX * TY / 32];
    const int an = block / 2;
    const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY;
    const int tid = threadIdx.y * TX + threadIdx.x;
    const int sw = TX + block + 1, sh = TY + block + 1;      // source tile actually used
    const int dw = TX + block - 1, dh = TY + block - 1;
    const int sx0 = x0 - an - 1, sy0 = y0 - an - 1;           // virtual origin of the source tile
    for (int i = tid; i < sw * sh; i += TX * TY) {
        const int ty = i / sw, tx = i - ty * sw;
        src[ty][tx] = img[(size_t)reflect101(sy0 + ty, h) * pitch + reflect101(sx0 + tx, w)];
    }
    __syncthreads();
    // derivative at the REFLECTED position of every tile entry (the box filter mirrors the product
    // images, not the source), read through the tile by actual coordinates
    for (int i = tid; i < dw * dh; i += TX * TY) {
        const int ty = i / dw, tx = i - ty * dw;
        const int rx = reflect101(x0 - an + tx, w), ry = reflect101(y0 - an + ty, h);
        const int xm = min(max(reflect101(rx - 1, w) - sx0, 0), sw - 1), xc = min(max(rx - sx0, 0), sw - 1),
                  xp = min(max(reflect101(rx + 1, w) - sx0, 0), sw - 1);
        const int ym = min(max(reflect101(ry - 1, h) - sy0, 0), sh - 1), yc = min(max(ry - sy0, 0), sh - 1),
                  yp = min(max(reflect101(ry + 1, h) - sy0, 0), sh - 1);
        const float s0 = (float)((int)src[ym][xp] - (int)src[ym][xm]);
        const float s1 = (float)((int)src[yc][xp] - (int)src[yc][xm]);
        const float s2 = (float)((int)src[yp][xp] - (int)src[yp][xm]);
        dxs[ty][tx] = __fmaf_rn(__fadd_rn(s0, s2), k1, __fmul_rn(s1, k0));
        float rt, rb;
        {
            const float a = (float)src[ym][xm], b = (float)src[ym][xc], c = (float)src[ym][xp];
            rt = rx < vec_cols ? __fmaf_rn(c, k1, __fmaf_rn(b, k0, __fmul_rn(a, k1)))
                               : __fadd_rn(__fadd_rn(__fmul_rn(a, k1), __fmul_rn(b, k0)), __fmul_rn(c, k1));
        }
        {
            const float a = (float)src[yp][xm], b = (float)src[yp][xc], c = (float)src[yp][xp];
            rb = rx < vec_cols ? __fmaf_rn(c, k1, __fmaf_rn(b, k0, __fmul_rn(a, k1)))
                               : __fadd_rn(__fadd_rn(__fmul_rn(a, k1), __fmul_rn(b, k0)), __fmul_rn(c, k1));
        }
        dys[ty][tx] = __fsub_rn(rb, rt);
    }
    __syncthreads();
    const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
    unsigned key = 0;
    if (x < w && y < h) {
        double sxx = 0., sxy = 0., syy = 0.;
        for (int j = 0; j < block; j++) {
            double rxx = 0., rxy = 0., ryy = 0.;
            for (int i = 0; i < block; i++) {
                const float dx = dxs[threadIdx.y + j][threadIdx.x + i], dy = dys[threadIdx.y + j][threadIdx.x + i];
                rxx += (double)__fmul_rn(dx, dx); rxy += (double)__fmul_rn(dx, dy); ryy += (double)__fmul_rn(dy, dy);
            }
            sxx += rxx; sxy += rxy; syy += ryy;
        }
        const float a = __fmul_rn((float)sxx, 0.5f), b = (float)sxy, c = __fmul_rn((float)syy, 0.5f);
        const float t = __fsub_rn(a, c);
        const float v = __fsub_rn(__fadd_rn(a, c), __fsqrt_rn(__fadd_rn(__fmul_rn(t, t), __fmul_rn(b, b))));
        const size_t idx = (size_t)y * w + x;
        eig[idx] = v;
        if (!mask || mask[idx]) key = ordered_key(v);
    }
    key = __reduce_max_sync(0xffffffffu, key);
    if ((tid & 31) == 0) wmax[tid >> 5] = key;
    __syncthreads();
    if (tid < 32) {
        key = tid < TX * TY / 32 ? wmax[tid] : 0u;
        key = __reduce_max_sync(0xffffffffu, key);
        if (tid == 0 && key) atomicMax(&sc->max_key, key);
    }
}

// mask := 255 everywhere except the truncated positions of the tracked candidates (reference
// vo_pipeline.py:371-378: int() of the PREVIOUS-frame positions of the candidates KLT kept)
__global__ void shi_mask_points_kernel(uint8_t* __restrict__ mask, int w, int h, const float2* __restrict__ pts,
                                       const uint8_t* __restrict__ status, const int32_t* __restrict__ n_dev, int n_upper)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = n_dev ? min(*n_dev, n_upper) : n_upper;
    if (i >= n || (status && !status[i])) return;
    const float2 p = pts[i];
    // Python int(): truncation toward zero (so (-1, 0) maps to pixel 0)
    const long long x = (long long)truncf(p.x), y = (long long)truncf(p.y);
    if (x >= 0 && x < w && y >= 0 && y < h) mask[(size_t)y * w + x] = 0;
}

__global__ void shi_nms_kernel(const float* __restrict__ eig, int w, int h, const uint8_t* __restrict__ mask,
                               double quality, unsigned long long* __restrict__ keys, DetScalars* sc)
{
    __shared__ int sCount, sBase;
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    if (tid == 0) sCount = 0;
    __syncthreads();
    const unsigned mk = sc->max_key;
    const float maxv = mk ? key_value(mk) : 0.f;
    const float thr = (float)((double)maxv * quality);
    bool keep = false;
    float v = 0.f;
    if (x >= 1 && x < w - 1 && y >= 1 && y < h - 1) {
        const size_t idx = (size_t)y * w + x;
        v = eig[idx];
        if (v > thr && v != 0.f && (!mask || mask[idx])) {
            keep = true;
#pragma unroll
            for (int j = -1; j <= 1; j++)
#pragma unroll
                for (int i = -1; i <= 1; i++) {
                    const float u = eig[idx + (ptrdiff_t)j * w + i];
                    if ((u > thr ? u : 0.f) > v) keep = false;   // 3x3 dilation of the thresholded map
                }
        }
    }
    int slot = 0;
    if (keep) slot = atomicAdd(&sCount, 1);
    __syncthreads();
    if (tid == 0 && sCount) sBase = atomicAdd(&sc->n_cand, sCount);
    __syncthreads();
    if (keep) keys[sBase + slot] = ((unsigned long long)ordered_key(v) << 32) | (unsigned)(y * w + x);
}

// descending bitonic sort of `n2` (power of two) keys by a single CTA
template <typename Ptr>
__device__ void bitonic_desc(Ptr a, int n2)
{
    for (int k = 2; k <= n2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < (n2 >> 1); t += blockDim.x) {
                const int i = 2 * t - (t & (j - 1));      // lower index of the pair
                const int p = i + j;
                const unsigned long long u = a[i], v = a[p];
                const bool desc = (i & k) == 0;
                if (desc ? (u < v) : (u > v)) { a[i] = v; a[p] = u; }
            }
            __syncthreads();
        }
}

__global__ void __launch_bounds__(1024, 1)
shi_select_kernel(unsigned long long* __restrict__ keys, int key_cap, int w, int h, int max_corners, float min_dist,
                  int cell, int gw, int gh, int* __restrict__ grid_cnt, unsigned* __restrict__ grid_pts,
                  float2* __restrict__ out, int out_cap, DetScalars* sc)
{
    extern __shared__ unsigned long long skeys[];
    const int n = min(sc->n_cand, key_cap);
    int n2 = 1;
    while (n2 < n) n2 <<= 1;
    const bool in_smem = n2 <= kSortSmem;
    const bool use_dist = min_dist >= 1.f;
    if (use_dist)
        for (int i = threadIdx.x; i < gw * gh; i += blockDim.x) grid_cnt[i] = 0;
    if (in_smem) {
        for (int i = threadIdx.x; i < n2; i += blockDim.x) skeys[i] = i < n ? keys[i] : 0ull;
        __syncthreads();
        bitonic_desc(skeys, n2);
    } else {
        for (int i = n + threadIdx.x; i < n2; i += blockDim.x) keys[i] = 0ull;   // key buffer is a power of two
        __syncthreads();
        bitonic_desc(keys, n2);
    }
    __syncthreads();
    if (threadIdx.x >= 32) return;
    const unsigned long long* sorted = in_smem ? skeys : keys;
    const int lane = threadIdx.x;
    const unsigned lt = (1u << lane) - 1u;
    const float md2 = min_dist * min_dist;
    const int limit = max_corners > 0 ? min(max_corners, out_cap) : out_cap;
    volatile const int* vcnt = grid_cnt;
    volatile const unsigned* vpts = grid_pts;
    int accepted = 0, overflow = 0;
    for (int pos = 0; pos < n && accepted < limit; pos += 32) {
        const int k = pos + lane;
        const bool valid = k < n;
        const unsigned idx = valid ? (unsigned)(sorted[k] & 0xffffffffull) : 0u;
        const int y = idx / w, x = idx - y * w;
        const int cx = x / cell, cy = y / cell;
        bool alive = valid;
        unsigned nb = 0;
        if (use_dist) {
            if (alive) {
                const int x1 = max(cx - 1, 0), x2 = min(cx + 1, gw - 1), y1 = max(cy - 1, 0), y2 = min(cy + 1, gh - 1);
                for (int yy = y1; yy <= y2 && alive; yy++)
                    for (int xx = x1; xx <= x2 && alive; xx++) {
                        const int c = yy * gw + xx;
                        const int cnt = min(vcnt[c], kCellCap);
                        for (int e = 0; e < cnt; e++) {
                            const unsigned p = vpts[c * kCellCap + e];
                            const float dx = (float)(x - (int)(p & 0xffffu)), dy = (float)(y - (int)(p >> 16));
                            if (dx * dx + dy * dy < md2) { alive = false; break; }
                        }
                    }
            }
            // earlier candidates of this batch that would block this one if accepted
            const unsigned me = (unsigned)x | ((unsigned)y << 16);
            for (int j = 0; j < 31; j++) {
                const unsigned pj = __shfl_sync(0xffffffffu, me, j);
                const int xj = pj & 0xffffu, yj = pj >> 16;
                const float dx = (float)(x - xj), dy = (float)(y - yj);
                // cv2 only looks into the 3x3 cells around the candidate
                const bool near = abs(xj / cell - cx) <= 1 && abs(yj / cell - cy) <= 1 && dx * dx + dy * dy < md2;
                if (j < lane && near) nb |= 1u << j;
            }
        }
        bool undecided = alive, acc = false;
        for (;;) {
            const unsigned accm = __ballot_sync(0xffffffffu, acc);
            const unsigned undm = __ballot_sync(0xffffffffu, undecided);
            if (!undm) break;
            if (undecided) {
                if (nb & accm) undecided = false;
                else if (!(nb & undm)) { acc = true; undecided = false; }
            }
        }
        const unsigned accm = __ballot_sync(0xffffffffu, acc);
        const int rank = __popc(accm & lt);
        const bool take = acc && accepted + rank < limit;
        if (take) {
            out[accepted + rank] = make_float2((float)x, (float)y);
            if (use_dist) {
                const int c = cy * gw + cx;
                const int slot = atomicAdd(&grid_cnt[c], 1);
                if (slot < kCellCap) grid_pts[c * kCellCap + slot] = (unsigned)x | ((unsigned)y << 16);
                else overflow = 1;
            }
        }
        accepted += __popc(__ballot_sync(0xffffffffu, take));
        __threadfence_block();
        __syncwarp();
    }
    overflow = __any_sync(0xffffffffu, overflow);
    if (lane == 0) { sc->n_out = accepted; sc->overflow = overflow; sc->n_sorted = n; }
}

__global__ void det_reset_kernel(DetScalars* sc)
{
    sc->max_key = 0; sc->n_cand = 0; sc->n_out = 0; sc->overflow = 0; sc->n_sorted = 0;
}

// corners (float2) -> the double2 staging layout cand_add_kernel reads
__global__ void corners_to_f64_kernel(const float2* __restrict__ in, const DetScalars* sc, int cap, double2* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < min(sc->n_out, cap)) out[i] = make_double2((double)in[i].x, (double)in[i].y);
}

}  // namespace

void gftt_defaults(vo_gftt_params* p)
{
    p->max_corners = 500; p->quality_level = 0.03; p->min_distance = 10.0; p->block_size = 4; p->vec_cols = -1;
}

int detector_free(Ctx* c)
{
    Detector& d = c->det;
    cudaFree(d.eig.p); cudaFree(d.mask.p); cudaFree(d.keys.p); cudaFree(d.grid_cnt.p); cudaFree(d.grid_pts.p);
    cudaFree(d.out.p); cudaFree(d.d_sc);
    if (d.h_sc) cudaFreeHost(d.h_sc);
    d = Detector();
    return VO_OK;
}

// Enqueues the detector on c->stream.  d_mask: w*h u8 on the device or nullptr.  Results stay on the
// device: d.out (float2 corners in cv2's order), d.d_sc->n_out.
int launch_gftt(Ctx* c, const FrameSlot& f, const uint8_t* d_mask, const vo_gftt_params& prm)
{
    Detector& d = c->det;
    const int w = f.w, h = f.h;
    if (prm.block_size < 1 || prm.block_size > kMaxBlock)
        return set_error(c, VO_ERR_INVALID, "goodFeaturesToTrack blockSize %d outside 1..%d", prm.block_size, kMaxBlock);
    if (w < 3 || h < 3 || w > 65535 || h > 65535) return set_error(c, VO_ERR_INVALID, "frame %dx%d not supported by the detector", w, h);
    if (!(prm.quality_level > 0.) || prm.min_distance < 0.)
        return set_error(c, VO_ERR_INVALID, "goodFeaturesToTrack needs qualityLevel > 0 and minDistance >= 0");
    const size_t px = (size_t)w * h;
    size_t kcap = 1;
    while (kcap < px) kcap <<= 1;
    VO_TRY(ensure(c, d.eig, px));
    VO_TRY(ensure(c, d.keys, kcap));
    const int out_cap = prm.max_corners > 0 ? prm.max_corners : (int)px;
    VO_TRY(ensure(c, d.out, (size_t)out_cap * 2));
    const int cell = prm.min_distance >= 1. ? (int)lrint(prm.min_distance) : 1;
    const int gw = (w + cell - 1) / cell, gh = (h + cell - 1) / cell;
    VO_TRY(ensure(c, d.grid_cnt, (size_t)gw * gh));
    VO_TRY(ensure(c, d.grid_pts, (size_t)gw * gh * kCellCap));
    if (!d.d_sc) {
        VO_CUDA(c, cudaMalloc((void**)&d.d_sc, sizeof(DetScalars)));
        VO_CUDA(c, cudaMallocHost((void**)&d.h_sc, sizeof(DetScalars)));
        VO_CUDA(c, cudaFuncSetAttribute(shi_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSortSmem * 8));
    }
    const double scale = 1.0 / ((double)(1 << 2) * prm.block_size) / 255.0;
    const int vec_cols = prm.vec_cols < 0 ? (w & ~31) : prm.vec_cols;
    det_reset_kernel<<<1, 1, 0, c->stream>>>(d.d_sc);
    shi_eig_kernel<<<dim3((w + TX - 1) / TX, (h + TY - 1) / TY), dim3(TX, TY), 0, c->stream>>>(
        f.raw, w, h, w, prm.block_size, vec_cols, (float)scale, (float)(2.0 * scale), d_mask, d.eig.p, d.d_sc);
    shi_nms_kernel<<<dim3((w + 31) / 32, (h + 7) / 8), dim3(32, 8), 0, c->stream>>>(
        d.eig.p, w, h, d_mask, prm.quality_level, (unsigned long long*)d.keys.p, d.d_sc);
    shi_select_kernel<<<1, 1024, kSortSmem * 8, c->stream>>>(
        (unsigned long long*)d.keys.p, (int)kcap, w, h, prm.max_corners, (float)prm.min_distance, cell, gw, gh,
        d.grid_cnt.p, d.grid_pts.p, (float2*)d.out.p, out_cap, d.d_sc);
    c->launches += 4;
    VO_CUDA(c, cudaGetLastError());
    d.out_cap = out_cap;
    return VO_OK;
}

// mask on the device from a point list (status may be null): 255 everywhere, 0 at int(x), int(y)
int build_point_mask(Ctx* c, int w, int h, const float* d_pts, const uint8_t* d_status, const int32_t* d_n, int n_upper)
{
    Detector& d = c->det;
    VO_TRY(ensure(c, d.mask, (size_t)w * h));
    VO_CUDA(c, cudaMemsetAsync(d.mask.p, 255, (size_t)w * h, c->stream));
    if (n_upper > 0) {
        shi_mask_points_kernel<<<(n_upper + 255) / 256, 256, 0, c->stream>>>(d.mask.p, w, h, (const float2*)d_pts, d_status, d_n, n_upper);
        c->launches++;
        VO_CUDA(c, cudaGetLastError());
    }
    return VO_OK;
}

int corners_to_stage(Ctx* c, double* d_stage)
{
    Detector& d = c->det;
    corners_to_f64_kernel<<<(d.out_cap + 255) / 256, 256, 0, c->stream>>>((const float2*)d.out.p, d.d_sc, d.out_cap, (double2*)d_stage);
    c->launches++;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

}  // namespace vo

using namespace vo;

extern "C" {

void vo_gftt_params_default(vo_gftt_params* p) { gftt_defaults(p); }

int vo_good_features_to_track(vo_ctx* c, int64_t frame_key, const uint8_t* mask, const float* exclude_pts, int n_exclude,
                              const vo_gftt_params* params, float* corners, int cap, int* n_corners, float* eig_out,
                              int32_t info[4])
{
    if (!c || cap < 0 || (cap > 0 && !corners) || n_exclude < 0 || (n_exclude > 0 && !exclude_pts)) return VO_ERR_INVALID;
    if (mask && n_exclude > 0) return set_error(c, VO_ERR_INVALID, "pass either a mask image or a point list, not both");
    vo_gftt_params prm;
    if (params) prm = *params; else gftt_defaults(&prm);
    FrameSlot* f = get_frame(c, frame_key);
    if (!f) return set_error(c, VO_ERR_NOFRAME, "frame %lld not resident", (long long)frame_key);
    Detector& d = c->det;
    const size_t px = (size_t)f->w * f->h;
    const uint8_t* d_mask = nullptr;
    if (mask) {
        VO_TRY(ensure(c, d.mask, px));
        VO_CUDA(c, cudaMemcpyAsync(d.mask.p, mask, px, cudaMemcpyDefault, c->stream));
        d_mask = d.mask.p;
    } else if (n_exclude > 0) {
        VO_TRY(ensure(c, c->u_pts, (size_t)n_exclude * 2));
        VO_CUDA(c, cudaMemcpyAsync(c->u_pts.p, exclude_pts, sizeof(float) * 2 * n_exclude, cudaMemcpyDefault, c->stream));
        VO_TRY(build_point_mask(c, f->w, f->h, c->u_pts.p, nullptr, nullptr, n_exclude));
        d_mask = d.mask.p;
    }
    VO_TRY(launch_gftt(c, *f, d_mask, prm));
    VO_CUDA(c, cudaMemcpyAsync(d.h_sc, d.d_sc, sizeof(DetScalars), cudaMemcpyDeviceToHost, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    const int n = d.h_sc->n_out;
    if (n_corners) *n_corners = n;
    if (info) { info[0] = d.h_sc->n_sorted; info[1] = d.h_sc->overflow; info[2] = (int32_t)d.h_sc->max_key; info[3] = 0; }
    if (d.h_sc->overflow) return set_error(c, VO_ERR_CAPACITY, "accepted-corner grid cell overflow");
    if (n > cap) return set_error(c, VO_ERR_CAPACITY, "%d corners, buffer %d", n, cap);
    if (n > 0) VO_CUDA(c, cudaMemcpyAsync(corners, d.out.p, sizeof(float) * 2 * n, cudaMemcpyDefault, c->stream));
    if (eig_out) VO_CUDA(c, cudaMemcpyAsync(eig_out, d.eig.p, sizeof(float) * px, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

}  // extern "C"
