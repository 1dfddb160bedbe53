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
//                      order cv2 sorts in), then cv2's greedy minimum-distance acceptance evaluated for
//                      all candidates in parallel rounds over a cell-bucketed candidate list (a
//                      candidate depends on earlier ones only, so the fixed point is the sequential scan).
#include <cooperative_groups.h>
#include "ctx.cuh"

namespace vo {

namespace {

constexpr int TX = 32, TY = 16;
constexpr int kMaxBlock = 8;
constexpr int SW = TX + kMaxBlock + 1, SH = TY + kMaxBlock + 1;   // source tile (virtual coordinates)
constexpr int DW = TX + kMaxBlock - 1, DH = TY + kMaxBlock - 1;   // derivative tile

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

// exclusive prefix sum over the CTA (blockDim.x == 1024); *total = sum of all v
__device__ __forceinline__ int block_excl_scan(int v, int* total)
{
    __shared__ int wsum[32];
    __shared__ int tot;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = wsum[lane], winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += t;
        }
        wsum[lane] = winc - w;
        if (lane == 31) tot = winc;
    }
    __syncthreads();
    const int r = wsum[wid] + inc - v;
    *total = tot;
    __syncthreads();      // wsum/tot may be reused by the next call
    return r;
}

struct SelArrays {
    unsigned* xy;      // (x | y << 16) in sorted order
    uint8_t* state;    // 0 undecided, 1 accepted, 2 rejected
    int* items;        // candidate ranks bucketed by cell
    int* cell_cnt;     // gw*gh
    int* cell_start;   // gw*gh + 1
};

// cv2's greedy minimum-distance acceptance - "take a candidate unless an accepted one lies closer than
// minDistance inside the 3x3 minDistance-cells around it", scanned in sorted order - evaluated for all
// n candidates in parallel: candidates are bucketed by cell (counting sort) and, in rounds, every
// undecided candidate looks at the EARLIER candidates around it: one of them accepted => rejected;
// none of them undecided => accepted.  A candidate's fate depends on earlier ones only, so the fixed
// point is exactly the sequential result whatever the schedule.  Writes the first `limit` accepted
// candidates (sorted order) to out and returns how many were accepted (counting stops at limit).
// A.xy must hold the n positions; all pointers may be shared or global memory.
__device__ int greedy_select(int n, const SelArrays& A, bool use_dist, int cell, int gw, int gh, float md2, int limit,
                             float2* __restrict__ out, int* rounds_out)
{
    const int T = threadIdx.x, NT = blockDim.x;
    const int ncell = gw * gh;
    int rounds = 0;
    if (use_dist) {
        for (int i = T; i < ncell; i += NT) A.cell_cnt[i] = 0;
        __syncthreads();
        for (int i = T; i < n; i += NT) {
            const unsigned p = A.xy[i];
            A.state[i] = 0;
            atomicAdd(&A.cell_cnt[(int)((p >> 16) / cell) * gw + (int)((p & 0xffffu) / cell)], 1);
        }
        __syncthreads();
        int running = 0;
        for (int base = 0; base < ncell; base += NT) {
            const int c = base + T;
            const int v = c < ncell ? A.cell_cnt[c] : 0;
            int total;
            const int ex = block_excl_scan(v, &total);
            if (c < ncell) A.cell_start[c] = running + ex;
            running += total;
        }
        if (T == 0) A.cell_start[ncell] = running;
        __syncthreads();
        for (int i = T; i < n; i += NT) {
            const unsigned p = A.xy[i];
            const int c = (int)((p >> 16) / cell) * gw + (int)((p & 0xffffu) / cell);
            A.items[A.cell_start[c] + atomicSub(&A.cell_cnt[c], 1) - 1] = i;
        }
        __syncthreads();
        int pending;
        do {
            pending = 0;
            rounds++;
            for (int i = T; i < n; i += NT) {
                if (A.state[i]) continue;
                const unsigned p = A.xy[i];
                const int x = p & 0xffffu, y = p >> 16;
                const int cx = x / cell, cy = y / cell;
                const int x1 = max(cx - 1, 0), x2 = min(cx + 1, gw - 1), y1 = max(cy - 1, 0), y2 = min(cy + 1, gh - 1);
                int verdict = 1;      // accepted unless an earlier neighbour says otherwise
                for (int yy = y1; yy <= y2 && verdict != 2; yy++) {
                    // the cells of one grid row are adjacent in the item list
                    const int pb = A.cell_start[yy * gw + x1], pe = A.cell_start[yy * gw + x2 + 1];
                    for (int q = pb; q < pe; q++) {
                        const int j = A.items[q];
                        if (j >= i) continue;
                        const unsigned pj = A.xy[j];
                        const float dx = (float)(x - (int)(pj & 0xffffu)), dy = (float)(y - (int)(pj >> 16));
                        if (dx * dx + dy * dy < md2) {
                            const int sj = A.state[j];
                            if (sj == 1) { verdict = 2; break; }
                            if (sj == 0) verdict = 0;
                        }
                    }
                }
                if (verdict) A.state[i] = (uint8_t)verdict;
                else pending = 1;
            }
        } while (__syncthreads_or(pending));
    }
    int running = 0;
    for (int base = 0; base < n && running < limit; base += NT) {
        const int i = base + T;
        const int a = i < n && (!use_dist || A.state[i] == 1);
        int total;
        const int ex = block_excl_scan(a, &total);
        if (a && running + ex < limit) {
            const unsigned p = A.xy[i];
            out[running + ex] = make_float2((float)(p & 0xffffu), (float)(p >> 16));
        }
        running += total;
    }
    *rounds_out = rounds;
    return running;
}

constexpr int kTopSmem = 8192;     // candidates the shared-memory path sorts and selects (64 KB of keys)
constexpr int kHistBins = 2048;

// One CTA.  Sorts the candidate keys descending (= cv2's order: value, then address) and runs the
// greedy selection.  Only the best-valued candidates can end up among the first maxCorners accepted,
// so a first attempt takes the top ~4 x maxCorners candidates by value (threshold from a histogram of
// the keys), sorts and selects them in shared memory; the result is exact when it yields maxCorners
// corners.  Otherwise (few corners, or no limit) everything is redone on all candidates - in shared
// memory if they fit, else in global memory.
__global__ void __launch_bounds__(1024, 1)
shi_select_kernel(unsigned long long* __restrict__ keys, int key_cap, int w, int h, int max_corners, float min_dist,
                  double quality, int cell, int gw, int gh, int cells_in_smem, SelArrays G,
                  float2* __restrict__ out, int out_cap, DetScalars* sc)
{
    extern __shared__ unsigned long long skeys[];
    __shared__ int hist[kHistBins];
    __shared__ int s_m, s_bin;
    const int T = threadIdx.x, NT = blockDim.x;
    if (!sc->fallback) return;         // the cluster kernel has produced the result
    const int n = min(sc->n_cand, key_cap);
    const bool use_dist = min_dist >= 1.f;
    const int ncell = gw * gh;
    const float md2 = min_dist * min_dist;
    const int limit = max_corners > 0 ? min(max_corners, out_cap) : out_cap;
    const long long t0 = clock64();
    // shared-memory arrays: [keys 8 B x kTopSmem -> reused as xy + items][state][cell_cnt][cell_start]
    SelArrays S;
    S.xy = reinterpret_cast<unsigned*>(skeys);
    S.items = reinterpret_cast<int*>(skeys) + kTopSmem;
    S.state = reinterpret_cast<uint8_t*>(skeys + kTopSmem);
    S.cell_cnt = cells_in_smem ? reinterpret_cast<int*>(S.state + kTopSmem) : G.cell_cnt;
    S.cell_start = cells_in_smem ? S.cell_cnt + ncell : G.cell_start;

    unsigned long long th = 0ull;      // keys >= th take part in this attempt
    int m = n;
    if (n > 1024 && limit < n) {
        const unsigned kmax = sc->max_key;
        const float maxv = kmax ? key_value(kmax) : 0.f;
        const unsigned kmin = ordered_key((float)((double)maxv * quality));
        int shift = 0;
        while (((kmax > kmin ? kmax - kmin : 0u) >> shift) >= (unsigned)kHistBins) shift++;
        for (int i = T; i < kHistBins; i += NT) hist[i] = 0;
        __syncthreads();
        for (int i = T; i < n; i += NT) {
            const unsigned k32 = (unsigned)(keys[i] >> 32);
            atomicAdd(&hist[min((k32 > kmin ? k32 - kmin : 0u) >> shift, (unsigned)kHistBins - 1u)], 1);
        }
        __syncthreads();
        if (T < 32) {      // lowest bin whose suffix count reaches the target
            const int target = min(max(4 * limit, 1024), kTopSmem * 3 / 4);
            constexpr int SEG = kHistBins / 32;
            const int hi = kHistBins - 1 - T * SEG;          // lane 0 owns the top segment
            int seg = 0;
            for (int b = 0; b < SEG; b++) seg += hist[hi - b];
            int inc = seg;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, d);
                if (T >= d) inc += t;
            }
            const unsigned reach = __ballot_sync(0xffffffffu, inc >= target);
            if (reach == 0u) { if (T == 0) { s_bin = 0; s_m = n; } }
            else if (T == __ffs(reach) - 1) {
                int cum = inc - seg, b = 0;
                for (; b < SEG; b++) { cum += hist[hi - b]; if (cum >= target) break; }
                s_bin = hi - b; s_m = cum;
            }
        }
        __syncthreads();
        if (s_bin > 0 && s_m <= kTopSmem) { th = (unsigned long long)(kmin + ((unsigned)s_bin << shift)) << 32; m = s_m; }
    }
    int rounds = 0, accepted = 0;
    long long t1 = t0, t2 = t0;
    for (int attempt = 0; attempt < 2; attempt++) {
        if (m <= kTopSmem) {
            int m2 = 1;
            while (m2 < m) m2 <<= 1;
            if (T == 0) s_m = 0;
            __syncthreads();
            if (th == 0ull) { for (int i = T; i < n; i += NT) skeys[i] = keys[i]; }
            else {
                for (int i = T; i < n; i += NT) {
                    const unsigned long long k = keys[i];
                    if (k >= th) skeys[atomicAdd(&s_m, 1)] = k;
                }
            }
            for (int i = m + T; i < m2; i += NT) skeys[i] = 0ull;
            __syncthreads();
            bitonic_desc(skeys, m2);
            t1 = clock64();
            // keys -> positions, in place (every key is in registers before anything is overwritten)
            unsigned pos[kTopSmem / 1024];
#pragma unroll
            for (int r = 0; r < kTopSmem / 1024; r++) {
                const int i = T + r * 1024;
                if (i < m) {
                    const unsigned idx = (unsigned)(skeys[i] & 0xffffffffull);
                    const unsigned y = idx / (unsigned)w;
                    pos[r] = (idx - y * w) | (y << 16);
                }
            }
            __syncthreads();
#pragma unroll
            for (int r = 0; r < kTopSmem / 1024; r++) {
                const int i = T + r * 1024;
                if (i < m) S.xy[i] = pos[r];
            }
            __syncthreads();
            accepted = greedy_select(m, S, use_dist, cell, gw, gh, md2, limit, out, &rounds);
        } else {
            int n2 = 1;
            while (n2 < n) n2 <<= 1;
            for (int i = n + T; i < n2; i += NT) keys[i] = 0ull;   // the key buffer is a power of two
            __syncthreads();
            bitonic_desc(keys, n2);
            t1 = clock64();
            for (int i = T; i < n; i += NT) {
                const unsigned idx = (unsigned)(keys[i] & 0xffffffffull);
                const unsigned y = idx / (unsigned)w;
                G.xy[i] = (idx - y * w) | (y << 16);
            }
            __syncthreads();
            accepted = greedy_select(n, G, use_dist, cell, gw, gh, md2, limit, out, &rounds);
        }
        t2 = clock64();
        if (accepted >= limit || m >= n) break;
        th = 0ull; m = n;       // the best candidates did not yield maxCorners corners: take them all
        __syncthreads();
    }
    if (T == 0) {
        sc->n_out = min(accepted, limit); sc->overflow = 0; sc->n_sorted = m; sc->rounds = rounds;
        sc->cyc[0] = (int)(t1 - t0); sc->cyc[1] = 0; sc->cyc[2] = (int)(t2 - t1);
    }
}

// ---- fast path: the same selection on a thread-block cluster of 8 CTAs (8 SMs) -------------------
// The single-CTA kernel above is instruction-bound on one SM (~190 warp instructions per candidate).
// Here the candidates are partitioned by VALUE into 8 slabs of about n/8 keys (boundaries from a
// 2048-bin histogram of the keys, like a sample sort); CTA r sorts slab r in its shared memory, so the
// concatenation of the slabs is the globally sorted list; positions, states and per-cell linked lists
// (head/next) live in global memory (L2), and the parallel rounds of the greedy selection run on all
// 8 SMs with a cluster barrier (which also invalidates L1) between rounds.  Anything outside its
// limits (no maxCorners, no distance, a slab that does not fit, > 60 rounds) is left to the single-CTA
// kernel, which runs right after and returns at once unless sc->fallback is set.
constexpr int kCl = 8;
constexpr int kSlabCap = 8192;

__global__ void __cluster_dims__(kCl, 1, 1) __launch_bounds__(1024, 1)
shi_select_cluster_kernel(const unsigned long long* __restrict__ keys, int key_cap, int w, int h, int max_corners,
                          float min_dist, double quality, int cell, int gw, int gh, SelArrays G, int* __restrict__ ghist,
                          float2* __restrict__ out, int out_cap, DetScalars* sc)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ unsigned long long skeys[];
    __shared__ int hist[kHistBins];
    __shared__ uint8_t slab_of[kHistBins];
    __shared__ int slab_cnt[kCl], slab_off[kCl];
    __shared__ int s_m;
    const int T = threadIdx.x, NT = blockDim.x;
    const int r = (int)cluster.block_rank();
    const int gT = r * NT + T, GN = kCl * NT;
    const int n = min(sc->n_cand, key_cap);
    const int ncell = gw * gh;
    const int limit = max_corners > 0 ? min(max_corners, out_cap) : out_cap;
    int* head = G.cell_cnt;      // per cell: rank of the most recently linked candidate, -1 = empty
    int* next = G.items;         // per candidate: the next one in its cell
    int* gflag = ghist + kHistBins;   // 3 rotating "somebody is still undecided" flags
    const long long t0 = clock64();
    if (max_corners <= 0 || min_dist < 1.f || n > kCl * kSlabCap) {   // uniform over the cluster
        if (gT == 0) sc->fallback = 1;
        return;
    }
    const unsigned kmax = sc->max_key;
    const float maxv = kmax ? key_value(kmax) : 0.f;
    const unsigned kmin = ordered_key((float)((double)maxv * quality));
    int shift = 0;
    while (((kmax > kmin ? kmax - kmin : 0u) >> shift) >= (unsigned)kHistBins) shift++;
    // ---- histogram of the keys (ghist was zeroed by det_reset_kernel) and empty cell lists
    for (int i = T; i < kHistBins; i += NT) hist[i] = 0;
    if (T < kCl) { slab_cnt[T] = 0; slab_off[T] = 0x7fffffff; }
    __syncthreads();
    for (int i = gT; i < n; i += GN) {
        const unsigned k32 = (unsigned)(keys[i] >> 32);
        atomicAdd(&hist[min((k32 > kmin ? k32 - kmin : 0u) >> shift, (unsigned)kHistBins - 1u)], 1);
    }
    for (int i = gT; i < ncell; i += GN) head[i] = -1;
    __syncthreads();
    for (int i = T; i < kHistBins; i += NT)
        if (hist[i]) atomicAdd(&ghist[i], hist[i]);
    cluster.sync();
    // ---- slabs: bin b belongs to slab min(7, (#keys in bins above b) / ceil(n/8)); every CTA computes all
    const int per = max((n + kCl - 1) / kCl, 1);
    {
        int running = 0;
        for (int base = 0; base < kHistBins; base += NT) {
            const int b = kHistBins - 1 - (base + T);          // descending bins
            const int v = __ldcg(&ghist[b]);
            int total;
            const int above = running + block_excl_scan(v, &total);
            const int sl = min(kCl - 1, above / per);
            slab_of[b] = (uint8_t)sl;
            hist[b] = v;
            if (v) { atomicAdd(&slab_cnt[sl], v); atomicMin(&slab_off[sl], above); }
            running += total;
        }
    }
    if (T == 0) s_m = 0;
    __syncthreads();
    bool bad = false;
    for (int k = 0; k < kCl; k++) bad |= slab_cnt[k] > kSlabCap;
    if (bad) {                                                  // same decision in every CTA
        if (gT == 0) sc->fallback = 1;
        return;
    }
    const int m = slab_cnt[r], off = m ? slab_off[r] : 0;
    // ---- my slab: gather, sort in shared memory, publish positions, link into the cell lists
    for (int i = T; i < n; i += NT) {
        const unsigned long long k = keys[i];
        const unsigned k32 = (unsigned)(k >> 32);
        if (slab_of[min((k32 > kmin ? k32 - kmin : 0u) >> shift, (unsigned)kHistBins - 1u)] == r) skeys[atomicAdd(&s_m, 1)] = k;
    }
    int m2 = 1;
    while (m2 < m) m2 <<= 1;
    for (int i = m + T; i < m2; i += NT) skeys[i] = 0ull;
    __syncthreads();
    bitonic_desc(skeys, m2);
    for (int i = T; i < m; i += NT) {
        const unsigned idx = (unsigned)(skeys[i] & 0xffffffffull);
        const unsigned y = idx / (unsigned)w, x = idx - y * w;
        G.xy[off + i] = x | (y << 16);
        G.state[off + i] = 0;
        next[off + i] = atomicExch(&head[(int)(y / cell) * gw + (int)(x / cell)], off + i);
    }
    const long long t1 = clock64();
    cluster.sync();
    // ---- rounds on all 8 SMs
    const float md2 = min_dist * min_dist;
    int rounds = 0;
    bool give_up = false;
    for (;;) {
        int pending = 0;
        if (gT == 0) gflag[(rounds + 1) % 3] = 0;
        for (int i = gT; i < n; i += GN) {
            if (G.state[i]) continue;
            const unsigned p = G.xy[i];
            const int x = p & 0xffffu, y = p >> 16;
            const int cx = x / cell, cy = y / cell;
            const int x1 = max(cx - 1, 0), x2 = min(cx + 1, gw - 1), y1 = max(cy - 1, 0), y2 = min(cy + 1, gh - 1);
            int verdict = 1;
            for (int yy = y1; yy <= y2 && verdict != 2; yy++)
                for (int xx = x1; xx <= x2 && verdict != 2; xx++)
                    for (int j = head[yy * gw + xx]; j >= 0; j = next[j]) {
                        if (j >= i) continue;
                        const unsigned pj = G.xy[j];
                        const float dx = (float)(x - (int)(pj & 0xffffu)), dy = (float)(y - (int)(pj >> 16));
                        if (dx * dx + dy * dy < md2) {
                            const int sj = __ldcg(&G.state[j]);
                            if (sj == 1) { verdict = 2; break; }
                            if (sj == 0) verdict = 0;
                        }
                    }
            if (verdict) G.state[i] = (uint8_t)verdict;
            else pending = 1;
        }
        pending = __syncthreads_or(pending);
        if (pending && T == 0) atomicOr(&gflag[rounds % 3], 1);
        cluster.sync();
        const int any = __ldcg(&gflag[rounds % 3]);
        rounds++;
        if (!any) break;
        if (rounds >= 60) { give_up = true; break; }
    }
    if (give_up) {
        if (gT == 0) sc->fallback = 1;
        return;
    }
    if (r != 0) return;
    // ---- CTA 0: the first `limit` accepted candidates in sorted order
    int running = 0;
    for (int base = 0; base < n && running < limit; base += NT) {
        const int i = base + T;
        const int a = i < n && __ldcg(&G.state[i]) == 1;
        int total;
        const int ex = block_excl_scan(a, &total);
        if (a && running + ex < limit) {
            const unsigned p = G.xy[i];
            out[running + ex] = make_float2((float)(p & 0xffffu), (float)(p >> 16));
        }
        running += total;
    }
    if (T == 0) {
        sc->n_out = min(running, limit); sc->overflow = 0; sc->n_sorted = n; sc->rounds = rounds; sc->fallback = 0;
        sc->cyc[0] = (int)(t1 - t0); sc->cyc[1] = 0; sc->cyc[2] = (int)(clock64() - t1);
    }
}

__global__ void det_reset_kernel(DetScalars* sc, int* __restrict__ ghist, int fallback)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { sc->max_key = 0; sc->n_cand = 0; sc->n_out = 0; sc->overflow = 0; sc->n_sorted = 0; sc->rounds = 0; sc->fallback = fallback; }
    if (i < kHistBins + 8) ghist[i] = 0;
}

// corners (float2) -> the double2 staging layout cand_add_kernel reads
__global__ void corners_to_f64_kernel(const float2* __restrict__ in, const DetScalars* sc, int cap, double2* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < min(sc->n_out, cap)) out[i] = make_double2((double)in[i].x, (double)in[i].y);
}

}  // namespace

constexpr int kSelSmemMax = 200 * 1024;

void gftt_defaults(vo_gftt_params* p)
{
    p->max_corners = 500; p->quality_level = 0.03; p->min_distance = 10.0; p->block_size = 4; p->vec_cols = -1;
    p->use_cluster = 1;
}

int detector_free(Ctx* c)
{
    Detector& d = c->det;
    cudaFree(d.eig.p); cudaFree(d.mask.p); cudaFree(d.keys.p); cudaFree(d.cell_cnt.p); cudaFree(d.cell_start.p); cudaFree(d.items.p); cudaFree(d.xy.p); cudaFree(d.state.p); cudaFree(d.ghist.p);
    cudaFree(d.out.p); cudaFree(d.d_sc);
    if (d.h_sc) cudaFreeHost(d.h_sc);
    d = Detector();
    return VO_OK;
}

// Enqueues the detector on c->stream.  d_mask: w*h u8 on the device or nullptr.  Results stay on the
// device: d.out (float2 corners in cv2's order), d.d_sc->n_out.
int launch_gftt(Ctx* c, const FrameSlot& f, const uint8_t* d_mask, const vo_gftt_params& prm)
{
    VO_NVTX("vo:shi_tomasi");
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
    VO_TRY(ensure(c, d.cell_cnt, (size_t)gw * gh));
    VO_TRY(ensure(c, d.cell_start, (size_t)gw * gh + 1));
    VO_TRY(ensure(c, d.items, kcap));
    VO_TRY(ensure(c, d.xy, kcap));
    VO_TRY(ensure(c, d.state, kcap));
    if (!d.d_sc) {
        VO_CUDA(c, cudaMalloc((void**)&d.d_sc, sizeof(DetScalars)));
        VO_CUDA(c, cudaMallocHost((void**)&d.h_sc, sizeof(DetScalars)));
        VO_CUDA(c, cudaFuncSetAttribute(shi_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSelSmemMax));
        VO_CUDA(c, cudaFuncSetAttribute(shi_select_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSlabCap * 8));
    }
    const double scale = 1.0 / ((double)(1 << 2) * prm.block_size) / 255.0;
    const int vec_cols = prm.vec_cols < 0 ? (w & ~31) : prm.vec_cols;
    VO_TRY(ensure(c, d.ghist, (size_t)kHistBins + 8));
    const bool use_cluster = prm.use_cluster != 0;
    det_reset_kernel<<<(kHistBins + 8 + 255) / 256, 256, 0, c->stream>>>(d.d_sc, d.ghist.p, use_cluster ? 0 : 1);
    shi_eig_kernel<<<dim3((w + TX - 1) / TX, (h + TY - 1) / TY), dim3(TX, TY), 0, c->stream>>>(
        f.raw, w, h, w, prm.block_size, vec_cols, (float)scale, (float)(2.0 * scale), d_mask, d.eig.p, d.d_sc);
    shi_nms_kernel<<<dim3((w + 31) / 32, (h + 7) / 8), dim3(32, 8), 0, c->stream>>>(
        d.eig.p, w, h, d_mask, prm.quality_level, (unsigned long long*)d.keys.p, d.d_sc);
    // dynamic shared memory: keys (reused as positions + bucket items), states, and the cell arrays if they fit
    const size_t base_smem = (size_t)kTopSmem * 8 + kTopSmem;
    const size_t cell_smem = ((size_t)gw * gh * 2 + 1) * sizeof(int);
    const int cells_in_smem = base_smem + cell_smem <= (size_t)kSelSmemMax;
    SelArrays G{ d.xy.p, d.state.p, d.items.p, d.cell_cnt.p, d.cell_start.p };
    if (use_cluster) {
        shi_select_cluster_kernel<<<kCl, 1024, (size_t)kSlabCap * 8, c->stream>>>(
            (const unsigned long long*)d.keys.p, (int)kcap, w, h, prm.max_corners, (float)prm.min_distance, prm.quality_level,
            cell, gw, gh, G, d.ghist.p, (float2*)d.out.p, out_cap, d.d_sc);
        c->launches++;
    }
    shi_select_kernel<<<1, 1024, base_smem + (cells_in_smem ? cell_smem : 0), c->stream>>>(
        (unsigned long long*)d.keys.p, (int)kcap, w, h, prm.max_corners, (float)prm.min_distance, prm.quality_level,
        cell, gw, gh, cells_in_smem, G, (float2*)d.out.p, out_cap, d.d_sc);
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
    if (info) { info[0] = d.h_sc->n_cand; info[1] = d.h_sc->n_sorted + (d.h_sc->fallback ? 1 << 30 : 0); info[2] = d.h_sc->rounds; info[3] = d.h_sc->cyc[0] / 1000 + 1000 * (d.h_sc->cyc[1] / 1000) + 1000000 * (d.h_sc->cyc[2] / 1000); }
    if (d.h_sc->overflow) return set_error(c, VO_ERR_CAPACITY, "accepted-corner grid cell overflow");
    if (n > cap) return set_error(c, VO_ERR_CAPACITY, "%d corners, buffer %d", n, cap);
    if (n > 0) VO_CUDA(c, cudaMemcpyAsync(corners, d.out.p, sizeof(float) * 2 * n, cudaMemcpyDefault, c->stream));
    if (eig_out) VO_CUDA(c, cudaMemcpyAsync(eig_out, d.eig.p, sizeof(float) * px, cudaMemcpyDefault, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));
    return VO_OK;
}

}  // extern "C"
