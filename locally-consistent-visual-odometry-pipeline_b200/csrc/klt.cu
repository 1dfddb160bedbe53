// K2: pyramidal Lucas-Kanade tracking, one warp per track, all levels in one launch.
//
// Replaces cv2.calcOpticalFlowPyrLK(old, new, pts, None) at
//   /root/reference/src/vo_pipeline.py:215 (landmark tracks P) and :501 (candidate tracks C),
// following the behaviour recorded in SURVEY.md App. A.1 (OpenCV LKTrackerInvoker): 14-bit fixed
// point bilinear weights, int16 window with 5 fractional bits, Scharr derivatives that are zero
// outside the image, float32 2x2 solve, both break rules, level-0-only status writes and the
// post-loop error pass that can still clear status.
//
// Mapping: the 21x21 window is cut into 63 horizontal runs of 7 px (3 per row); lane l owns runs
// l and l+32 (98 % lane use).  A run needs 8+8 bytes of the image (two rows), fetched as 3+3
// aligned 32-bit words and re-aligned with funnel shifts, and its 7 bilinear values cost 14
// DP2A (two 14-bit weights x two u8 taps per instruction) instead of 28 byte loads + 28 IMAD.
// The template window I and its derivatives stay in registers for the whole level; only the
// moving window of the next image is gathered each Gauss-Newton step - from a per-warp shared-memory
// tile holding its neighbourhood (see Tile below), from global memory when it leaves the tile.  The
// per-step sums are exact: products fit int32 per lane and the cross-lane sum is two REDUX.SUM
// instructions (16-bit split), so results do not depend on the lane mapping or summation order.
#include <stdlib.h>
#include "ctx.cuh"

namespace vo {

namespace {

constexpr int kWarpsPerBlock = 4;

__device__ __forceinline__ long long warp_sum_i64(int v)
{
    const unsigned lo = __reduce_add_sync(0xffffffffu, (unsigned)(v & 0xFFFF));
    const int hi = __reduce_add_sync(0xffffffffu, v >> 16);
    return ((long long)hi << 16) + (long long)lo;
}

struct Weights { int w00, w01, w10, w11; };

// iw00=round((1-a)(1-b)2^14) ... with float32 products evaluated in the reference's order
__device__ __forceinline__ Weights make_weights(float a, float b)
{
    const float na = __fsub_rn(1.f, a), nb = __fsub_rn(1.f, b);
    Weights w;
    w.w00 = __float2int_rn(__fmul_rn(__fmul_rn(na, nb), 16384.f));
    w.w01 = __float2int_rn(__fmul_rn(__fmul_rn(a, nb), 16384.f));
    w.w10 = __float2int_rn(__fmul_rn(__fmul_rn(na, b), 16384.f));
    w.w11 = 16384 - w.w00 - w.w01 - w.w10;   // may round to -1: weights are signed 16-bit halves
    return w;
}

// d = c + a.s16[0]*b.u8[0] + a.s16[1]*b.u8[1]   (lo)   /   ... b.u8[2], b.u8[3]   (hi)
__device__ __forceinline__ int dp2a_lo(int a, unsigned b, int c)
{
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi(int a, unsigned b, int c)
{
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int pack_w(int lo, int hi) { return (lo & 0xFFFF) | (hi << 16); }

constexpr int kRun = 7;   // px per run

// 8 consecutive bytes at an arbitrary address as two words (the third aligned word is read even
// when p is aligned: every level row has >= 32 px of border + stride slack behind it)
__device__ __forceinline__ void load8(const uint8_t* __restrict__ p, unsigned& W0, unsigned& W1)
{
    const unsigned* a = reinterpret_cast<const unsigned*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const unsigned sh = ((unsigned)reinterpret_cast<uintptr_t>(p) & 3u) * 8u;
    const unsigned a0 = __ldg(a), a1 = __ldg(a + 1), a2 = __ldg(a + 2);
    W0 = __funnelshift_r(a0, a1, sh);
    W1 = __funnelshift_r(a1, a2, sh);
}

// acc[x] += wl * byte[x] + wr * byte[x+1] for the 7 px of a run (bytes 0..7 in W0:W1)
__device__ __forceinline__ void row7(unsigned W0, unsigned W1, int wpair, int* acc)
{
    const unsigned S1 = __funnelshift_r(W0, W1, 8);   // bytes 1..4
    const unsigned S5 = W1 >> 8;                      // bytes 5..7
    acc[0] = dp2a_lo(wpair, W0, acc[0]);
    acc[1] = dp2a_lo(wpair, S1, acc[1]);
    acc[2] = dp2a_hi(wpair, W0, acc[2]);
    acc[3] = dp2a_hi(wpair, S1, acc[3]);
    acc[4] = dp2a_lo(wpair, W1, acc[4]);
    acc[5] = dp2a_lo(wpair, S5, acc[5]);
    acc[6] = dp2a_hi(wpair, W1, acc[6]);
}

// the 7 bilinear samples ((sum + 2^8) >> 9, 5 fractional bits) of the run starting at p
__device__ __forceinline__ void sample_run(const uint8_t* __restrict__ p, int stride, int wtop, int wbot, int* v)
{
    unsigned T0, T1, B0, B1;
    load8(p, T0, T1);
    load8(p + stride, B0, B1);
#pragma unroll
    for (int x = 0; x < kRun; x++) v[x] = 1 << 8;
    row7(T0, T1, wtop, v);
    row7(B0, B1, wbot, v);
#pragma unroll
    for (int x = 0; x < kRun; x++) v[x] >>= 9;
}

// ---- the moving window's neighbourhood in shared memory ------------------------------------------
// A Gauss-Newton step reads 12 words per lane from 22 different rows of the next image: one warp-wide
// request touches >= 21 L1 lines, so the L1 tag stage (one line per cycle, shared by every warp of the
// SM), not the arithmetic, paces the step (ncu set g: long-scoreboard 39 % of the stall samples, 20
// sectors per request).  Per level the warp copies the (WIN+11) x 48 B neighbourhood of the starting
// position into its own shared-memory tile once (12 requests of ~4 lines) and samples from there while
// the window stays inside it (+-5 rows, -12..+10 columns); outside - or at the very edge of the padded
// level - the step reads global memory as before.  Same bytes either way: results do not change.
template <int WIN> struct Tile {
    static constexpr int kWords = (WIN + 18) / 4 + 3;           // 12 words (48 B) per row
    static constexpr int kPitch = kWords | 1;                    // odd: rows spread over the banks
    static constexpr int kRows = WIN + 1 + 10;                   // 32
    static constexpr int kOyMax = kRows - (WIN + 1);             // 10
    static constexpr int kOxMax = 4 * (kWords - 3) + 3 - (WIN - kRun);   // 25: last run's third word is in the row
    static constexpr int kOx0 = (kOxMax / 2) | 3;                // start offset 12..15 after the alignment
    static constexpr int kOy0 = kOyMax / 2;
    // the template window of a level: (WIN+1) rows of the image (word aligned, <= 7 words) and of the
    // derivative plane ((WIN+1) short2 per row), staged the same way for the per-level set-up
    static constexpr int kIWords = (3 + WIN + 1 + 3) / 4 + 1;     // 8 (7 used + the over-read word of load8_s)
    static constexpr int kIPitch = kIWords | 1;                   // 9
    static constexpr int kDPitch = (WIN + 1) | 1;                 // 23
    static constexpr int kSetupWords = (WIN + 1) * (kIPitch + kDPitch);
    static constexpr int kTileWords = kRows * kPitch;
    static constexpr int kSmemWords = kSetupWords > kTileWords ? kSetupWords : kTileWords;
};

__device__ __forceinline__ void load8_s(const unsigned* __restrict__ row, int xb, unsigned& W0, unsigned& W1)
{
    const unsigned* a = row + (xb >> 2);
    const unsigned sh = ((unsigned)xb & 3u) * 8u;
    const unsigned a0 = a[0], a1 = a[1], a2 = a[2];
    W0 = __funnelshift_r(a0, a1, sh);
    W1 = __funnelshift_r(a1, a2, sh);
}

__device__ __forceinline__ void sample_run_s(const unsigned* __restrict__ row, int pitch, int xb, int wtop, int wbot, int* v)
{
    unsigned T0, T1, B0, B1;
    load8_s(row, xb, T0, T1);
    load8_s(row + pitch, xb, B0, B1);
#pragma unroll
    for (int x = 0; x < kRun; x++) v[x] = 1 << 8;
    row7(T0, T1, wtop, v);
    row7(B0, B1, wbot, v);
#pragma unroll
    for (int x = 0; x < kRun; x++) v[x] >>= 9;
}

// ---- TMA staging of the window tile (A/B switch VO_KLT_TMA=1) ---------------------------------------------------------
// The (WIN+11)-row window tile is a regular 2-D box of the padded level: one elected lane issues a single
// cp.async.bulk.tensor.2d (64 B x 32 rows) into the warp's 128 B aligned tile and the warp waits on its own mbarrier,
// instead of twelve __ldg + st.shared requests per lane.
struct KltTmaps { CUtensorMap m[4]; };

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%1], %0;" ::"r"(count), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %0;" ::"r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned phase)
{
    asm volatile(
        "{\n\t.reg .pred P1;\n\tWAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, unsigned long long* bar, int x, int y)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
                 "l"(reinterpret_cast<unsigned long long>(map)), "r"(smem_u32(bar)), "r"(x), "r"(y)
                 : "memory");
}

// TILE: 0 window gathers from global memory, 1 shared-memory tile filled with __ldg, 2 filled by TMA - a template
// parameter so that the tile's pitch and bounds are constants in the Gauss-Newton step's address arithmetic
template <int WIN, int MINB, int TILE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, MINB)
klt_track_kernel(const PyramidDesc I, const PyramidDesc J, const __grid_constant__ KltTmaps maps, const int top,
                 const KltSegs segs,
                 float2* __restrict__ out, uint8_t* __restrict__ status_out, float* __restrict__ err_out,
                 int32_t* __restrict__ iters_out, const int max_iters, const double eps2,
                 const double min_eig_thr, const int32_t* __restrict__ gate)
{
    static_assert(WIN % kRun == 0, "window side must be a multiple of the run length");
    constexpr int RPR = WIN / kRun;            // runs per row
    constexpr int NRUNS = WIN * RPR;           // 63
    constexpr int NR = (NRUNS + 31) / 32;      // runs per lane
    using TL = Tile<WIN>;
    static_assert((TL::kSmemWords * 4) % 128 == 0, "per-warp tiles must keep the 128 B alignment TMA needs");
    __shared__ __align__(128) unsigned s_tile[kWarpsPerBlock][TL::kSmemWords];
    __shared__ __align__(8) unsigned long long s_bar[kWarpsPerBlock];
    pdl_prologue();
    if (gate && !*gate) return;   // enqueued ahead of a pose that did not come: leave every buffer as it is
    const int lane = threadIdx.x & 31;
    unsigned* const tile = s_tile[threadIdx.x >> 5];
    unsigned long long* const bar = &s_bar[threadIdx.x >> 5];
    constexpr int use_tile = TILE;
    constexpr bool tma = TILE == 2;
    // TMA wants the box origin on a 16 B boundary (an unaligned one is an illegal instruction) and writes dense rows: the
    // box is 64 B wide to cover every alignment, rows 16 words apart - rows two apart share their banks
    constexpr int kWordsT = 16, kOx0T = 13, kOxMaxT = 4 * (kWordsT - 3) + 3 - (WIN - kRun);   // 41
    constexpr int jpitch = tma ? kWordsT : TL::kPitch;
    constexpr int oxmax = tma ? kOxMaxT : TL::kOxMax;
    unsigned phase = 0;
    if (tma) {
        if (lane == 0) {
            mbar_init(bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
    }
    int track = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    // several track lists share one launch (landmarks + candidates; or the lists a tracking step ahead of the
    // pruning covers): list k holds up to n_upper[k] rows (the live count on the device), results at out_off[k]
    float2 pt;
    {
        int k = 0;
        while (k < segs.n - 1 && track >= segs.n_upper[k]) { track -= segs.n_upper[k]; k++; }
        const int32_t* nd = segs.n_dev[k];
        const int n = nd ? min(*nd, segs.n_upper[k]) : segs.n_upper[k];
        if (track >= n) return;
        if (segs.f64[k]) {
            const double2 d = reinterpret_cast<const double2*>(segs.pts[k])[track];
            pt = make_float2((float)d.x, (float)d.y);
        } else {
            pt = segs.pts[k][track];
        }
        track += segs.out_off[k] + (segs.off_dev[k] ? *segs.off_dev[k] : 0);
    }

    const float half = (WIN - 1) * 0.5f;
    float cx = 0.f, cy = 0.f;  // stored nextPts (window centre)
    int st = 1, iters = 0;
    float err = 0.f;

    // window-relative origin of this lane's runs (a slot past the window recomputes the last run
    // and is masked out of every sum)
    int rrow[NR], rcol[NR];
    bool rvalid[NR];
#pragma unroll
    for (int r = 0; r < NR; r++) {
        const int j = lane + 32 * r;
        rvalid[r] = j < NRUNS;
        const int jj = rvalid[r] ? j : NRUNS - 1;
        rrow[r] = jj / RPR;
        rcol[r] = (jj - rrow[r] * RPR) * kRun;
    }

    int Iw[NR][kRun];              // template window, 5 fractional bits
    int dIx[NR][kRun], dIy[NR][kRun];   // its derivatives

    for (int level = top; level >= 0; level--) {
        const LevelDesc LI = I.lv[level];
        const LevelDesc LJ = J.lv[level];
        const float scale = 1.f / (float)(1 << level);
        float px = __fmul_rn(pt.x, scale), py = __fmul_rn(pt.y, scale);
        if (level == top) { cx = px; cy = py; }
        else { cx = __fmul_rn(cx, 2.f); cy = __fmul_rn(cy, 2.f); }

        px = __fsub_rn(px, half); py = __fsub_rn(py, half);
        const int ix = __float2int_rd(px), iy = __float2int_rd(py);
        if (ix < -WIN || ix >= LI.w || iy < -WIN || iy >= LI.h) {
            if (level == 0) { st = 0; err = 0.f; }
            continue;
        }
        Weights w = make_weights(__fsub_rn(px, (float)ix), __fsub_rn(py, (float)iy));
        const int stride = LI.stride;
        const size_t base = (size_t)(iy + kPad) * stride + (ix + kPad);
        int a11 = 0, a12 = 0, a22 = 0;
        // stage the (WIN+1)^2 neighbourhood of image and derivative plane (always inside the padded level:
        // ix + kPad >= kPad - WIN and the row has kPad + slack behind the image)
        unsigned* const sI = tile;
        short2* const sD = reinterpret_cast<short2*>(tile + (WIN + 1) * TL::kIPitch);
        const int xal = (ix + kPad) & 3;
        if (use_tile) {
            const unsigned* srcI = reinterpret_cast<const unsigned*>(LI.img + base - xal);
            const short2* srcD = LI.deriv + base;
            const int wstride = stride >> 2;
            __syncwarp();   // the previous level's iteration tile is no longer read
#pragma unroll
            for (int i = 0; i < ((WIN + 1) * TL::kIWords + 31) / 32; i++) {
                const int idx = i * 32 + lane;
                const int row = idx / TL::kIWords, col = idx - row * TL::kIWords;
                if (idx < (WIN + 1) * TL::kIWords) sI[row * TL::kIPitch + col] = __ldg(srcI + (size_t)row * wstride + col);
            }
#pragma unroll
            for (int i = 0; i < ((WIN + 1) * (WIN + 1) + 31) / 32; i++) {
                const int idx = i * 32 + lane;
                const int row = idx / (WIN + 1), col = idx - row * (WIN + 1);
                if (idx < (WIN + 1) * (WIN + 1)) sD[row * TL::kDPitch + col] = srcD[(size_t)row * stride + col];
            }
            __syncwarp();
        }
        // one run's derivatives: bilinear blend of the Scharr plane, structure-tensor sums
        auto deriv_run = [&](const short2* d, const int dstride, const int r) {
            short2 t = d[0], b = d[dstride];
#pragma unroll
            for (int x = 0; x < kRun; x++) {
                const short2 tn = d[x + 1], bn = d[dstride + x + 1];
                int ixv = (t.x * w.w00 + tn.x * w.w01 + b.x * w.w10 + bn.x * w.w11 + (1 << 13)) >> 14;
                int iyv = (t.y * w.w00 + tn.y * w.w01 + b.y * w.w10 + bn.y * w.w11 + (1 << 13)) >> 14;
                if (!rvalid[r]) { ixv = 0; iyv = 0; }
                dIx[r][x] = ixv; dIy[r][x] = iyv;
                a11 += ixv * ixv; a12 += ixv * iyv; a22 += iyv * iyv;
                t = tn; b = bn;
            }
        };
        if (use_tile) {
#pragma unroll
            for (int r = 0; r < NR; r++) {
                sample_run_s(sI + rrow[r] * TL::kIPitch, TL::kIPitch, xal + rcol[r], pack_w(w.w00, w.w01), pack_w(w.w10, w.w11), Iw[r]);
                deriv_run(sD + rrow[r] * TL::kDPitch + rcol[r], TL::kDPitch, r);
            }
        } else {
#pragma unroll
            for (int r = 0; r < NR; r++) {
                const int off = rrow[r] * stride + rcol[r];
                sample_run(LI.img + base + off, stride, pack_w(w.w00, w.w01), pack_w(w.w10, w.w11), Iw[r]);
                deriv_run(LI.deriv + base + off, stride, r);
            }
        }
        const float FLT_SCALE = 1.f / (1 << 20);
        const float A11 = __fmul_rn(__ll2float_rn(warp_sum_i64(a11)), FLT_SCALE);
        const float A12 = __fmul_rn(__ll2float_rn(warp_sum_i64(a12)), FLT_SCALE);
        const float A22 = __fmul_rn(__ll2float_rn(warp_sum_i64(a22)), FLT_SCALE);
        float D = __fsub_rn(__fmul_rn(A11, A22), __fmul_rn(A12, A12));
        const float dd = __fsub_rn(A11, A22);
        const float disc = __fadd_rn(__fmul_rn(dd, dd), __fmul_rn(__fmul_rn(4.f, A12), A12));
        const float min_eig = __fdiv_rn(__fsub_rn(__fadd_rn(A22, A11), __fsqrt_rn(disc)), (float)(2 * WIN * WIN));
        if ((double)min_eig < min_eig_thr || D < 1.1920928955078125e-07f) {
            if (level == 0) st = 0;
            continue;
        }
        D = __fdiv_rn(1.f, D);

        float qx = __fsub_rn(cx, half), qy = __fsub_rn(cy, half);  // nextPt (window top-left)
        float pdx = 0.f, pdy = 0.f;
        const int jstride = LJ.stride;
        int joff[NR];
#pragma unroll
        for (int r = 0; r < NR; r++) joff[r] = rrow[r] * jstride + rcol[r];
        // stage the neighbourhood of the starting position (padded coordinates, x origin word aligned)
        const int X0 = __float2int_rd(qx) + kPad, Y0 = __float2int_rd(qy) + kPad;
        const int tx0 = tma ? ((X0 - kOx0T) & ~15) : ((X0 - TL::kOx0 + 3) & ~3), ty0 = Y0 - TL::kOy0;
        const bool tile_ok = use_tile && max_iters > 0 && X0 >= 0 && X0 < jstride && tx0 >= 0 &&
                             tx0 + 4 * (tma ? kWordsT : TL::kWords) <= jstride && ty0 >= 0 && ty0 + TL::kRows <= LJ.h + 2 * kPad;
        if (tile_ok && tma) {
            // every lane's generic-proxy accesses to the tile (the set-up staging above, the previous level's samples)
            // are ordered before the async-proxy write that is about to land in the same bytes
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                mbar_expect_tx(bar, TL::kRows * kWordsT * 4);
                tma_load_2d(tile, &maps.m[level], bar, tx0, ty0);
            }
            mbar_wait(bar, phase);
            phase ^= 1;
            __syncwarp();
        } else if (tile_ok) {
            const unsigned* src = reinterpret_cast<const unsigned*>(LJ.img + (size_t)ty0 * jstride + tx0);
            const int wstride = jstride >> 2;
            __syncwarp();   // the previous level's readers are done
#pragma unroll
            for (int i = 0; i < (TL::kRows * TL::kWords + 31) / 32; i++) {
                const int idx = i * 32 + lane;
                const int row = idx / TL::kWords, col = idx - row * TL::kWords;
                if (idx < TL::kRows * TL::kWords) tile[row * TL::kPitch + col] = __ldg(src + (size_t)row * wstride + col);
            }
            __syncwarp();
        }
        for (int j = 0; j < max_iters; j++) {
            const int jx = __float2int_rd(qx), jy = __float2int_rd(qy);
            if (jx < -WIN || jx >= LJ.w || jy < -WIN || jy >= LJ.h) {
                if (level == 0) st = 0;
                break;
            }
            iters++;
            w = make_weights(__fsub_rn(qx, (float)jx), __fsub_rn(qy, (float)jy));
            const int wtop = pack_w(w.w00, w.w01), wbot = pack_w(w.w10, w.w11);
            int b1 = 0, b2 = 0;
            const int ox = jx + kPad - tx0, oy = jy + kPad - ty0;
            if (tile_ok && (unsigned)ox <= (unsigned)oxmax && (unsigned)oy <= (unsigned)TL::kOyMax) {
#pragma unroll
                for (int r = 0; r < NR; r++) {
                    int v[kRun];
                    sample_run_s(tile + (oy + rrow[r]) * jpitch, jpitch, ox + rcol[r], wtop, wbot, v);
#pragma unroll
                    for (int x = 0; x < kRun; x++) {
                        const int diff = v[x] - Iw[r][x];
                        b1 += diff * dIx[r][x];
                        b2 += diff * dIy[r][x];
                    }
                }
            } else {
                const uint8_t* jb = LJ.img + (size_t)(jy + kPad) * jstride + (jx + kPad);
#pragma unroll
                for (int r = 0; r < NR; r++) {
                    int v[kRun];
                    sample_run(jb + joff[r], jstride, wtop, wbot, v);
#pragma unroll
                    for (int x = 0; x < kRun; x++) {
                        const int diff = v[x] - Iw[r][x];
                        b1 += diff * dIx[r][x];
                        b2 += diff * dIy[r][x];
                    }
                }
            }
            const float fb1 = __fmul_rn(__ll2float_rn(warp_sum_i64(b1)), FLT_SCALE);
            const float fb2 = __fmul_rn(__ll2float_rn(warp_sum_i64(b2)), FLT_SCALE);
            const float dx = __fmul_rn(__fsub_rn(__fmul_rn(A12, fb2), __fmul_rn(A22, fb1)), D);
            const float dy = __fmul_rn(__fsub_rn(__fmul_rn(A12, fb1), __fmul_rn(A11, fb2)), D);
            qx = __fadd_rn(qx, dx); qy = __fadd_rn(qy, dy);
            cx = __fadd_rn(qx, half); cy = __fadd_rn(qy, half);
            if ((double)dx * dx + (double)dy * dy <= eps2) break;
            if (j > 0 && fabsf(__fadd_rn(dx, pdx)) < 0.01 && fabsf(__fadd_rn(dy, pdy)) < 0.01) {
                cx = __fsub_rn(cx, __fmul_rn(dx, 0.5f));
                cy = __fsub_rn(cy, __fmul_rn(dy, 0.5f));
                break;
            }
            pdx = dx; pdy = dy;
        }

        if (level == 0 && st) {
            const float ex = __fsub_rn(cx, half), ey = __fsub_rn(cy, half);
            const int jx = __float2int_rd(ex), jy = __float2int_rd(ey);
            if (jx < -WIN || jx >= LJ.w || jy < -WIN || jy >= LJ.h) {
                st = 0;
            } else {
                w = make_weights(__fsub_rn(ex, (float)jx), __fsub_rn(ey, (float)jy));
                const uint8_t* jb = LJ.img + (size_t)(jy + kPad) * jstride + (jx + kPad);
                int e = 0;
#pragma unroll
                for (int r = 0; r < NR; r++) {
                    int v[kRun];
                    sample_run(jb + joff[r], jstride, pack_w(w.w00, w.w01), pack_w(w.w10, w.w11), v);
                    if (rvalid[r]) {
#pragma unroll
                        for (int x = 0; x < kRun; x++) e += abs(v[x] - Iw[r][x]);
                    }
                }
                e = __reduce_add_sync(0xffffffffu, e);
                err = __fmul_rn((float)e, 1.f / (32 * WIN * WIN));
            }
        }
    }
    if (lane == 0) {
        out[track] = make_float2(cx, cy);
        status_out[track] = (uint8_t)st;
        if (err_out) err_out[track] = err;
        if (iters_out) iters_out[track] = iters;
    }
}

__global__ void fb_compare_kernel(const float2* __restrict__ pts, const float2* __restrict__ back,
                                  const uint8_t* __restrict__ back_status, const int32_t* __restrict__ n_dev,
                                  int n_upper, float thr2, uint8_t* __restrict__ status, const int32_t* __restrict__ gate)
{
    if (gate && !*gate) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = n_dev ? min(*n_dev, n_upper) : n_upper;
    if (i >= n) return;
    const float dx = back[i].x - pts[i].x, dy = back[i].y - pts[i].y;
    if (!back_status[i] || dx * dx + dy * dy > thr2) status[i] = 0;
}

}  // namespace

// Forward-backward check (BASELINE.json north_star; the reference has none, so it is off by default): the
// tracked positions are tracked back into the previous frame and a track keeps its status only if the way
// back succeeds and ends within fb_thresh px of where it started.  `which` selects the scratch (0: main
// stream / landmarks, 1: side stream / candidates).
int klt_fb_check(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const float* d_pts,
                 const int32_t* d_n, int n_upper, const float* d_out, uint8_t* d_status,
                 const vo_klt_params& prm, cudaStream_t st, int which, const int32_t* gate)
{
    if (n_upper <= 0) return VO_OK;
    if (!st) st = c->stream;
    DevBuf<float>& bp = which ? c->fb_pts2 : c->fb_pts;
    DevBuf<uint8_t>& bs = which ? c->fb_status2 : c->fb_status;
    VO_TRY(ensure(c, bp, (size_t)n_upper * 2));
    VO_TRY(ensure(c, bs, (size_t)n_upper));
    VO_TRY(launch_klt2(c, next, prev, d_out, d_n, n_upper, nullptr, nullptr, 0, 0, bp.p, bs.p, nullptr, nullptr, prm, st, gate));
    fb_compare_kernel<<<(n_upper + 255) / 256, 256, 0, st>>>(
        (const float2*)d_pts, (const float2*)bp.p, bs.p, d_n, n_upper,
        (float)(prm.fb_thresh * prm.fb_thresh), d_status, gate);
    c->launches++;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

int launch_klt(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const float* d_pts,
               const int32_t* d_n, int n_upper, float* d_out, uint8_t* d_status, float* d_err,
               int32_t* d_iters, const vo_klt_params& prm)
{
    return launch_klt2(c, prev, next, d_pts, d_n, n_upper, nullptr, nullptr, 0, 0, d_out, d_status, d_err,
                       d_iters, prm);
}

int launch_klt2(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const float* d_pts,
                const int32_t* d_n, int n_upper, const float* d_ptsB, const int32_t* d_nB, int nB_upper,
                int b_off, float* d_out, uint8_t* d_status, float* d_err, int32_t* d_iters,
                const vo_klt_params& prm, cudaStream_t st, const int32_t* gate)
{
    if (n_upper < 0) n_upper = 0;
    if (nB_upper < 0) nB_upper = 0;
    KltSegs segs{};
    segs.n = 2;
    segs.pts[0] = (const float2*)d_pts; segs.n_dev[0] = d_n; segs.n_upper[0] = n_upper; segs.out_off[0] = 0;
    segs.pts[1] = (const float2*)d_ptsB; segs.n_dev[1] = d_nB; segs.n_upper[1] = nB_upper; segs.out_off[1] = b_off;
    return launch_klt_segs(c, prev, next, segs, d_out, d_status, d_err, d_iters, prm, st, gate);
}

int launch_klt_segs(Ctx* c, const PyramidDesc& prev, const PyramidDesc& next, const KltSegs& segs, float* d_out,
                    uint8_t* d_status, float* d_err, int32_t* d_iters, const vo_klt_params& prm, cudaStream_t st,
                    const int32_t* gate)
{
    VO_NVTX("vo:klt_track");
    if (!st) st = c->stream;
    int total = 0;
    for (int k = 0; k < segs.n; k++) total += segs.n_upper[k];
    if (total <= 0) return VO_OK;
    if (prm.win != 21) return set_error(c, VO_ERR_INVALID, "KLT window %d not built (only 21)", prm.win);
    if (prev.nlevels != next.nlevels || prev.lv[0].w != next.lv[0].w || prev.lv[0].h != next.lv[0].h)
        return set_error(c, VO_ERR_INVALID, "KLT frames differ in size");
    // effective top level, as buildOpticalFlowPyramid: stop before a level that is <= winSize
    int top = 0;
    for (int l = 1; l <= prm.max_level && l < prev.nlevels; l++) {
        if (prev.lv[l].w <= prm.win || prev.lv[l].h <= prm.win) break;
        top = l;
    }
    int max_iters = prm.max_iters < 0 ? 0 : (prm.max_iters > 100 ? 100 : prm.max_iters);
    const double eps = prm.epsilon < 0. ? 0. : (prm.epsilon > 10. ? 10. : prm.epsilon);
    const double eps2 = eps * eps;
    static const bool no_tile = getenv("VO_KLT_NO_TILE") != nullptr;   // A/B switches (same results either way)
    const char* e_tma = getenv("VO_KLT_TMA");                         // read per launch (tests toggle it)
    const bool want_tma = e_tma && e_tma[0] == '1' && next.tmaps != nullptr && top <= 3;
    KltTmaps maps;
    if (want_tma) memcpy(&maps, next.tmaps, sizeof(maps)); else memset(&maps, 0, sizeof(maps));
    const int blocks = (total + kWarpsPerBlock - 1) / kWarpsPerBlock;
    // CTAs per SM the register allocation aims at: 4 (128 registers), 5 (96, no spill) or 6 (80, 40 B of spill);
    // VO_KLT_MINB: A/B switch (same results)
    static const int minb = getenv("VO_KLT_MINB") ? atoi(getenv("VO_KLT_MINB")) : 4;
    const int tile_mode = no_tile ? 0 : (want_tma ? 2 : 1);
    // (programmatic dependent launch: the CTAs are made resident while the kernel in front - the frame tail, one CTA - is
    // still running, and start the moment it ends; VO_KLT_NO_PDL=1: plain launch, A/B switch)
    static const bool no_pdl = getenv("VO_KLT_NO_PDL") != nullptr;
#define VO_KLT_LAUNCH(MB, TM) do { if (no_pdl) klt_track_kernel<21, MB, TM><<<blocks, kWarpsPerBlock * 32, 0, st>>>( \
        prev, next, maps, top, segs, (float2*)d_out, d_status, d_err, d_iters, max_iters, eps2, prm.min_eig_thr, gate); \
    else VO_CUDA(c, launch_pdl(klt_track_kernel<21, MB, TM>, dim3(blocks), dim3(kWarpsPerBlock * 32), st, prev, next, maps, top, segs, \
        (float2*)d_out, d_status, d_err, d_iters, max_iters, eps2, prm.min_eig_thr, gate)); } while (0)
    if (tile_mode == 0) VO_KLT_LAUNCH(4, 0);
    else if (tile_mode == 2) VO_KLT_LAUNCH(4, 2);
    else if (minb == 5) VO_KLT_LAUNCH(5, 1);
    else if (minb == 6) VO_KLT_LAUNCH(6, 1);
    else VO_KLT_LAUNCH(4, 1);
#undef VO_KLT_LAUNCH
    c->launches++;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

}  // namespace vo
