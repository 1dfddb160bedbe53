// Minimal-set generation for the PnP-RANSAC (cv::RNG emulation), shared by the kernels that learn
// the point count (post-KLT compaction, point packing) and by the RANSAC kernels themselves.
#pragma once
#include "ctx.cuh"

namespace vo {

constexpr int kPnpChunk0 = 128;   // hypotheses of the first RANSAC chunk (sets pre-generated)

// ---- cv::RNG -----------------------------------------------------------------------------------
// cv::RNG is a multiply-with-carry generator: strictly sequential, but its raw 32-bit output
// stream does not depend on the data (the seed is fixed at 2^64-1 inside RANSAC's run()).  The
// host tabulates the stream once (rng_table_host); the device turns it into minimal sets with
// `% n` and the duplicate re-draws of getSubset, 32 sets per step.
struct RngTab { const uint32_t* tab; int len; };

// rows [sc->sets_ready, to) of the minimal-set table, exactly as getSubset draws them.
// Called by one full warp.  Optimistic step: lane l assumes sets h..h+l-1 needed no re-draw and
// reads draws pos+4l..pos+4l+3; the prefix of lanes for which that held is committed, the first
// set with a duplicate is drawn serially by lane 0.
__device__ inline void gen_sets_warp(DevScalars* sc, int32_t* sets, RngTab rt, int n, int to)
{
    const int lane = threadIdx.x & 31;
    int h = sc->sets_ready;
    long long pos = (long long)sc->rng;
    int exhausted = 0;
    // the raw draws of the NEXT step are requested before this step's are looked at (as if no set of this step
    // needed a re-draw, which is the rule): the table is cold in L1 and a step is otherwise one L2 round trip long
    auto fetch = [&](long long p0, uint32_t* d) {
        const bool in = p0 + 3 < rt.len;
#pragma unroll
        for (int k = 0; k < 4; k++) d[k] = in ? __ldg(rt.tab + p0 + k) : 0u;
        return in;
    };
    uint32_t cur[4], nxt[4];
    bool cur_in = fetch(pos + 4 * lane, cur);
    while (h < to && !exhausted) {
        const int cnt = min(32, to - h);
        const bool nxt_in = fetch(pos + 128 + 4 * lane, nxt);
        bool ok = false;
        int4 v = make_int4(0, 0, 0, 0);
        if (lane < cnt && cur_in) {
            v.x = (int)(cur[0] % (unsigned)n);
            v.y = (int)(cur[1] % (unsigned)n);
            v.z = (int)(cur[2] % (unsigned)n);
            v.w = (int)(cur[3] % (unsigned)n);
            ok = v.y != v.x && v.z != v.x && v.z != v.y && v.w != v.x && v.w != v.y && v.w != v.z;
        }
        const unsigned bad = __ballot_sync(0xffffffffu, lane < cnt && !ok);
        const int good = bad ? __ffs(bad) - 1 : cnt;
        if (lane < good) reinterpret_cast<int4*>(sets)[h + lane] = v;
        h += good; pos += 4 * good;
        if (good == 32) {   // the usual case: the prefetched draws are the next step's
#pragma unroll
            for (int k = 0; k < 4; k++) cur[k] = nxt[k];
            cur_in = nxt_in;
            continue;
        }
        if (good < cnt) {   // set h: serial draw with re-draws (or the table ran out)
            long long np = pos;
            int fail = 0;
            if (lane == 0) {
                int idx[kMinimal];
#pragma unroll
                for (int i = 0; i < kMinimal; i++) {
                    for (;;) {
                        if (np >= rt.len) { fail = 1; break; }
                        const int x = (int)(rt.tab[np++] % (unsigned)n);
                        bool dup = false;
#pragma unroll
                        for (int j = 0; j < kMinimal; j++) dup |= (j < i) && idx[j] == x;
                        if (!dup) { idx[i] = x; break; }
                    }
                    if (fail) break;
                }
                if (!fail) reinterpret_cast<int4*>(sets)[h] = make_int4(idx[0], idx[1], idx[2], idx[3]);
            }
            exhausted = __shfl_sync(0xffffffffu, fail, 0);
            pos = __shfl_sync(0xffffffffu, np, 0);
            if (!exhausted) h++;
        }
        if (h < to && !exhausted) cur_in = fetch(pos + 4 * lane, cur);   // the draw position moved: fetch again
    }
    __syncwarp();
    if (lane == 0) {
        sc->rng = (uint64_t)pos;
        sc->sets_ready = exhausted ? to : h;   // rows past h stay unset; rng_exhausted aborts the run
        if (exhausted) sc->rng_exhausted = 1;
    }
}

// The first kPnpChunk0 sets start at draw 0, so their raw draws can be fetched before the point count
// is known (the table is cold in L2 after anything large has run): lane l, batch b holds draws
// 128 b + 4 l .. + 3.
struct RngPreload { uint4 v[kPnpChunk0 / 32]; };

__device__ inline void pnp_preload_draws(RngTab rt, RngPreload& pre)
{
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int b = 0; b < kPnpChunk0 / 32; b++) {
        pre.v[b] = make_uint4(0, 0, 0, 0);
        if (rt.tab && 128 * b + 4 * lane + 3 < rt.len) pre.v[b] = reinterpret_cast<const uint4*>(rt.tab)[32 * b + lane];
    }
}

// Called by ONE full warp of the kernel that fixes the point count n of the next RANSAC: resets the
// draw position and pre-generates the minimal sets of the first chunk, so the RANSAC needs no
// set-up launch of its own.  Batches are served from the preloaded draws as long as no set needed
// a re-draw (then the draw position is no longer 128 b and the generic path takes over).
__device__ inline void pnp_prepare_sets(DevScalars* sc, int32_t* sets, RngTab rt, int n, const RngPreload& pre)
{
    const int lane = threadIdx.x & 31;
    __syncwarp();
    int h = 0;
    if (n > kMinimal && rt.tab && rt.len >= 4 * kPnpChunk0) {
#pragma unroll
        for (int b = 0; b < kPnpChunk0 / 32; b++) {
            if (h != 32 * b) break;
            const uint4 r = pre.v[b];
            const int4 v = make_int4((int)(r.x % (unsigned)n), (int)(r.y % (unsigned)n), (int)(r.z % (unsigned)n),
                                     (int)(r.w % (unsigned)n));
            const bool ok = v.y != v.x && v.z != v.x && v.z != v.y && v.w != v.x && v.w != v.y && v.w != v.z;
            const unsigned bad = __ballot_sync(0xffffffffu, !ok);
            const int good = bad ? __ffs(bad) - 1 : 32;
            if (lane < good) reinterpret_cast<int4*>(sets)[h + lane] = v;
            h += good;
        }
    }
    if (lane == 0) {
        sc->sets_ready = h; sc->rng = (uint64_t)(4 * h); sc->rng_exhausted = 0;
        if (n == kMinimal) {   // cv2: count == modelPoints -> one direct solve on all points
            reinterpret_cast<int4*>(sets)[0] = make_int4(0, 1, 2, 3);
            sc->sets_ready = 0x7fffffff;
        }
    }
    __syncwarp();
    if (n > kMinimal && rt.tab && h < kPnpChunk0) gen_sets_warp(sc, sets, rt, n, kPnpChunk0);
}

}  // namespace vo
