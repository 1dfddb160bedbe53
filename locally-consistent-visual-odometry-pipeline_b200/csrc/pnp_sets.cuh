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
    while (h < to && !exhausted) {
        const int cnt = min(32, to - h);
        bool ok = false;
        int4 v = make_int4(0, 0, 0, 0);
        const long long p0 = pos + 4 * lane;
        if (lane < cnt && p0 + 3 < rt.len) {
            v.x = (int)(rt.tab[p0] % (unsigned)n);
            v.y = (int)(rt.tab[p0 + 1] % (unsigned)n);
            v.z = (int)(rt.tab[p0 + 2] % (unsigned)n);
            v.w = (int)(rt.tab[p0 + 3] % (unsigned)n);
            ok = v.y != v.x && v.z != v.x && v.z != v.y && v.w != v.x && v.w != v.y && v.w != v.z;
        }
        const unsigned bad = __ballot_sync(0xffffffffu, lane < cnt && !ok);
        const int good = bad ? __ffs(bad) - 1 : cnt;
        if (lane < good) reinterpret_cast<int4*>(sets)[h + lane] = v;
        h += good; pos += 4 * good;
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
    }
    __syncwarp();
    if (lane == 0) {
        sc->rng = (uint64_t)pos;
        sc->sets_ready = exhausted ? to : h;   // rows past h stay unset; rng_exhausted aborts the run
        if (exhausted) sc->rng_exhausted = 1;
    }
}

// Called by ONE full warp of the kernel that fixes the point count n of the next RANSAC: resets the
// draw position and pre-generates the minimal sets of the first chunk, so the RANSAC needs no
// set-up launch of its own.
__device__ inline void pnp_prepare_sets(DevScalars* sc, int32_t* sets, RngTab rt, int n)
{
    __syncwarp();
    if ((threadIdx.x & 31) == 0) {
        sc->sets_ready = 0; sc->rng = 0; sc->rng_exhausted = 0;
        if (n == kMinimal) {   // cv2: count == modelPoints -> one direct solve on all points
            reinterpret_cast<int4*>(sets)[0] = make_int4(0, 1, 2, 3);
            sc->sets_ready = 0x7fffffff;
        }
    }
    __syncwarp();
    if (n > kMinimal && rt.tab) gen_sets_warp(sc, sets, rt, n, kPnpChunk0);
}

}  // namespace vo
