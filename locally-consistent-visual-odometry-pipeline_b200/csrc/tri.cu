// K7/K8: state compaction after tracking, the bearing-angle baseline test fused with batched DLT
// triangulation and the C -> P/X promotion, candidate append.
//
// Replaces, in /root/reference/src/vo_pipeline.py:
//   :217-225  boolean-mask selection of tracked landmarks (associateKeypointsToLandmarks)
//   :503-545  update of tracked candidates, removal of lost rows of C/F/T
//   :564-591  append of new candidates (F := C, T := new pose)
//   :608-642  bearing-angle test (Python loop over all candidates, raw pixel bearing vectors)
//   :658-674  cv2.triangulatePoints per validated candidate + np.concatenate onto P and X
//   :718-721  np.delete of the validated rows
// cv2.triangulatePoints behaviour per SURVEY.md App. A.3: unnormalised 4x4 DLT, right singular
// vector of the smallest singular value; here one-sided Jacobi in FP64 (FP32 misses the 1e-4 gate).
// All compactions are order-preserving (prefix sums), so row order equals the reference's.
#include "ctx.cuh"
#include "dlt.cuh"
#include "pnp_sets.cuh"

namespace vo {

namespace {

constexpr int kScanThreads = 1024;

// block-wide exclusive rank of `flag` among the CTA's threads for the current tile; `base` is the
// running total (shared), advanced by the tile's total.  All threads must call.
template <int NT>
__device__ __forceinline__ int tile_rank(bool flag, int* sWarp, int* sBase)
{
    const int tid = threadIdx.x;
    const unsigned bal = __ballot_sync(0xffffffffu, flag);
    const int wpre = __popc(bal & ((1u << (tid & 31)) - 1));
    if ((tid & 31) == 0) sWarp[tid >> 5] = __popc(bal);
    __syncthreads();
    int off = *sBase;
    for (int w = 0; w < (tid >> 5); w++) off += sWarp[w];
    __syncthreads();
    if (tid == 0) {
        int s = 0;
        for (int w = 0; w < NT / 32; w++) s += sWarp[w];
        *sBase += s;
    }
    __syncthreads();
    return off + wpre;
}

// after KLT on the landmarks: keep status==1 rows (new position, X), pack the float32 PnP inputs.
// One CTA, one barrier: 31 warps own contiguous segments of the rows (count, barrier, then
// order-preserving writes from warp-local ranks); the 32nd warp, which after the barrier knows the
// point count of the coming PnP-RANSAC, draws its first minimal sets meanwhile.
__global__ void __launch_bounds__(kScanThreads)
landmark_compact_kernel(DevScalars* sc, int n_upper, const float2* __restrict__ klt_out,
                        const uint8_t* __restrict__ status, const int32_t* __restrict__ iters,
                        const double* __restrict__ X_in, float2* __restrict__ P_out, double* __restrict__ X_out,
                        float4* __restrict__ q4, float* __restrict__ qv, float2* __restrict__ lost,
                        int32_t* __restrict__ sets, RngTab rt, const int32_t* __restrict__ gate,
                        const int32_t* __restrict__ src)
{
    constexpr int NW = kScanThreads / 32 - 1;   // data warps
    __shared__ int sKeepW[32], sLostW[32], sIterW[32];
    pdl_prologue();
    if (gate && !*gate) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(sc->n_land, n_upper);
    const int seg = ((n + NW - 1) / NW + 31) & ~31;   // rows per data warp, a multiple of the warp width
    const int lo = warp * seg, hi = min(n, lo + seg);
    int ck = 0, cl = 0, it = 0;
    RngPreload pre;
    if (warp == NW) pnp_preload_draws(rt, pre);   // in flight while the data warps count
    if (warp < NW) {
        for (int i0 = lo; i0 < hi; i0 += 32) {
            const int i = i0 + lane;
            const bool in = i < hi;
            const int t = in ? (src ? src[i] : i) : 0;   // the row's track (tracking ahead of the pruning: a superset)
            const bool keep = in && status[t] != 0;
            if (in && iters) it += iters[t];
            const unsigned bk = __ballot_sync(0xffffffffu, keep);
            const unsigned bi = __ballot_sync(0xffffffffu, in);
            ck += __popc(bk); cl += __popc(bi & ~bk);
        }
        it = __reduce_add_sync(0xffffffffu, it);
    }
    if (lane == 0) { sKeepW[warp] = ck; sLostW[warp] = cl; sIterW[warp] = it; }
    __syncthreads();
    // every warp: exclusive scan of the per-warp counts across lanes
    const int kw = sKeepW[lane], lw = sLostW[lane];
    int ks = kw, ls = lw;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int a = __shfl_up_sync(0xffffffffu, ks, o), b = __shfl_up_sync(0xffffffffu, ls, o);
        if (lane >= o) { ks += a; ls += b; }
    }
    const int n_keep = __shfl_sync(0xffffffffu, ks, 31), n_lost = __shfl_sync(0xffffffffu, ls, 31);
    int ok_ = __shfl_sync(0xffffffffu, ks - kw, warp), ol_ = __shfl_sync(0xffffffffu, ls - lw, warp);
    if (warp == NW) {
        // the point count of the coming PnP-RANSAC is known: draw its first minimal sets here
        pnp_prepare_sets(sc, sets, rt, n_keep, pre);
        return;
    }
    for (int i0 = lo; i0 < hi; i0 += 32) {
        const int i = i0 + lane;
        const bool in = i < hi;
        const int t = in ? (src ? src[i] : i) : 0;
        const bool keep = in && status[t] != 0;
        const unsigned bk = __ballot_sync(0xffffffffu, keep);
        const unsigned bi = __ballot_sync(0xffffffffu, in);
        const unsigned below = (1u << lane) - 1;
        if (keep) {
            const int o = ok_ + __popc(bk & below);
            const float2 p = klt_out[t];
            const double x = X_in[3 * (size_t)i], y = X_in[3 * (size_t)i + 1], z = X_in[3 * (size_t)i + 2];
            P_out[o] = p;
            X_out[3 * (size_t)o] = x; X_out[3 * (size_t)o + 1] = y; X_out[3 * (size_t)o + 2] = z;
            q4[o] = make_float4((float)x, (float)y, (float)z, p.x);   // cv2 converts X to float32
            qv[o] = p.y;
        } else if (in) {
            lost[ol_ + __popc(bi & ~bk & below)] = klt_out[t];
        }
        ok_ += __popc(bk); ol_ += __popc(bi & ~bk);
    }
    if (warp == 0) {
        int its = sIterW[lane];
        its = __reduce_add_sync(0xffffffffu, its);
        if (lane == 0) {
            sc->klt_iters_cand = 0; sc->prev2_n_land = sc->prev_n_land; sc->prev_n_land = n; sc->n_land = n_keep; sc->n_tracked = n_keep; sc->n_lost = n_lost;
            sc->klt_iters = its;
        }
    }
}

// M = K(3x3) @ T[0:3,:] (T 4x4 row-major)
__device__ __forceinline__ void k_times_pose(const double* K, const double* T, double* M)
{
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) M[4 * r + c] = K[3 * r] * T[c] + K[3 * r + 1] * T[4 + c] + K[3 * r + 2] * T[8 + c];
}

// vo_pipeline.py:621-642 evaluated as written (float64): raw-pixel bearing vectors, normalised,
// cos clipped, arccos, degrees
__device__ __forceinline__ double bearing_angle_deg(double uc, double vc, double uf, double vf)
{
    const double wc = sqrt(uc * uc + vc * vc + 1.), wf = sqrt(uf * uf + vf * vf + 1.);
    const double a0 = uc / wc, a1 = vc / wc, a2 = 1. / wc, b0 = uf / wf, b1 = vf / wf, b2 = 1. / wf;
    const double na = sqrt(a0 * a0 + a1 * a1 + a2 * a2), nb = sqrt(b0 * b0 + b1 * b1 + b2 * b2);
    double cs = (a0 * b0 + a1 * b1 + a2 * b2) / (na * nb);
    cs = fmin(fmax(cs, -1.), 1.);
    return acos(cs) * (180. / 3.14159265358979323846);
}

struct PromoteArgs {
    double K[9];
    double Tnew[16];
    double alpha_deg;
    int use_dev_T;     // take the new pose from sc->T (fused per-frame call) instead of Tnew
    int use_klt;       // candidates were tracked: positions/status at klt_out/klt_status + b_off
    int b_off;
    int ncand_upper;
};

// Promotion step, two launches of ceil(n/256) CTAs (a single CTA was the longest serial piece of
// the frame after RANSAC):
//   cand_classify_kernel  per candidate: lost / stays / promoted (bearing test) + per-CTA counts
//   cand_apply_kernel     per-CTA offsets from the counts (order-preserving), C := new position
//                         (F too while C/F alias, App. B-5), DLT with (T_first, T_new) for the
//                         promoted rows appended to P/X, the rest compacted into the other
//                         candidate buffers; the last CTA to finish publishes the new sizes.
constexpr int kCandThreads = 256;
enum : uint8_t { kCandLost = 0, kCandStay = 1, kCandPromote = 2, kCandNone = 255 };

struct CandRow { float2 c; double2 f; bool keep; };

__device__ __forceinline__ CandRow load_cand(const PromoteArgs& a, bool alias, int i, const float2* __restrict__ klt_out,
                                             const uint8_t* __restrict__ status, const float2* __restrict__ C_in,
                                             const double2* __restrict__ F_in)
{
    CandRow r;
    r.keep = true;
    if (a.use_klt) { r.keep = status[a.b_off + i] != 0; r.c = klt_out[a.b_off + i]; }
    else r.c = C_in[i];
    r.f = (alias && a.use_klt) ? make_double2((double)r.c.x, (double)r.c.y) : F_in[i];
    return r;
}

__global__ void __launch_bounds__(kCandThreads)
cand_classify_kernel(DevScalars* sc, PromoteArgs a, const float2* __restrict__ klt_out,
                     const uint8_t* __restrict__ status, const float2* __restrict__ C_in,
                     const double2* __restrict__ F_in, uint8_t* __restrict__ flag, int32_t* __restrict__ blk_counts,
                     const int32_t* __restrict__ klt_iters, int early, const int32_t* __restrict__ gate)
{
    pdl_prologue();
    if (gate && !*gate) return;
    const int n = min(sc->n_cand, a.ncand_upper);
    const bool alias = sc->cf_alias != 0;
    // early (side stream, before the PnP): a pose is assumed; the host takes the step back if none comes
    const bool have_pose = early ? true : (a.use_dev_T ? sc->ok != 0 : true);
    const int i = blockIdx.x * kCandThreads + threadIdx.x;
    uint8_t f = kCandNone;
    if (i < n) {
        const CandRow r = load_cand(a, alias, i, klt_out, status, C_in, F_in);
        f = kCandLost;
        if (r.keep) {
            f = kCandStay;
            if (have_pose && bearing_angle_deg((double)r.c.x, (double)r.c.y, r.f.x, r.f.y) > a.alpha_deg) f = kCandPromote;
        }
        flag[i] = f;
    }
    if (a.use_klt && klt_iters) {   // roofline accounting: iterations spent on the candidates
        int it = i < n ? klt_iters[a.b_off + i] : 0;
        it = __reduce_add_sync(0xffffffffu, it);
        if ((threadIdx.x & 31) == 0 && it) atomicAdd(early ? &sc->klt_iters_cand2 : &sc->klt_iters_cand, it);
    }
    const int n_stay = __syncthreads_count(f == kCandStay);
    const int n_prom = __syncthreads_count(f == kCandPromote);
    if (threadIdx.x == 0) { blk_counts[2 * blockIdx.x] = n_stay; blk_counts[2 * blockIdx.x + 1] = n_prom; }
}

__global__ void __launch_bounds__(kCandThreads)
cand_apply_kernel(DevScalars* sc, PromoteArgs a, const float2* __restrict__ klt_out, const uint8_t* __restrict__ status,
                  const float2* __restrict__ C_in, const double2* __restrict__ F_in, const int32_t* __restrict__ Ti_in,
                  const uint8_t* __restrict__ flag, const int32_t* __restrict__ blk_counts,
                  float2* __restrict__ C_out, double2* __restrict__ F_out, int32_t* __restrict__ Ti_out,
                  const double* __restrict__ poses, float2* __restrict__ P, double* __restrict__ X, int land_cap,
                  int early, double* __restrict__ prom_rows, int32_t* __restrict__ prom_ti,
                  const int32_t* __restrict__ gate, float2* __restrict__ prom_P)
{
    __shared__ int sWarpStay[kCandThreads / 32], sWarpProm[kCandThreads / 32];
    __shared__ int sBaseStay, sBaseProm, sLast;
    __shared__ double sM2[12];
    pdl_prologue();
    if (gate && !*gate) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(sc->n_cand, a.ncand_upper);
    const int n_land = sc->n_land;
    const bool alias = sc->cf_alias != 0;
    if (warp == 0) {   // offsets of this CTA = counts of the CTAs before it
        int st = 0, pr = 0;
        for (int b = lane; b < (int)blockIdx.x; b += 32) { st += blk_counts[2 * b]; pr += blk_counts[2 * b + 1]; }
        st = __reduce_add_sync(0xffffffffu, st);
        pr = __reduce_add_sync(0xffffffffu, pr);
        if (lane == 0) { sBaseStay = st; sBaseProm = pr; }
    } else if (warp == 1 && lane == 0 && !early) {
        k_times_pose(a.K, a.use_dev_T ? sc->T : a.Tnew, sM2);
    }
    const int i = blockIdx.x * kCandThreads + tid;
    const uint8_t f = i < n ? flag[i] : (uint8_t)kCandNone;
    const unsigned bs = __ballot_sync(0xffffffffu, f == kCandStay);
    const unsigned bp = __ballot_sync(0xffffffffu, f == kCandPromote);
    if (lane == 0) { sWarpStay[warp] = __popc(bs); sWarpProm[warp] = __popc(bp); }
    __syncthreads();
    int os = sBaseStay, op = sBaseProm;
    for (int w = 0; w < warp; w++) { os += sWarpStay[w]; op += sWarpProm[w]; }
    const unsigned below = (1u << lane) - 1;
    os += __popc(bs & below); op += __popc(bp & below);
    if (f == kCandStay || f == kCandPromote) {
        const CandRow r = load_cand(a, alias, i, klt_out, status, C_in, F_in);
        if (f == kCandStay) { C_out[os] = r.c; F_out[os] = r.f; Ti_out[os] = Ti_in[i]; }
        else if (early) {   // the row waits for the pose: promote_finish_kernel triangulates and appends it
            prom_rows[4 * (size_t)op] = r.f.x; prom_rows[4 * (size_t)op + 1] = r.f.y;
            prom_rows[4 * (size_t)op + 2] = (double)r.c.x; prom_rows[4 * (size_t)op + 3] = (double)r.c.y;
            prom_ti[op] = Ti_in[i];
            if (prom_P) prom_P[op] = r.c;
        }
        else if (n_land + op < land_cap) {
            double M1[12], Xh[4];
            k_times_pose(a.K, poses + 16 * (size_t)Ti_in[i], M1);
            dlt_point(M1, sM2, r.f.x, r.f.y, (double)r.c.x, (double)r.c.y, Xh);
            const int o = n_land + op;
            P[o] = r.c;
            X[3 * (size_t)o] = Xh[0] / Xh[3]; X[3 * (size_t)o + 1] = Xh[1] / Xh[3]; X[3 * (size_t)o + 2] = Xh[2] / Xh[3];
        }
    }
    // the last CTA to get here publishes the sizes: every CTA has read the old ones by then
    __syncthreads();
    int32_t* ticket = early ? &sc->ticket2 : &sc->ticket;   // early: on the side stream, under main-stream kernels
    if (tid == 0) {
        __threadfence();
        sLast = atomicAdd(ticket, 1) == (int)gridDim.x - 1;
    }
    __syncthreads();
    if (!sLast) return;
    if (warp == 0) {
        int st = 0, pr = 0;
        for (int b = lane; b < (int)gridDim.x; b += 32) { st += blk_counts[2 * b]; pr += blk_counts[2 * b + 1]; }
        st = __reduce_add_sync(0xffffffffu, st);
        pr = __reduce_add_sync(0xffffffffu, pr);
        if (lane == 0) {
            const int kept = st + pr, n_lost = n - kept;
            // (early: the landmark side - spec_n_land, n_land - belongs to the main stream's chain, which is
            // pruning the landmarks right now; promote_finish_kernel sets it)
            if (!early) sc->spec_n_land = n_land;
            sc->spec_n_cand = sc->n_cand; sc->spec_cf_alias = sc->cf_alias;
            sc->spec_cand_was_empty = sc->cand_was_empty;
            sc->n_cand = st;
            if (early) sc->n_cand_early = st;
            sc->n_promoted = pr;
            sc->n_cand_lost = n_lost;
            if (!early) sc->n_land = min(n_land + pr, land_cap);
            sc->cand_was_empty = kept == 0;                    // C empty after tracking/removal (vo:579)
            sc->cf_alias = alias && n_lost == 0 && pr == 0;    // np.delete re-allocates: aliasing ends
            *ticket = 0;
        }
    }
}

// The pose-dependent rest of an early classification: DLT of the promoted rows with (T_first, T_new),
// appended behind the live landmarks.  One CTA; each row a serial FP64 Jacobi SVD.
//   mode 0  positions, 3-D points and counters in one go (synchronous paths);
//   mode 1  positions and counters only - all the next tracking step needs (main stream, pipelined frames);
//   mode 2  the 3-D points of the rows mode 1 appended, once the refined pose is there (pose stream), and the
//           pose-table row of the frame's new candidates.
__global__ void __launch_bounds__(128)
promote_finish_kernel(DevScalars* sc, PromoteArgs a, const double* __restrict__ prom_rows,
                      const int32_t* __restrict__ prom_ti, double* __restrict__ poses, float2* __restrict__ P,
                      double* __restrict__ X, int land_cap, int mode, int pose_idx)
{
    __shared__ double sM2[12];
    pdl_prologue();
    const int tid = threadIdx.x;
    const bool ok = a.use_dev_T ? sc->ok != 0 : true;
    if (mode == 1) {
        // counters as of the pose, for the host (the block is read back while this stream goes on)
        const int n_land = sc->n_land;
        const int n_prom = ok ? sc->n_promoted : 0;
        __syncthreads();
        if (tid == 0) {
            sc->snap_n_land = n_land; sc->snap_n_cand = sc->n_cand; sc->snap_cf_alias = sc->cf_alias;
            sc->snap_n_appended = sc->n_appended;
            sc->prom_base = n_land;
            sc->spec_n_land = n_land;
            sc->n_land = min(n_land + n_prom, land_cap);
            sc->klt_iters_cand = sc->klt_iters_cand2;
            sc->klt_iters_cand2 = 0;
        }
        for (int i = tid; i < n_prom; i += blockDim.x) {
            const int o = n_land + i;
            if (o >= land_cap) break;
            P[o] = make_float2((float)prom_rows[4 * (size_t)i + 2], (float)prom_rows[4 * (size_t)i + 3]);
        }
        return;
    }
    if (mode == 2 && !ok) return;
    const int n_land = mode == 2 ? sc->prom_base : sc->n_land;
    const int n_prom = sc->n_promoted;
    if (tid == 0) k_times_pose(a.K, a.use_dev_T ? sc->T : a.Tnew, sM2);
    if (mode == 2 && pose_idx >= 0 && tid >= 32 && tid < 48) poses[16 * (size_t)pose_idx + (tid - 32)] = sc->T[tid - 32];
    __syncthreads();
    for (int i = tid; i < n_prom; i += blockDim.x) {
        const int o = n_land + i;
        if (o >= land_cap) break;
        double M1[12], Xh[4];
        k_times_pose(a.K, poses + 16 * (size_t)prom_ti[i], M1);
        const double fx = prom_rows[4 * (size_t)i], fy = prom_rows[4 * (size_t)i + 1];
        const double cx = prom_rows[4 * (size_t)i + 2], cy = prom_rows[4 * (size_t)i + 3];
        dlt_point(M1, sM2, fx, fy, cx, cy, Xh);
        if (mode == 0) P[o] = make_float2((float)cx, (float)cy);
        X[3 * (size_t)o] = Xh[0] / Xh[3]; X[3 * (size_t)o + 1] = Xh[1] / Xh[3]; X[3 * (size_t)o + 2] = Xh[2] / Xh[3];
    }
    if (mode != 0) return;
    __syncthreads();
    if (tid == 0) {
        sc->spec_n_land = n_land;
        sc->n_land = min(n_land + n_prom, land_cap);
        sc->klt_iters_cand = sc->klt_iters_cand2;
        sc->klt_iters_cand2 = 0;
    }
}

// take an early classification back: the candidate-side scalars as the KLT left them
__global__ void early_undo_kernel(DevScalars* sc)
{
    sc->n_cand = sc->spec_n_cand; sc->cf_alias = sc->spec_cf_alias; sc->cand_was_empty = sc->spec_cand_was_empty;
    sc->n_promoted = 0; sc->n_cand_lost = 0; sc->klt_iters_cand2 = 0;
}

// append n_new candidates: C := F := pts, T := pose (vo_pipeline.py:574-591); the last CTA to
// finish publishes the new size (every CTA has read the old one by then)
__global__ void cand_add_kernel(DevScalars* sc, const double2* __restrict__ pts, int n_new_upper,
                                const int32_t* __restrict__ n_new_dev, int use_dev_T,
                                const double* __restrict__ Thost, float2* __restrict__ C, double2* __restrict__ F,
                                int32_t* __restrict__ Ti, double* __restrict__ poses, int pose_idx, int cand_cap,
                                int side, int write_pose)
{
    __shared__ int sLast;
    pdl_prologue();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n0 = sc->n_cand;
    int32_t* ticket = side ? &sc->ticket2 : &sc->ticket;
    const int n_new = n_new_dev ? min(*n_new_dev, n_new_upper) : n_new_upper;
    const bool skip = use_dev_T && !sc->ok;   // no pose this frame: nothing is added
    if (!skip) {
        // (write_pose = 0: the refined pose is still being computed on the pose stream, which writes the row itself)
        if (i < 16 && write_pose) poses[16 * (size_t)pose_idx + i] = use_dev_T ? sc->T[i] : Thost[i];
        if (i < n_new && n0 + i < cand_cap) {
            const double2 p = pts[i];
            C[n0 + i] = make_float2((float)p.x, (float)p.y);
            F[n0 + i] = p;
            Ti[n0 + i] = pose_idx;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        sLast = atomicAdd(ticket, 1) == (int)gridDim.x - 1;
        if (sLast) {
            *ticket = 0;
            if (!skip && n_new > 0) {
                // C empty when the batch arrives: the reference binds C and F to one ndarray (App. B-5);
                // otherwise np.concatenate allocates fresh arrays and any aliasing ends
                sc->cf_alias = sc->cand_was_empty ? 1 : 0;
                sc->n_cand = min(n0 + n_new, cand_cap);
                sc->cand_was_empty = 0;
            }
        }
    }
}

// Append externally supplied landmarks behind P/X (reference vo_pipeline.py:318-320: the re-bootstrap
// stacks its fresh landmarks onto associations['P'] / ['X']; main.py:198-200 sets the first ones).
// target > 0: top-up - only as many of the offered rows as bring |P| to `target` (decided here, on the
// device: the host does not know |P| without a round trip).  One CTA.
__global__ void __launch_bounds__(1024)
land_append_kernel(DevScalars* sc, const float2* __restrict__ srcP, const double* __restrict__ srcX, int n_rows,
                   int target, float2* __restrict__ P, double* __restrict__ X, int cap, const int32_t* gate)
{
    pdl_prologue();
    if (gate && !*gate) return;
    const int n0 = sc->n_land;
    int n_add = target > 0 ? min(n_rows, max(target - n0, 0)) : n_rows;
    n_add = max(min(n_add, cap - n0), 0);
    for (int i = threadIdx.x; i < n_add; i += blockDim.x) {
        P[n0 + i] = srcP[i];
        const size_t o = 3 * (size_t)(n0 + i), s = 3 * (size_t)i;
        X[o] = srcX[s]; X[o + 1] = srcX[s + 1]; X[o + 2] = srcX[s + 2];
    }
    __syncthreads();
    if (threadIdx.x == 0) { sc->n_land = n0 + n_add; sc->n_appended = n_add; }
}

// Pipelined frames: everything the main stream still owes the frame once the RANSAC kernel has ended, in ONE launch
// (one CTA) instead of four - the promoted rows' positions and the counters snapshot (promote_finish_kernel mode 1),
// the append of the frame's new candidates (cand_add_kernel; their pose-table row is written on the pose stream) and
// the landmark batches (land_append_kernel), in that order.  Same results as the separate kernels.
struct TailArgs {
    const double* prom_rows; float2* P; double* X; int land_cap;
    const double2* feed; int n_new; float2* C; double2* F; int32_t* Ti; int pose_idx; int cand_cap;
    int n_segs; const float2* segP[4]; const double* segX[4]; int segN[4]; int segTarget[4];
    int32_t* src; int src_prom_base; int src_seg_base[4];   // superset track index of every appended row (Pipeline::Lsrc)
};

__global__ void __launch_bounds__(1024) frame_tail_kernel(DevScalars* sc, TailArgs a)
{
    pdl_prologue();
    const int tid = threadIdx.x;
    const bool ok = sc->ok != 0;
    const int n_land0 = sc->n_land, n_cand0 = sc->n_cand;
    const int n_prom = ok ? sc->n_promoted : 0;
    const int cf_alias0 = sc->cf_alias, was_empty0 = sc->cand_was_empty, n_app0 = sc->n_appended, iters2 = sc->klt_iters_cand2;
    __syncthreads();   // every thread has read the scalars thread 0 is about to rewrite
    // promoted rows: positions now, 3-D points when the refined pose is there (pose stream)
    int n_land = min(n_land0 + n_prom, a.land_cap);
    for (int i = tid; i < n_prom && n_land0 + i < a.land_cap; i += blockDim.x) {
        a.P[n_land0 + i] = make_float2((float)a.prom_rows[4 * (size_t)i + 2], (float)a.prom_rows[4 * (size_t)i + 3]);
        if (a.src) a.src[n_land0 + i] = a.src_prom_base + i;
    }
    // new candidates (no pose this frame: none)
    const bool add = ok && a.n_new > 0;
    if (add)
        for (int i = tid; i < a.n_new && n_cand0 + i < a.cand_cap; i += blockDim.x) {
            const double2 p = a.feed[i];
            a.C[n_cand0 + i] = make_float2((float)p.x, (float)p.y);
            a.F[n_cand0 + i] = p;
            a.Ti[n_cand0 + i] = a.pose_idx;
        }
    // landmark batches, in call order
    int n_app = n_app0;
    for (int k = 0; k < a.n_segs; k++) {
        int n_add = a.segTarget[k] > 0 ? min(a.segN[k], max(a.segTarget[k] - n_land, 0)) : a.segN[k];
        n_add = max(min(n_add, a.land_cap - n_land), 0);
        for (int i = tid; i < n_add; i += blockDim.x) {
            a.P[n_land + i] = a.segP[k][i];
            const size_t o = 3 * (size_t)(n_land + i), s = 3 * (size_t)i;
            a.X[o] = a.segX[k][s]; a.X[o + 1] = a.segX[k][s + 1]; a.X[o + 2] = a.segX[k][s + 2];
            if (a.src) a.src[n_land + i] = a.src_seg_base[k] + i;
        }
        n_land += n_add;
        n_app = n_add;
    }
    if (tid == 0) {
        sc->snap_n_land = n_land0; sc->snap_n_cand = n_cand0; sc->snap_cf_alias = cf_alias0; sc->snap_n_appended = n_app0;
        sc->prom_base = n_land0; sc->spec_n_land = n_land0;
        sc->klt_iters_cand = iters2; sc->klt_iters_cand2 = 0;
        sc->n_land = n_land; sc->n_appended = n_app;
        sc->gate_ok = ok ? 1 : 0;
        if (add) { sc->cf_alias = was_empty0 ? 1 : 0; sc->n_cand = min(n_cand0 + a.n_new, a.cand_cap); sc->cand_was_empty = 0; }
    }
}

__global__ void triangulate_kernel(const double* __restrict__ M1, const double* __restrict__ M2, int per_point,
                                   const double2* __restrict__ p1, const double2* __restrict__ p2, int n,
                                   double* __restrict__ X4)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double m1[12], m2[12], Xh[4];
    const double* a = M1 + (per_point ? 12 * (size_t)i : 0);
#pragma unroll
    for (int k = 0; k < 12; k++) { m1[k] = a[k]; m2[k] = M2[k]; }
    dlt_point(m1, m2, p1[i].x, p1[i].y, p2[i].x, p2[i].y, Xh);
#pragma unroll
    for (int k = 0; k < 4; k++) X4[4 * (size_t)i + k] = Xh[k];
}

}  // namespace

int launch_landmark_compact(Ctx* c, int n_upper, const int32_t* gate, bool from_superset)
{
    VO_NVTX("vo:landmark_compact");
    Pipeline* p = c->pipe;
    VO_TRY(join_pose_stream(c));   // the 3-D points of rows promoted by a pipelined frame come from the pose stream
    const size_t n = n_upper > 0 ? n_upper : 1;
    const int nxt = (p->cur + 1) % kLandBufs;
    VO_TRY(ensure(c, p->P[nxt], n * 2));
    VO_TRY(ensure(c, p->X[nxt], n * 3));
    VO_TRY(ensure(c, p->q4, n * 4));
    VO_TRY(ensure(c, p->qv, n));
    VO_TRY(ensure(c, p->lost, n * 2));
    VO_TRY(ensure_rng_table(c, p->hyp_capacity));
    // the tracks: this frame's own KLT launch, or (from_superset) the launch that ran ahead of the pruning on the
    // tracking stream - rows find their track through Lsrc
    const int f = p->lk_flip;
    const float* out = from_superset ? p->lk_out[f].p : p->klt_out.p;
    const uint8_t* st = from_superset ? p->lk_status[f].p : p->klt_status.p;
    const int32_t* it = from_superset ? p->lk_iters[f].p : p->klt_iters.p;
    const int32_t* src = from_superset ? p->Lsrc[p->cur].p : nullptr;
    if (from_superset) VO_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_trk, 0));
    VO_CUDA(c, launch_pdl(landmark_compact_kernel, dim3(1), dim3(kScanThreads), c->stream,
        p->d_sc, n_upper, (const float2*)out, st, it,
        (const double*)p->X[p->cur].p, (float2*)p->P[nxt].p, p->X[nxt].p, (float4*)p->q4.p, p->qv.p, (float2*)p->lost.p,
        p->sets.p, RngTab{ p->rng_tab.p, p->rng_tab_len }, gate, src));
    c->launches++;
    VO_CUDA(c, cudaEventRecord(c->ev_compact, c->stream));
    p->prev2 = p->prev == nxt ? -1 : p->prev; p->prev2_n_upper = p->prev_n_upper;
    p->prev = p->cur;
    p->prev_n_upper = n_upper;
    p->cur = nxt;
    p->b1_buf = nxt; p->b1_upper = n_upper;
    p->src_valid = false;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

int early_undo(Ctx* c)
{
    Pipeline* p = c->pipe;
    if (!p->early_valid) return VO_OK;
    p->early_valid = false;
    VO_TRY(join_side_stream(c));
    early_undo_kernel<<<1, 1, 0, c->stream>>>(p->d_sc);
    c->launches++;
    VO_CUDA(c, cudaGetLastError());
    p->ccur ^= 1;                 // the buffers the classification read from are intact
    p->mask_src_valid = false;
    // a scalar read-back in between may have tightened the bound to the classified count
    if (p->n_cand_upper < p->early_ncand_upper) p->n_cand_upper = p->early_ncand_upper;
    return VO_OK;
}

// Candidate side of the promotion step on the side stream, behind the candidates' KLT that track_both
// has just enqueued there (vo_pipeline.py:503-545 and the bearing test :621-642 need no pose): runs
// under the landmark chain and the PnP.  launch_promote finishes it when it is called with the same
// alpha, or takes it back.
int launch_promote_early(Ctx* c, double alpha_deg, int ncand_upper, int b_off, const int32_t* gate)
{
    VO_NVTX("vo:promote_early(classify)");
    Pipeline* p = c->pipe;
    if (ncand_upper <= 0) return VO_OK;
    PromoteArgs a;
    memset(&a, 0, sizeof(a));
    a.alpha_deg = alpha_deg; a.use_dev_T = 1; a.use_klt = 1; a.b_off = b_off; a.ncand_upper = ncand_upper;
    const size_t nc = ncand_upper;
    VO_TRY(ensure(c, p->C[p->ccur ^ 1], nc * 2));
    VO_TRY(ensure(c, p->F[p->ccur ^ 1], nc * 2));
    VO_TRY(ensure(c, p->Ti[p->ccur ^ 1], nc));
    VO_TRY(ensure(c, p->prom_rows, nc * 4));
    VO_TRY(ensure(c, p->prom_ti, nc));
    VO_TRY(ensure(c, p->prom_P, nc * 2));
    const int nb = (ncand_upper + kCandThreads - 1) / kCandThreads;
    VO_TRY(ensure(c, p->cand_flag, nc));
    VO_TRY(ensure(c, p->blk_counts, (size_t)nb * 2));
    VO_CUDA(c, launch_pdl(cand_classify_kernel, dim3(nb), dim3(kCandThreads), c->stream2,
        p->d_sc, a, (const float2*)p->klt_out.p, (const uint8_t*)p->klt_status.p, (const float2*)p->C[p->ccur].p,
        (const double2*)p->F[p->ccur].p, p->cand_flag.p, p->blk_counts.p, (const int32_t*)p->klt_iters.p, 1, gate));
    VO_CUDA(c, launch_pdl(cand_apply_kernel, dim3(nb), dim3(kCandThreads), c->stream2,
        p->d_sc, a, (const float2*)p->klt_out.p, (const uint8_t*)p->klt_status.p, (const float2*)p->C[p->ccur].p,
        (const double2*)p->F[p->ccur].p, (const int32_t*)p->Ti[p->ccur].p, (const uint8_t*)p->cand_flag.p,
        (const int32_t*)p->blk_counts.p, (float2*)p->C[p->ccur ^ 1].p, (double2*)p->F[p->ccur ^ 1].p,
        p->Ti[p->ccur ^ 1].p, (const double*)p->poses.p, (float2*)nullptr, (double*)nullptr, 0, 1, p->prom_rows.p,
        p->prom_ti.p, gate, (float2*)p->prom_P.p));
    c->launches += 2;
    VO_CUDA(c, cudaGetLastError());
    p->ccur ^= 1;
    p->mask_src_valid = true; p->mask_src_off = b_off;
    p->early_valid = true; p->early_alpha = alpha_deg; p->early_ncand_upper = ncand_upper;
    c->side_writes_counts = true;   // n_cand is being rewritten over there
    return VO_OK;
}

int launch_promote(Ctx* c, const double K[9], const double* Tnew, double alpha_deg, int ncand_upper, bool use_klt,
                   int b_off, bool split, int pose_idx)
{
    VO_NVTX("vo:promote");
    Pipeline* p = c->pipe;
    if (!(split && p->early_valid && use_klt && alpha_deg == p->early_alpha)) split = false;
    if (split) {   // (the pose stream is in order: no need to wait for its previous frame here)
        if (c->side_inflight) {
            c->side_inflight = false; c->side_writes_counts = false;
            VO_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
        }
    } else {
        VO_TRY(join_pose_stream(c));   // pose table rows / 3-D points a pipelined frame left to the pose stream
        VO_TRY(join_side_stream(c));   // the candidates' KLT results come from the side stream
    }
    if (p->early_valid) {
        if (use_klt && alpha_deg == p->early_alpha) {
            // the candidate side is done: triangulate the promoted rows with the pose and append them
            p->early_valid = false;
            PromoteArgs a;
            memcpy(a.K, K, sizeof(a.K));
            if (Tnew) memcpy(a.Tnew, Tnew, sizeof(a.Tnew)); else memset(a.Tnew, 0, sizeof(a.Tnew));
            a.alpha_deg = alpha_deg; a.use_dev_T = Tnew == nullptr; a.use_klt = 1; a.b_off = b_off;
            a.ncand_upper = p->early_ncand_upper;
            const size_t need = (size_t)p->n_land_upper + (size_t)p->early_ncand_upper + 1;
            VO_TRY(ensure(c, p->P[p->cur], need * 2, true));
            VO_TRY(ensure(c, p->X[p->cur], need * 3, true));
            const int land_cap = (int)(p->P[p->cur].cap / 2);
            if (!split) {
                VO_CUDA(c, launch_pdl(promote_finish_kernel, dim3(1), dim3(128), c->stream, p->d_sc, a,
                                      (const double*)p->prom_rows.p, (const int32_t*)p->prom_ti.p, p->poses.p,
                                      (float2*)p->P[p->cur].p, p->X[p->cur].p, land_cap, 0, -1));
                c->launches++;
            } else {
                // main stream: positions + counters (and the snapshot the host reads with the pose)
                VO_CUDA(c, launch_pdl(promote_finish_kernel, dim3(1), dim3(128), c->stream, p->d_sc, a,
                                      (const double*)p->prom_rows.p, (const int32_t*)p->prom_ti.p, p->poses.p,
                                      (float2*)p->P[p->cur].p, p->X[p->cur].p, land_cap, 1, -1));
                VO_CUDA(c, cudaEventRecord(c->ev_snap, c->stream));
                // pose stream, behind the refit: scalar block back to the host, then the 3-D points
                VO_CUDA(c, cudaStreamWaitEvent(c->stream4, c->ev_snap, 0));
                VO_CUDA(c, cudaMemcpyAsync(p->h_sc, p->d_sc, sizeof(DevScalars), cudaMemcpyDeviceToHost, c->stream4));
                VO_CUDA(c, cudaEventRecord(p->ev_pose, c->stream4));
                VO_CUDA(c, launch_pdl(promote_finish_kernel, dim3(1), dim3(128), c->stream4, p->d_sc, a,
                                      (const double*)p->prom_rows.p, (const int32_t*)p->prom_ti.p, p->poses.p,
                                      (float2*)p->P[p->cur].p, p->X[p->cur].p, land_cap, 2, pose_idx));
                VO_CUDA(c, cudaEventRecord(c->ev_pose_done, c->stream4));
                c->pose_inflight = true;
                c->launches += 2;
            }
            p->n_land_upper += p->early_ncand_upper;
            VO_CUDA(c, cudaGetLastError());
            return VO_OK;
        }
        VO_TRY(early_undo(c));   // other arguments than assumed: back to the tracked candidates
        if (ncand_upper < p->n_cand_upper) ncand_upper = p->n_cand_upper;   // (the caller's bound may be the classified count)
    }
    PromoteArgs a;
    memcpy(a.K, K, sizeof(a.K));
    if (Tnew) memcpy(a.Tnew, Tnew, sizeof(a.Tnew)); else memset(a.Tnew, 0, sizeof(a.Tnew));
    a.alpha_deg = alpha_deg;
    a.use_dev_T = Tnew == nullptr;
    a.use_klt = use_klt;
    a.b_off = b_off;
    a.ncand_upper = ncand_upper;
    const size_t nc = ncand_upper > 0 ? ncand_upper : 1;
    VO_TRY(ensure(c, p->C[p->ccur ^ 1], nc * 2));
    VO_TRY(ensure(c, p->F[p->ccur ^ 1], nc * 2));
    VO_TRY(ensure(c, p->Ti[p->ccur ^ 1], nc));
    // promoted rows are appended in place behind the live landmarks
    const size_t need = (size_t)p->n_land_upper + nc;
    VO_TRY(ensure(c, p->P[p->cur], need * 2, true));
    VO_TRY(ensure(c, p->X[p->cur], need * 3, true));
    const int land_cap = (int)(p->P[p->cur].cap / 2);
    const int nb = ncand_upper > 0 ? (ncand_upper + kCandThreads - 1) / kCandThreads : 1;
    VO_TRY(ensure(c, p->cand_flag, nc));
    VO_TRY(ensure(c, p->blk_counts, (size_t)nb * 2));
    VO_CUDA(c, launch_pdl(cand_classify_kernel, dim3(nb), dim3(kCandThreads), c->stream,
        p->d_sc, a, (const float2*)p->klt_out.p, (const uint8_t*)p->klt_status.p, (const float2*)p->C[p->ccur].p,
        (const double2*)p->F[p->ccur].p, p->cand_flag.p, p->blk_counts.p, (const int32_t*)p->klt_iters.p, 0,
        (const int32_t*)nullptr));
    VO_CUDA(c, launch_pdl(cand_apply_kernel, dim3(nb), dim3(kCandThreads), c->stream,
        p->d_sc, a, (const float2*)p->klt_out.p, (const uint8_t*)p->klt_status.p, (const float2*)p->C[p->ccur].p,
        (const double2*)p->F[p->ccur].p, (const int32_t*)p->Ti[p->ccur].p, (const uint8_t*)p->cand_flag.p,
        (const int32_t*)p->blk_counts.p, (float2*)p->C[p->ccur ^ 1].p, (double2*)p->F[p->ccur ^ 1].p,
        p->Ti[p->ccur ^ 1].p, (const double*)p->poses.p, (float2*)p->P[p->cur].p, p->X[p->cur].p, land_cap, 0,
        (double*)nullptr, (int32_t*)nullptr, (const int32_t*)nullptr, (float2*)nullptr));
    c->launches++;
    c->launches++;
    p->ccur ^= 1;
    p->mask_src_valid = use_klt; p->mask_src_off = b_off;
    p->n_land_upper += ncand_upper;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

int launch_cand_add(Ctx* c, const double* d_pts, int n_new, const double* d_Thost, const int32_t* d_n_new,
                    bool on_side_stream, bool write_pose)
{
    VO_NVTX("vo:cand_add");
    Pipeline* p = c->pipe;
    if (!on_side_stream) VO_TRY(join_side_stream(c));
    const size_t need = (size_t)p->n_cand_upper + (n_new > 0 ? n_new : 0) + 1;
    VO_TRY(ensure(c, p->C[p->ccur], need * 2, true));
    VO_TRY(ensure(c, p->F[p->ccur], need * 2, true));
    VO_TRY(ensure(c, p->Ti[p->ccur], need, true));
    VO_TRY(ensure(c, p->poses, (size_t)(p->n_poses + 1) * 16, true));
    const int cap = (int)(p->C[p->ccur].cap / 2);
    const int threads = 256, blocks = ((n_new > 16 ? n_new : 16) + threads - 1) / threads;
    VO_CUDA(c, launch_pdl(cand_add_kernel, dim3(blocks), dim3(threads), on_side_stream ? c->stream2 : c->stream, p->d_sc,
                          (const double2*)d_pts, n_new, d_n_new, (int)(d_Thost == nullptr), d_Thost, (float2*)p->C[p->ccur].p,
                          (double2*)p->F[p->ccur].p, p->Ti[p->ccur].p, p->poses.p, p->n_poses, cap, (int)on_side_stream,
                          (int)write_pose));
    c->launches++;
    p->n_poses++;
    p->n_cand_upper += n_new > 0 ? n_new : 0;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

// d_P / d_X must be device-accessible (device memory or pinned host memory, read over the unified
// address space).  The host-side bound of |P| follows the same rule as the kernel.
int launch_land_append(Ctx* c, const float* d_P, const double* d_X, int n_rows, int target, const int32_t* gate)
{
    VO_NVTX("vo:land_append");
    Pipeline* p = c->pipe;
    if (n_rows <= 0) return VO_OK;
    int upper = p->n_land_upper + n_rows;
    if (target > 0 && upper > target) upper = target > p->n_land_upper ? target : p->n_land_upper;
    VO_TRY(ensure(c, p->P[p->cur], (size_t)upper * 2 + 2, true));
    VO_TRY(ensure(c, p->X[p->cur], (size_t)upper * 3 + 3, true));
    const int cap = (int)(p->P[p->cur].cap / 2);
    VO_CUDA(c, launch_pdl(land_append_kernel, dim3(1), dim3(1024), c->stream, p->d_sc, (const float2*)d_P, d_X, n_rows,
                          target, (float2*)p->P[p->cur].p, p->X[p->cur].p, cap, gate));
    c->launches++;
    p->n_land_upper = upper;
    p->src_valid = false;   // (rows without a track index)
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

// See frame_tail_kernel.  The early classification of this frame must be pending with this alpha (the caller checked);
// segments: device-accessible rows.  Afterwards: scalar block read-back and the promoted rows' DLT on the pose stream.
int launch_frame_tail(Ctx* c, const double K[9], double alpha_deg, int b_off, const double* d_feed, int n_new,
                      const Pipeline::LandSeg* segs, int n_segs)
{
    VO_NVTX("vo:frame_tail");
    Pipeline* p = c->pipe;
    if (c->side_inflight) {   // the candidates' classification comes from the side stream
        c->side_inflight = false; c->side_writes_counts = false;
        VO_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    }
    p->early_valid = false;
    PromoteArgs pa;
    memcpy(pa.K, K, sizeof(pa.K));
    memset(pa.Tnew, 0, sizeof(pa.Tnew));
    pa.alpha_deg = alpha_deg; pa.use_dev_T = 1; pa.use_klt = 1; pa.b_off = b_off; pa.ncand_upper = p->early_ncand_upper;
    // capacities: promoted rows + every batch behind the landmarks; the new candidates behind the candidates
    int upper = p->n_land_upper + p->early_ncand_upper;
    for (int k = 0; k < n_segs; k++) {
        int up = upper + segs[k].n;
        if (segs[k].target > 0 && up > segs[k].target) up = segs[k].target > upper ? segs[k].target : upper;
        upper = up;
    }
    VO_TRY(ensure(c, p->P[p->cur], (size_t)upper * 2 + 2, true));
    VO_TRY(ensure(c, p->X[p->cur], (size_t)upper * 3 + 3, true));
    const size_t cneed = (size_t)p->n_cand_upper + (n_new > 0 ? n_new : 0) + 1;
    VO_TRY(ensure(c, p->C[p->ccur], cneed * 2, true));
    VO_TRY(ensure(c, p->F[p->ccur], cneed * 2, true));
    VO_TRY(ensure(c, p->Ti[p->ccur], cneed, true));
    VO_TRY(ensure(c, p->poses, (size_t)(p->n_poses + 1) * 16, true));
    TailArgs a;
    memset(&a, 0, sizeof(a));
    a.prom_rows = p->prom_rows.p; a.P = (float2*)p->P[p->cur].p; a.X = p->X[p->cur].p; a.land_cap = (int)(p->P[p->cur].cap / 2);
    a.feed = (const double2*)d_feed; a.n_new = n_new > 0 ? n_new : 0;
    a.C = (float2*)p->C[p->ccur].p; a.F = (double2*)p->F[p->ccur].p; a.Ti = p->Ti[p->ccur].p;
    a.pose_idx = p->n_poses; a.cand_cap = (int)(p->C[p->ccur].cap / 2);
    a.n_segs = n_segs;
    for (int k = 0; k < n_segs; k++) { a.segP[k] = (const float2*)segs[k].P; a.segX[k] = segs[k].X; a.segN[k] = segs[k].n; a.segTarget[k] = segs[k].target; }
    // where each appended row's track sits, should the next frame's KLT run ahead of this frame's pruning (Pipeline::src_*):
    // [0, b1_upper) the rows of the last compaction, then the parked rows, then every offered row of every batch
    if (p->b1_buf >= 0 && p->Lsrc[p->cur].p) {
        VO_TRY(ensure(c, p->Lsrc[p->cur], (size_t)upper + 1, true));
        a.src = p->Lsrc[p->cur].p;
        p->src_prom_base = p->b1_upper; p->src_prom_upper = p->early_ncand_upper;
        int base = p->src_prom_base + p->src_prom_upper;
        p->src_n_segs = n_segs;
        for (int k = 0; k < n_segs; k++) { p->src_seg_base[k] = base; p->src_seg_n[k] = segs[k].n; p->src_seg_P[k] = segs[k].P; base += segs[k].n; }
        p->src_total = base;
        a.src_prom_base = p->src_prom_base;
        for (int k = 0; k < n_segs; k++) a.src_seg_base[k] = p->src_seg_base[k];
    }
    VO_CUDA(c, launch_pdl(frame_tail_kernel, dim3(1), dim3(1024), c->stream, p->d_sc, a));
    p->src_valid = a.src != nullptr;
    VO_CUDA(c, cudaEventRecord(c->ev_snap, c->stream));
    // pose stream, behind the refit: scalar block back to the host, then the 3-D points and the pose-table row
    VO_CUDA(c, cudaStreamWaitEvent(c->stream4, c->ev_snap, 0));
    VO_CUDA(c, cudaMemcpyAsync(p->h_sc, p->d_sc, sizeof(DevScalars), cudaMemcpyDeviceToHost, c->stream4));
    VO_CUDA(c, cudaEventRecord(p->ev_pose, c->stream4));
    VO_CUDA(c, launch_pdl(promote_finish_kernel, dim3(1), dim3(128), c->stream4, p->d_sc, pa, (const double*)p->prom_rows.p,
                          (const int32_t*)p->prom_ti.p, p->poses.p, (float2*)p->P[p->cur].p, p->X[p->cur].p, a.land_cap, 2,
                          n_new > 0 ? p->n_poses : -1));
    VO_CUDA(c, cudaEventRecord(c->ev_pose_done, c->stream4));
    c->pose_inflight = true;
    c->launches += 2;
    if (n_new > 0) { p->n_poses++; p->n_cand_upper += n_new; }
    p->n_land_upper = upper;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

int launch_triangulate(Ctx* c, const double* d_M1, const double* d_M2, int per_point, const double* d_p1,
                       const double* d_p2, int n, double* d_X4)
{
    if (n <= 0) return VO_OK;
    triangulate_kernel<<<(n + 127) / 128, 128, 0, c->stream>>>(d_M1, d_M2, per_point, (const double2*)d_p1,
                                                              (const double2*)d_p2, n, d_X4);
    c->launches++;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

}  // namespace vo
