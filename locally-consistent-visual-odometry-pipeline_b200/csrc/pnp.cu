// K3-K6: PnP-RANSAC with P3P hypotheses and an LM refit, device-resident.
//
// Replaces cv2.solvePnPRansac(X, P, K, zeros(4), flags=SOLVEPNP_ITERATIVE, confidence,
// reprojectionError=2) at /root/reference/src/vo_pipeline.py:258-264 plus the inlier pruning of
// :327-331 and the 4x4 assembly of :333 (behaviour: SURVEY.md App. A.2).  As BASELINE.json's
// north_star specifies, hypotheses come from a P3P minimal solver (3 points + 1 disambiguation
// point); everything else follows OpenCV's loop exactly so that, for the same minimal sets, the
// inlier mask is bit-identical to cv2.solvePnPRansac(flags=SOLVEPNP_P3P):
//   - minimal sets: cv::RNG(2^64-1) emulated on the device (the point count is only known on
//     the device, so the table cannot come from the host without a sync) or an explicit table;
//   - scoring: float64 projection rounded to float32, float32 dx*dx+dy*dy <= (float)(thr*thr);
//   - sequential semantics (first strictly-better model wins, shrinking iteration budget) are
//     emulated over parallel chunks of hypotheses: a chunk is solved+scored by one CTA per
//     hypothesis, then one warp scans the counts in order; later chunks skip every hypothesis
//     whose index is already beyond the budget;
//   - refit: LM on SE(3) over the winner's inlier set to the converged optimum (what
//     cv2.solvePnP(SOLVEPNP_ITERATIVE) returns), single CTA, deterministic tree reductions.
// No tensor cores: FP64 CUDA-core work (P3P, projection) and FP32 compares.
#include <float.h>
#include "ctx.cuh"

namespace vo {

namespace {

constexpr int kHypThreads = 256;
constexpr int kFinThreads = 512;

struct PnpArgs {
    double K[9];
    double confidence;
    float thr2;          // (float)(thr*thr)
    int max_iters;
    int refine_iters;
    int n_upper;
    int explicit_sets;
    int compact;
};

// ---- cv::RNG -----------------------------------------------------------------------------------
__device__ __forceinline__ unsigned rng_next(uint64_t& s)
{
    s = (uint64_t)(unsigned)s * 4164903690ULL + (unsigned)(s >> 32);
    return (unsigned)s;
}

// rows [sc->sets_ready, to) of the minimal-set table, exactly as getSubset draws them
__device__ void gen_sets(DevScalars* sc, int32_t* sets, int n, int to)
{
    uint64_t s = sc->rng;
    for (int h = sc->sets_ready; h < to; h++) {
        int idx[kMinimal];
#pragma unroll
        for (int i = 0; i < kMinimal; i++) {
            for (;;) {
                const int v = (int)(rng_next(s) % (unsigned)n);
                bool dup = false;
#pragma unroll
                for (int j = 0; j < kMinimal; j++) dup |= (j < i) && idx[j] == v;
                if (!dup) { idx[i] = v; break; }
            }
        }
        reinterpret_cast<int4*>(sets)[h] = make_int4(idx[0], idx[1], idx[2], idx[3]);
    }
    sc->rng = s;
    sc->sets_ready = to;
}

// RANSACUpdateNumIters
__device__ int update_num_iters(double p, double ep, int model_points, int max_iters)
{
    p = fmin(fmax(p, 0.), 1.);
    ep = fmin(fmax(ep, 0.), 1.);
    double num = fmax(1. - p, DBL_MIN);
    double denom = 1. - pow(1. - ep, (double)model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)rint(num / denom);
}

// ---- polynomial real roots by bracketing between the critical points ---------------------------
__device__ __forceinline__ double peval(const double* c, int deg, double x)
{
    double v = c[deg];
    for (int i = deg - 1; i >= 0; i--) v = v * x + c[i];
    return v;
}

__device__ double bracket_root(const double* c, int deg, double lo, double hi, double flo)
{
    double d[4];
    for (int i = 1; i <= deg; i++) d[i - 1] = c[i] * i;
    double x = 0.5 * (lo + hi);
    for (int it = 0; it < 200; it++) {
        const double f = peval(c, deg, x);
        if (f == 0.) return x;
        if ((f < 0.) == (flo < 0.)) lo = x; else hi = x;
        const double df = peval(d, deg - 1, x);
        double xn = df != 0. ? x - f / df : 0.5 * (lo + hi);
        if (!(xn > lo && xn < hi)) xn = 0.5 * (lo + hi);
        if (fabs(xn - x) <= 4e-16 * fabs(x) || hi - lo <= 4e-16 * fabs(x)) return xn;
        x = xn;
    }
    return x;
}

// real roots of c[0..deg] (deg <= 4) in (lo, hi], ascending
__device__ int real_roots(const double* cin, int deg, double lo, double hi, double* roots)
{
    double big = 0.;
    for (int i = 0; i <= deg; i++) big = fmax(big, fabs(cin[i]));
    while (deg > 0 && fabs(cin[deg]) <= 1e-14 * big) deg--;
    if (deg == 0) return 0;
    double der[4][5];
    int degs[4];
    for (int i = 0; i <= deg; i++) der[0][i] = cin[i];
    degs[0] = deg;
    int nd = 1;
    while (degs[nd - 1] > 1) {
        for (int i = 1; i <= degs[nd - 1]; i++) der[nd][i - 1] = der[nd - 1][i] * i;
        degs[nd] = degs[nd - 1] - 1;
        nd++;
    }
    double r[4], prev[4];
    int nprev = 0;
    {
        const double x = -der[nd - 1][0] / der[nd - 1][1];
        if (x > lo && x <= hi) prev[nprev++] = x;
    }
    for (int k = nd - 2; k >= 0; k--) {
        const double* p = der[k];
        const int dg = degs[k];
        int nr = 0;
        double a = lo, fa = peval(p, dg, a);
        for (int i = 0; i <= nprev; i++) {
            const double b = i < nprev ? prev[i] : hi;
            const double fb = peval(p, dg, b);
            if (b > a) {
                if (fb == 0.) r[nr++] = b;
                else if (fa != 0. && (fa < 0.) != (fb < 0.)) r[nr++] = bracket_root(p, dg, a, b, fa);
            }
            a = b; fa = fb;
        }
        nprev = nr;
        for (int i = 0; i < nr; i++) prev[i] = r[i];
    }
    for (int i = 0; i < nprev; i++) roots[i] = prev[i];
    return nprev;
}

__device__ __forceinline__ void cross3(const double* a, const double* b, double* c)
{
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

// float64 projection in cv::projectPoints' operation order (explicitly unfused), rounded to
// float32, float32 squared error
__device__ __forceinline__ float reproj_err2(const double* R, const double* t, const double* K, float xf,
                                             float yf, float zf, float u, float v)
{
    const double X = xf, Y = yf, Z = zf;
    double x = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[0], X), __dmul_rn(R[1], Y)), __dmul_rn(R[2], Z)), t[0]);
    double y = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[3], X), __dmul_rn(R[4], Y)), __dmul_rn(R[5], Z)), t[1]);
    double z = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(R[6], X), __dmul_rn(R[7], Y)), __dmul_rn(R[8], Z)), t[2]);
    z = z != 0. ? __ddiv_rn(1., z) : 1.;
    x = __dmul_rn(x, z); y = __dmul_rn(y, z);
    const float pu = (float)__dadd_rn(__dmul_rn(x, K[0]), K[2]);
    const float pv = (float)__dadd_rn(__dmul_rn(y, K[4]), K[5]);
    const float dx = __fsub_rn(u, pu), dy = __fsub_rn(v, pv);
    return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
}

// P3P (Grunert quartic; roots isolated by bracketing; distances polished by Gauss-Newton; exact
// 3-point alignment).  X: 4 world points (rows), p: 4 pixel points.  Roots come from the first
// three correspondences, the fourth picks the root with the smallest pixel error.  Rt = [R|t]
// as 12 doubles (R row-major then t).
__device__ bool p3p4(const double* X, const double* p, const double* K, double* Rt)
{
    double j[9];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const double x = (p[2 * i] - K[2]) / K[0], y = (p[2 * i + 1] - K[5]) / K[4];
        const double n = sqrt(x * x + y * y + 1.);
        j[3 * i] = x / n; j[3 * i + 1] = y / n; j[3 * i + 2] = 1. / n;
    }
    const double* X1 = X; const double* X2 = X + 3; const double* X3 = X + 6;
    double d12[3], d13[3], d23[3];
#pragma unroll
    for (int i = 0; i < 3; i++) { d12[i] = X1[i] - X2[i]; d13[i] = X1[i] - X3[i]; d23[i] = X2[i] - X3[i]; }
    const double a2 = dot3(d23, d23), b2 = dot3(d13, d13), c2 = dot3(d12, d12);
    if (!(a2 > 0. && b2 > 0. && c2 > 0.)) return false;
    const double ca = dot3(j + 3, j + 6), cb = dot3(j, j + 6), cg = dot3(j, j + 3);
    double nX[3], XmInv[9];
    cross3(d12, d13, nX);
    {
        double bc[3], cA[3], ab[3];
        cross3(d13, nX, bc); cross3(nX, d12, cA); cross3(d12, d13, ab);
        const double det = dot3(d12, bc);
        const double scale = sqrt(dot3(d12, d12) * dot3(d13, d13) * dot3(nX, nX));
        if (!(fabs(det) > 1e-12 * scale) || scale == 0.) return false;
        const double id = 1. / det;
#pragma unroll
        for (int i = 0; i < 3; i++) { XmInv[i] = bc[i] * id; XmInv[3 + i] = cA[i] * id; XmInv[6 + i] = ab[i] * id; }
    }
    // u = N(v)/D(v); quartic  N^2 - 2 cg N D + (1 - k W) D^2 = 0  in v = s3/s1
    const double q = (a2 - c2) / b2, k = c2 / b2;
    const double N[3] = { 1. + q, -2. * q * cb, q - 1. };
    const double D[2] = { 2. * cg, -2. * ca };
    const double W[3] = { 1., -2. * cb, 1. };
    double NN[5] = { 0, 0, 0, 0, 0 }, ND[4] = { 0, 0, 0, 0 }, DD[3] = { 0, 0, 0 }, G[3], GDD[5] = { 0, 0, 0, 0, 0 }, Q[5];
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = 0; b < 3; b++) NN[a + b] += N[a] * N[b];
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = 0; b < 2; b++) ND[a + b] += N[a] * D[b];
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
        for (int b = 0; b < 2; b++) DD[a + b] += D[a] * D[b];
#pragma unroll
    for (int a = 0; a < 3; a++) G[a] = (a == 0 ? 1. : 0.) - k * W[a];
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = 0; b < 3; b++) GDD[a + b] += G[a] * DD[b];
    double big = 0.;
#pragma unroll
    for (int a = 0; a < 5; a++) { Q[a] = NN[a] + GDD[a] - (a < 4 ? 2. * cg * ND[a] : 0.); big = fmax(big, fabs(Q[a])); }
    if (!(big > 0.) || !isfinite(big)) return false;
    double vr[4];
    const int nv = real_roots(Q, 4, 0., 1e6, vr);

    bool found = false;
    double beste = 0.;
    for (int r = 0; r < nv; r++) {
        const double v = vr[r];
        const double Wv = 1. + v * (v - 2. * cb);
        if (!(Wv > 0.)) continue;
        const double Dv = D[0] + D[1] * v;
        double u;
        if (fabs(Dv) > 1e-9) {
            u = (N[0] + v * (N[1] + v * N[2])) / Dv;
        } else {
            const double cc = 1. - k * Wv, disc = cg * cg - cc;
            if (!(disc >= 0.)) continue;
            const double sq = sqrt(disc), u0 = cg + sq, u1 = cg - sq;
            const double e0 = fabs(u0 * u0 + v * v - 2. * u0 * v * ca - (a2 / b2) * Wv);
            const double e1 = fabs(u1 * u1 + v * v - 2. * u1 * v * ca - (a2 / b2) * Wv);
            u = e0 <= e1 ? u0 : u1;
        }
        if (!(u > 0.)) continue;
        double s1 = sqrt(b2 / Wv), s2 = u * s1, s3 = v * s1;
        for (int it = 0; it < 5; it++) {
            const double r12 = s1 * s1 + s2 * s2 - 2. * s1 * s2 * cg - c2;
            const double r13 = s1 * s1 + s3 * s3 - 2. * s1 * s3 * cb - b2;
            const double r23 = s2 * s2 + s3 * s3 - 2. * s2 * s3 * ca - a2;
            const double J00 = 2. * (s1 - s2 * cg), J01 = 2. * (s2 - s1 * cg);
            const double J10 = 2. * (s1 - s3 * cb), J12 = 2. * (s3 - s1 * cb);
            const double J21 = 2. * (s2 - s3 * ca), J22 = 2. * (s3 - s2 * ca);
            const double det = -J00 * J12 * J21 - J01 * J10 * J22;
            if (!(fabs(det) > 1e-300)) break;
            const double ds1 = (-J12 * J21 * r12 - J01 * J22 * r13 + J01 * J12 * r23) / det;
            const double ds2 = (-J10 * J22 * r12 + J00 * J22 * r13 - J00 * J12 * r23) / det;
            const double ds3 = (J10 * J21 * r12 - J00 * J21 * r13 - J01 * J10 * r23) / det;
            s1 -= ds1; s2 -= ds2; s3 -= ds3;
            if (fabs(ds1) + fabs(ds2) + fabs(ds3) <= 1e-15 * (s1 + s2 + s3)) break;
        }
        if (!(s1 > 0. && s2 > 0. && s3 > 0.)) continue;
        double e12[3], e13[3], nY[3], Y1[3];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            Y1[i] = s1 * j[i];
            e12[i] = Y1[i] - s2 * j[3 + i];
            e13[i] = Y1[i] - s3 * j[6 + i];
        }
        cross3(e12, e13, nY);
        double Rc[12];
        bool fin = true;
#pragma unroll
        for (int i = 0; i < 3; i++) {
#pragma unroll
            for (int c = 0; c < 3; c++) {
                Rc[3 * i + c] = e12[i] * XmInv[c] + e13[i] * XmInv[3 + c] + nY[i] * XmInv[6 + c];
                fin &= isfinite(Rc[3 * i + c]);
            }
            Rc[9 + i] = Y1[i] - (Rc[3 * i] * X1[0] + Rc[3 * i + 1] * X1[1] + Rc[3 * i + 2] * X1[2]);
        }
        if (!fin) continue;
        // 4th point: pixel reprojection error (float64)
        const double* X4 = X + 9;
        double x = Rc[0] * X4[0] + Rc[1] * X4[1] + Rc[2] * X4[2] + Rc[9];
        double y = Rc[3] * X4[0] + Rc[4] * X4[1] + Rc[5] * X4[2] + Rc[10];
        double z = Rc[6] * X4[0] + Rc[7] * X4[1] + Rc[8] * X4[2] + Rc[11];
        z = z != 0. ? 1. / z : 1.;
        const double du = x * z * K[0] + K[2] - p[6], dv = y * z * K[4] + K[5] - p[7];
        const double e = du * du + dv * dv;
        if (!found || e < beste) {
            found = true; beste = e;
#pragma unroll
            for (int i = 0; i < 12; i++) Rt[i] = Rc[i];
        }
    }
    return found;
}

// ---- kernels -----------------------------------------------------------------------------------
__global__ void pnp_init_kernel(DevScalars* sc, int32_t* sets, const int32_t* d_n, PnpArgs a, int chunk0_end)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int n = d_n ? min(*d_n, a.n_upper) : a.n_upper;
    sc->n_tracked = n;
    sc->niters = a.max_iters; sc->best_count = 0; sc->best_hyp = -1; sc->iters_run = 0; sc->done = 0;
    sc->sets_ready = 0; sc->rng = 0xFFFFFFFFFFFFFFFFULL; sc->ok = 0; sc->n_inliers = 0; sc->lm_iters = 0;
    if (n < kMinimal || a.max_iters <= 0) { sc->niters = 0; sc->done = 1; return; }
    if (n == kMinimal) {   // cv2: count == modelPoints -> one direct solve on all points
        reinterpret_cast<int4*>(sets)[0] = make_int4(0, 1, 2, 3);
        sc->niters = 1; sc->sets_ready = a.max_iters;
        return;
    }
    if (!a.explicit_sets) gen_sets(sc, sets, n, chunk0_end);
}

// one CTA per hypothesis: thread 0 solves P3P, the CTA scores all points
__global__ void __launch_bounds__(kHypThreads)
pnp_hyp_kernel(const DevScalars* __restrict__ sc, const int32_t* __restrict__ sets, const float4* __restrict__ q4,
               const float* __restrict__ qv, double* __restrict__ hyp, int32_t* __restrict__ counts, PnpArgs a, int h0)
{
    const int h = h0 + blockIdx.x;
    if (h >= sc->niters) return;   // beyond the (shrinking) iteration budget: never looked at
    const int n = sc->n_tracked;
    __shared__ double sRt[12];
    __shared__ int sValid;
    __shared__ int sCount[kHypThreads / 32];
    if (threadIdx.x == 0) {
        const int4 s = reinterpret_cast<const int4*>(sets)[h];
        const int idx[4] = { s.x, s.y, s.z, s.w };
        double X[12], p[8];
        bool ok = true;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            ok &= idx[k] >= 0 && idx[k] < n;
            const int i = ok ? idx[k] : 0;
            const float4 q = q4[i];
            X[3 * k] = q.x; X[3 * k + 1] = q.y; X[3 * k + 2] = q.z;
            p[2 * k] = q.w; p[2 * k + 1] = qv[i];
        }
        double Rt[12];
        ok = ok && p3p4(X, p, a.K, Rt);
        sValid = ok;
        if (ok) {
#pragma unroll
            for (int i = 0; i < 12; i++) { sRt[i] = Rt[i]; hyp[(size_t)h * 12 + i] = Rt[i]; }
        }
    }
    __syncthreads();
    if (!sValid) { if (threadIdx.x == 0) counts[h] = -1; return; }
    double R[9], t[3];
#pragma unroll
    for (int i = 0; i < 9; i++) R[i] = sRt[i];
#pragma unroll
    for (int i = 0; i < 3; i++) t[i] = sRt[9 + i];
    int good = 0;
    for (int i = threadIdx.x; i < n; i += kHypThreads) {
        const float4 q = q4[i];
        good += reproj_err2(R, t, a.K, q.x, q.y, q.z, q.w, qv[i]) <= a.thr2;
    }
    good = __reduce_add_sync(0xffffffffu, good);
    if ((threadIdx.x & 31) == 0) sCount[threadIdx.x >> 5] = good;
    __syncthreads();
    if (threadIdx.x == 0) {
        int s = 0;
#pragma unroll
        for (int w = 0; w < kHypThreads / 32; w++) s += sCount[w];
        counts[h] = s;
    }
}

// one warp: sequential RANSAC bookkeeping over the counts of hypotheses [h0, h1)
__global__ void pnp_select_kernel(DevScalars* sc, int32_t* sets, const int32_t* __restrict__ counts, PnpArgs a,
                                  int h0, int h1, int next_end, int last)
{
    if (sc->done) return;
    const int lane = threadIdx.x;
    const int n = sc->n_tracked;
    int niters = sc->niters, best = sc->best_count, best_hyp = sc->best_hyp;
    int h = h0, done = 0;
    while (h < h1) {
        if (h >= niters) { done = 1; break; }
        const int lim = min(h1, niters);
        const int c = (h + lane < lim) ? counts[h + lane] : -1;
        const unsigned m = __ballot_sync(0xffffffffu, c > max(best, kMinimal - 1));
        if (!m) { h += 32; continue; }
        const int first = __ffs(m) - 1;
        best = __shfl_sync(0xffffffffu, c, first);
        best_hyp = h + first;
        niters = n == kMinimal ? niters : update_num_iters(a.confidence, (double)(n - best) / n, kMinimal, niters);
        h = best_hyp + 1;
    }
    if (h >= niters || last) done = 1;
    if (lane == 0) {
        sc->niters = niters; sc->best_count = best; sc->best_hyp = best_hyp; sc->done = done;
        // iterations the sequential loop executes: it leaves at the first iter >= niters, and the
        // budget may have dropped below the iteration that lowered it
        sc->iters_run = h >= niters ? max(niters, best_hyp + 1) : h1;
        if (!done && !a.explicit_sets && next_end > sc->sets_ready) gen_sets(sc, sets, n, next_end);
    }
}

// Single CTA: inlier mask of the winner, LM refit on SE(3), pose output, optional pruning of the
// landmark buffers to the inliers (vo_pipeline.py:327-331).
struct FinalizeBufs {
    const float* P_in; const double* X_in; float* P_out; double* X_out;   // landmark ping-pong
};

__device__ __forceinline__ void so3_exp(const double* w, double* R)
{
    const double th2 = dot3(w, w), th = sqrt(th2);
    double A, B;
    if (th < 1e-8) { A = 1. - th2 / 6.; B = 0.5 - th2 / 24.; }
    else { A = sin(th) / th; B = (1. - cos(th)) / th2; }
    const double wx = w[0], wy = w[1], wz = w[2];
    R[0] = 1. - B * (wy * wy + wz * wz); R[1] = B * wx * wy - A * wz;       R[2] = B * wx * wz + A * wy;
    R[3] = B * wx * wy + A * wz;        R[4] = 1. - B * (wx * wx + wz * wz); R[5] = B * wy * wz - A * wx;
    R[6] = B * wx * wz - A * wy;        R[7] = B * wy * wz + A * wx;        R[8] = 1. - B * (wx * wx + wy * wy);
}

__device__ bool chol6_solve(double* A, double* b)
{
    for (int i = 0; i < 6; i++)
        for (int j = 0; j <= i; j++) {
            double s = A[6 * i + j];
            for (int k = 0; k < j; k++) s -= A[6 * i + k] * A[6 * j + k];
            if (i == j) { if (!(s > 0.)) return false; A[6 * i + i] = sqrt(s); }
            else A[6 * i + j] = s / A[6 * j + j];
        }
    for (int i = 0; i < 6; i++) { double s = b[i]; for (int k = 0; k < i; k++) s -= A[6 * i + k] * b[k]; b[i] = s / A[6 * i + i]; }
    for (int i = 5; i >= 0; i--) { double s = b[i]; for (int k = i + 1; k < 6; k++) s -= A[6 * k + i] * b[k]; b[i] = s / A[6 * i + i]; }
    return true;
}

constexpr int kAcc = 28;  // 21 (lower triangle of JtJ) + 6 (Jt r) + 1 (cost)

// accumulate JtJ, Jt r and the cost at pose Rt over the masked points; result in sAcc (all threads
// see it after the trailing __syncthreads)
__device__ void lm_accumulate(const double* Rt, const double* K, const float4* __restrict__ q4,
                              const float* __restrict__ qv, const uint8_t* __restrict__ mask, int n,
                              double (*sPart)[kAcc], double* sAcc)
{
    double acc[kAcc];
#pragma unroll
    for (int i = 0; i < kAcc; i++) acc[i] = 0.;
    for (int i = threadIdx.x; i < n; i += kFinThreads) {
        if (!mask[i]) continue;
        const float4 q = q4[i];
        const double X = q.x, Y = q.y, Z = q.z;
        const double rx = Rt[0] * X + Rt[1] * Y + Rt[2] * Z;
        const double ry = Rt[3] * X + Rt[4] * Y + Rt[5] * Z;
        const double rz = Rt[6] * X + Rt[7] * Y + Rt[8] * Z;
        const double x = rx + Rt[9], y = ry + Rt[10], z = rz + Rt[11];
        const double iz = 1. / z;
        const double ru = K[0] * x * iz + K[2] - (double)q.w;
        const double rv = K[4] * y * iz + K[5] - (double)qv[i];
        const double ux = K[0] * iz, uz = -K[0] * x * iz * iz;
        const double vy = K[4] * iz, vz = -K[4] * y * iz * iz;
        // d xc / d w = -[R X]x ; d xc / d t = I
        const double Ju[6] = { uz * ry, ux * rz - uz * rx, -ux * ry, ux, 0., uz };
        const double Jv[6] = { -vy * rz + vz * ry, -vz * rx, vy * rx, 0., vy, vz };
        int k = 0;
#pragma unroll
        for (int a = 0; a < 6; a++)
#pragma unroll
            for (int b = 0; b <= a; b++) acc[k++] += Ju[a] * Ju[b] + Jv[a] * Jv[b];
#pragma unroll
        for (int a = 0; a < 6; a++) acc[21 + a] += Ju[a] * ru + Jv[a] * rv;
        acc[27] += ru * ru + rv * rv;
    }
#pragma unroll
    for (int i = 0; i < kAcc; i++) {
        double v = acc[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        acc[i] = v;
    }
    __syncthreads();   // previous readers of sPart/sAcc are done
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int i = 0; i < kAcc; i++) sPart[threadIdx.x >> 5][i] = acc[i];
    }
    __syncthreads();
    if (threadIdx.x < kAcc) {
        double s = 0.;
#pragma unroll
        for (int w = 0; w < kFinThreads / 32; w++) s += sPart[w][threadIdx.x];
        sAcc[threadIdx.x] = s;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kFinThreads)
pnp_finalize_kernel(DevScalars* sc, const double* __restrict__ hyp, const float4* __restrict__ q4,
                    const float* __restrict__ qv, uint8_t* __restrict__ mask, PnpArgs a, FinalizeBufs fb)
{
    __shared__ double sPart[kFinThreads / 32][kAcc];
    __shared__ double sAcc[kAcc];
    __shared__ double sRt[12], sTrial[12], sH[21], sG[6];
    __shared__ double sCost, sLambda, sStep;
    __shared__ int sFails, sIters;
    __shared__ int sState;   // 0 = continue with trial, 1 = stop
    __shared__ int sWarpTot[kFinThreads / 32];
    __shared__ int sBase;
    const int tid = threadIdx.x;
    const int n = sc->n_tracked;
    const int best = sc->best_hyp;
    const bool ok = best >= 0;

    int ninl = 0;
    if (ok) {
        if (tid < 12) sRt[tid] = hyp[(size_t)best * 12 + tid];
        __syncthreads();
        double R[9], t[3];
#pragma unroll
        for (int i = 0; i < 9; i++) R[i] = sRt[i];
#pragma unroll
        for (int i = 0; i < 3; i++) t[i] = sRt[9 + i];
        for (int i = tid; i < n; i += kFinThreads) {
            const float4 q = q4[i];
            const uint8_t f = n == kMinimal ? 1 : (reproj_err2(R, t, a.K, q.x, q.y, q.z, q.w, qv[i]) <= a.thr2);
            mask[i] = f;
        }
        __syncthreads();

        // ---- LM: evaluate at the current pose, then trial steps ------------------------------------
        lm_accumulate(sRt, a.K, q4, qv, mask, n, sPart, sAcc);
        if (tid == 0) {
            for (int i = 0; i < 21; i++) sH[i] = sAcc[i];
            for (int i = 0; i < 6; i++) sG[i] = sAcc[21 + i];
            sCost = sAcc[27]; sLambda = 1e-3; sState = 0; sFails = 0; sIters = 0;
        }
        __syncthreads();
        while (sIters < a.refine_iters) {
            if (tid == 0) {
                double A[36], d[6];
                bool solved = false;
                while (!solved && sFails < 12) {
                    int k = 0;
                    for (int r = 0; r < 6; r++)
                        for (int c = 0; c <= r; c++) A[6 * r + c] = sH[k++];
                    for (int r = 0; r < 6; r++) { A[6 * r + r] *= 1. + sLambda; d[r] = -sG[r]; }
                    solved = chol6_solve(A, d);
                    if (!solved) { sLambda *= 10.; sFails++; }
                }
                if (!solved) sState = 1;
                else {
                    double dR[9];
                    so3_exp(d, dR);
                    for (int r = 0; r < 3; r++)
                        for (int c = 0; c < 3; c++)
                            sTrial[3 * r + c] = dR[3 * r] * sRt[c] + dR[3 * r + 1] * sRt[3 + c] + dR[3 * r + 2] * sRt[6 + c];
                    for (int r = 0; r < 3; r++) sTrial[9 + r] = sRt[9 + r] + d[3 + r];
                    double step = 0.;
                    for (int r = 0; r < 6; r++) step += d[r] * d[r];
                    sStep = step;
                    sState = 0;
                }
            }
            __syncthreads();
            if (sState) break;
            lm_accumulate(sTrial, a.K, q4, qv, mask, n, sPart, sAcc);
            if (tid == 0) {
                const double cn = sAcc[27];
                if (cn <= sCost) {
                    const double dec = sCost - cn;
                    for (int i = 0; i < 12; i++) sRt[i] = sTrial[i];
                    for (int i = 0; i < 21; i++) sH[i] = sAcc[i];
                    for (int i = 0; i < 6; i++) sG[i] = sAcc[21 + i];
                    sCost = cn; sLambda = fmax(sLambda * 0.1, 1e-12);
                    sFails = 0; sIters++;
                    sState = (sStep < 1e-24 || dec <= 1e-15 * cn) ? 1 : 0;   // 1 = converged
                } else {
                    sLambda *= 10.;
                    if (++sFails >= 12) sState = 1;
                }
            }
            __syncthreads();
            const int st = sState;
            __syncthreads();
            if (st) break;
        }
        // inlier count
        int c = 0;
        for (int i = tid; i < n; i += kFinThreads) c += mask[i];
        c = __reduce_add_sync(0xffffffffu, c);
        if ((tid & 31) == 0) sWarpTot[tid >> 5] = c;
        __syncthreads();
        for (int w = 0; w < kFinThreads / 32; w++) ninl += sWarpTot[w];
        __syncthreads();
        if (tid == 0) {
            for (int r = 0; r < 3; r++) {
                for (int cc = 0; cc < 3; cc++) { sc->T[4 * r + cc] = sRt[3 * r + cc]; }
                sc->T[4 * r + 3] = sRt[9 + r];
            }
            sc->T[12] = 0.; sc->T[13] = 0.; sc->T[14] = 0.; sc->T[15] = 1.;
            for (int i = 0; i < 12; i++) sc->Th[i] = hyp[(size_t)best * 12 + i];
            sc->cost = sCost; sc->lm_iters = sIters;
        }
    } else {
        for (int i = tid; i < n; i += kFinThreads) mask[i] = 0;
        if (tid < 16) sc->T[tid] = 0.;
        __syncthreads();
    }
    if (tid == 0) { sc->ok = ok; sc->n_inliers = ninl; }

    if (!a.compact) return;
    // ---- prune landmarks to the inliers, order preserved (no model: state left as it is) -------
    if (tid == 0) sBase = 0;
    __syncthreads();
    for (int i0 = 0; i0 < n; i0 += kFinThreads) {
        const int i = i0 + tid;
        const bool keep = i < n && (!ok || mask[i]);
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        const int wpre = __popc(bal & ((1u << (tid & 31)) - 1));
        if ((tid & 31) == 0) sWarpTot[tid >> 5] = __popc(bal);
        __syncthreads();
        int off = sBase;
        for (int w = 0; w < (tid >> 5); w++) off += sWarpTot[w];
        if (keep) {
            const int o = off + wpre;
            reinterpret_cast<float2*>(fb.P_out)[o] = reinterpret_cast<const float2*>(fb.P_in)[i];
            fb.X_out[3 * (size_t)o] = fb.X_in[3 * (size_t)i];
            fb.X_out[3 * (size_t)o + 1] = fb.X_in[3 * (size_t)i + 1];
            fb.X_out[3 * (size_t)o + 2] = fb.X_in[3 * (size_t)i + 2];
        }
        __syncthreads();
        if (tid == 0) { int s = 0; for (int w = 0; w < kFinThreads / 32; w++) s += sWarpTot[w]; sBase += s; }
        __syncthreads();
    }
    if (tid == 0) sc->n_land = sBase;
}

}  // namespace

int launch_pnp(Ctx* c, const PnpLaunch& L)
{
    Pipeline* p = c->pipe;
    PnpArgs a;
    memcpy(a.K, L.K, sizeof(a.K));
    a.confidence = L.prm.confidence;
    a.thr2 = (float)(L.prm.reproj_thresh * L.prm.reproj_thresh);
    a.max_iters = L.prm.max_iters < 0 ? 0 : L.prm.max_iters;
    a.refine_iters = L.prm.refine_iters < 0 ? 0 : L.prm.refine_iters;
    a.n_upper = L.n_upper;
    a.explicit_sets = L.d_sets_in != nullptr;
    a.compact = L.compact;
    const int H = a.max_iters;
    VO_TRY(ensure(c, p->hyp, (size_t)(H > 0 ? H : 1) * 12));
    VO_TRY(ensure(c, p->counts, (size_t)(H > 0 ? H : 1)));
    VO_TRY(ensure(c, p->mask, (size_t)(L.n_upper > 0 ? L.n_upper : 1)));
    int32_t* sets = const_cast<int32_t*>(L.d_sets_in);
    if (!sets) {
        VO_TRY(ensure(c, p->sets, (size_t)(H > 0 ? H : 1) * kMinimal));
        sets = p->sets.p;
    }
    // chunk boundaries: small first chunk (the adaptive stop usually ends inside it)
    int bounds[kPnpChunks + 1] = { 0, H < 128 ? H : 128, H < 512 ? H : 512, H };
    pnp_init_kernel<<<1, 32, 0, c->stream>>>(p->d_sc, sets, L.d_n, a, bounds[1]);
    c->launches++;
    for (int k = 0; k < kPnpChunks; k++) {
        const int h0 = bounds[k], h1 = bounds[k + 1];
        if (h1 <= h0) continue;
        const bool last = (h1 == H);
        pnp_hyp_kernel<<<h1 - h0, kHypThreads, 0, c->stream>>>(p->d_sc, sets, (const float4*)L.q4, L.qv, p->hyp.p,
                                                               p->counts.p, a, h0);
        pnp_select_kernel<<<1, 32, 0, c->stream>>>(p->d_sc, sets, p->counts.p, a, h0, h1,
                                                   k + 2 <= kPnpChunks ? bounds[k + 2] : H, last ? 1 : 0);
        c->launches += 2;
    }
    FinalizeBufs fb{};
    if (L.compact) {
        fb.P_in = p->P[p->cur].p; fb.X_in = p->X[p->cur].p;
        fb.P_out = p->P[p->cur ^ 1].p; fb.X_out = p->X[p->cur ^ 1].p;
    }
    pnp_finalize_kernel<<<1, kFinThreads, 0, c->stream>>>(p->d_sc, p->hyp.p, (const float4*)L.q4, L.qv, p->mask.p, a, fb);
    c->launches++;
    if (L.compact) p->cur ^= 1;
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

}  // namespace vo
