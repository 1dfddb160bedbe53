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
// cv::RNG is a multiply-with-carry generator: strictly sequential, but its raw 32-bit output
// stream does not depend on the data (the seed is fixed at 2^64-1 inside RANSAC's run()).  The
// host tabulates the stream once (rng_table_host); the device turns it into minimal sets with
// `% n` and the duplicate re-draws of getSubset, 32 sets per step.
struct RngTab { const uint32_t* tab; int len; };

// rows [sc->sets_ready, to) of the minimal-set table, exactly as getSubset draws them.
// Called by one full warp.  Optimistic step: lane l assumes sets h..h+l-1 needed no re-draw and
// reads draws pos+4l..pos+4l+3; the prefix of lanes for which that held is committed, the first
// set with a duplicate is drawn serially by lane 0.
__device__ void gen_sets_warp(DevScalars* sc, int32_t* sets, RngTab rt, int n, int to)
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

// ---- 4-lane cooperative root isolation -----------------------------------------------------------
// The serial chain of real_roots/p3p4 is what bounds a RANSAC chunk (every hypothesis CTA waits
// for its pose), so one hypothesis is solved by a quad of lanes: the <=4 bracketing intervals of a
// derivative level, and later the <=4 quartic roots, are independent and run one per lane.  The
// arithmetic per interval / per root is the same as in the serial code (same expressions, same
// order), so both give identical bits.
template <int DEG>
__device__ __forceinline__ double peval_t(const double* c, double x)
{
    double v = c[DEG];
#pragma unroll
    for (int i = DEG - 1; i >= 0; i--) v = v * x + c[i];
    return v;
}

// d = coefficients of the derivative of c (degree DEG-1)
template <int DEG>
__device__ __forceinline__ double bracket_root_t(const double* c, const double* d, double lo, double hi, double flo)
{
    double x = 0.5 * (lo + hi);
    for (int it = 0; it < 200; it++) {
        const double f = peval_t<DEG>(c, x);
        if (f == 0.) return x;
        if ((f < 0.) == (flo < 0.)) lo = x; else hi = x;
        const double df = peval_t<DEG - 1>(d, x);
        double xn = df != 0. ? x - f / df : 0.5 * (lo + hi);
        if (!(xn > lo && xn < hi)) xn = 0.5 * (lo + hi);
        if (fabs(xn - x) <= 4e-16 * fabs(x) || hi - lo <= 4e-16 * fabs(x)) return xn;
        x = xn;
    }
    return x;
}

__device__ __forceinline__ double sel4(int i, double a0, double a1, double a2, double a3)
{
    return i == 0 ? a0 : (i == 1 ? a1 : (i == 2 ? a2 : a3));
}

// one derivative level: lane q of the quad takes interval q of (lo, prev..., hi]; the roots found
// are gathered back, ascending, into prev[0..return)
template <int DEG>
__device__ __forceinline__ int quad_level(const double* p, const double* dp, double lo, double hi, double* prev,
                                          int nprev, int q, unsigned qmask)
{
    bool has = false;
    double root = 0.;
    if (q <= nprev) {
        const double a = q == 0 ? lo : sel4(q - 1, prev[0], prev[1], prev[2], prev[3]);
        const double b = q < nprev ? sel4(q, prev[0], prev[1], prev[2], prev[3]) : hi;
        const double fa = peval_t<DEG>(p, a), fb = peval_t<DEG>(p, b);
        if (b > a) {
            if (fb == 0.) { has = true; root = b; }
            else if (fa != 0. && (fa < 0.) != (fb < 0.)) { has = true; root = bracket_root_t<DEG>(p, dp, a, b, fa); }
        }
    }
    int n = 0;
#pragma unroll
    for (int s = 0; s < 4; s++) {
        const int hs = __shfl_sync(qmask, (int)has, s, 4);
        const double rs = __shfl_sync(qmask, root, s, 4);
        if (hs) {
            if (n == 0) prev[0] = rs; else if (n == 1) prev[1] = rs; else if (n == 2) prev[2] = rs; else prev[3] = rs;
            n++;
        }
    }
    return n;
}

// real roots of the quartic c[0..4] in (lo, hi], ascending, by a quad of lanes (all 4 lanes get
// the full result).  Same result as real_roots(c, 4, lo, hi, roots).
__device__ int real_roots4_quad(const double* cin, double lo, double hi, double* roots, int q, unsigned qmask)
{
    double big = 0.;
#pragma unroll
    for (int i = 0; i <= 4; i++) big = fmax(big, fabs(cin[i]));
    if (fabs(cin[4]) <= 1e-14 * big) {   // degenerate leading coefficient: rare, serial path on lane 0
        int nv = 0;
        double r[4] = { 0., 0., 0., 0. };
        if (q == 0) nv = real_roots(cin, 4, lo, hi, r);
        nv = __shfl_sync(qmask, nv, 0, 4);
#pragma unroll
        for (int i = 0; i < 4; i++) roots[i] = __shfl_sync(qmask, r[i], 0, 4);
        return nv;
    }
    // Cauchy bound: every (complex) root, hence by Gauss-Lucas every critical point of every
    // derivative, has modulus < 1 + max|c_i/c_4|.  Shrinking the search interval from 1e6 to the
    // bound keeps the root set and saves the ~40 Newton steps a start at 5e5 spends coming down.
    {
        double m = 0.;
#pragma unroll
        for (int i = 0; i < 4; i++) m = fmax(m, fabs(cin[i]));
        hi = fmin(hi, 1. + m / fabs(cin[4]));
    }
    double d1[4], d2[3], d3[2];
#pragma unroll
    for (int i = 1; i <= 4; i++) d1[i - 1] = cin[i] * i;
#pragma unroll
    for (int i = 1; i <= 3; i++) d2[i - 1] = d1[i] * i;
#pragma unroll
    for (int i = 1; i <= 2; i++) d3[i - 1] = d2[i] * i;
    double prev[4] = { 0., 0., 0., 0. };
    int nprev = 0;
    {
        const double x = -d3[0] / d3[1];
        if (x > lo && x <= hi) prev[nprev++] = x;
    }
    nprev = quad_level<2>(d2, d3, lo, hi, prev, nprev, q, qmask);
    nprev = quad_level<3>(d1, d2, lo, hi, prev, nprev, q, qmask);
    nprev = quad_level<4>(cin, d1, lo, hi, prev, nprev, q, qmask);
#pragma unroll
    for (int i = 0; i < 4; i++) roots[i] = prev[i];
    return nprev;
}

// P3P (Grunert quartic; roots isolated by bracketing; distances polished by Gauss-Newton; exact
// 3-point alignment).  X: 4 world points (rows), p: 4 pixel points.  Roots come from the first
// three correspondences, the fourth picks the root with the smallest pixel error.  Called by the
// 4 lanes of a quad (q = lane in quad) with identical inputs; returns true on the ONE lane that
// holds the selected solution in Rt = [R|t] (12 doubles: R row-major then t); *found is set on all
// four lanes.
__device__ bool p3p4_quad(const double* X, const double* p, const double* K, double* Rt, int q, unsigned qmask,
                          bool* found_any)
{
    *found_any = false;
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
    const double qq = (a2 - c2) / b2, k = c2 / b2;
    const double N[3] = { 1. + qq, -2. * qq * cb, qq - 1. };
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
    const int nv = real_roots4_quad(Q, 0., 1e6, vr, q, qmask);

    // lane q evaluates root q
    bool mine = false;
    double e = 0.;
    do {
        if (q >= nv) break;
        const double v = sel4(q, vr[0], vr[1], vr[2], vr[3]);
        const double Wv = 1. + v * (v - 2. * cb);
        if (!(Wv > 0.)) break;
        const double Dv = D[0] + D[1] * v;
        double u;
        if (fabs(Dv) > 1e-9) {
            u = (N[0] + v * (N[1] + v * N[2])) / Dv;
        } else {
            const double cc = 1. - k * Wv, disc = cg * cg - cc;
            if (!(disc >= 0.)) break;
            const double sq = sqrt(disc), u0 = cg + sq, u1 = cg - sq;
            const double e0 = fabs(u0 * u0 + v * v - 2. * u0 * v * ca - (a2 / b2) * Wv);
            const double e1 = fabs(u1 * u1 + v * v - 2. * u1 * v * ca - (a2 / b2) * Wv);
            u = e0 <= e1 ? u0 : u1;
        }
        if (!(u > 0.)) break;
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
        if (!(s1 > 0. && s2 > 0. && s3 > 0.)) break;
        double e12[3], e13[3], nY[3], Y1[3];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            Y1[i] = s1 * j[i];
            e12[i] = Y1[i] - s2 * j[3 + i];
            e13[i] = Y1[i] - s3 * j[6 + i];
        }
        cross3(e12, e13, nY);
        bool fin = true;
#pragma unroll
        for (int i = 0; i < 3; i++) {
#pragma unroll
            for (int c = 0; c < 3; c++) {
                Rt[3 * i + c] = e12[i] * XmInv[c] + e13[i] * XmInv[3 + c] + nY[i] * XmInv[6 + c];
                fin &= isfinite(Rt[3 * i + c]);
            }
            Rt[9 + i] = Y1[i] - (Rt[3 * i] * X1[0] + Rt[3 * i + 1] * X1[1] + Rt[3 * i + 2] * X1[2]);
        }
        if (!fin) break;
        // 4th point: pixel reprojection error (float64)
        const double* X4 = X + 9;
        double x = Rt[0] * X4[0] + Rt[1] * X4[1] + Rt[2] * X4[2] + Rt[9];
        double y = Rt[3] * X4[0] + Rt[4] * X4[1] + Rt[5] * X4[2] + Rt[10];
        double z = Rt[6] * X4[0] + Rt[7] * X4[1] + Rt[8] * X4[2] + Rt[11];
        z = z != 0. ? 1. / z : 1.;
        const double du = x * z * K[0] + K[2] - p[6], dv = y * z * K[4] + K[5] - p[7];
        e = du * du + dv * dv;
        mine = true;
    } while (false);
    // the serial loop keeps the first root (ascending) with the strictly smallest error
    int win = -1;
    double beste = 0.;
#pragma unroll
    for (int s = 0; s < 4; s++) {
        const int ms = __shfl_sync(qmask, (int)mine, s, 4);
        const double es = __shfl_sync(qmask, e, s, 4);
        if (ms && (win < 0 || es < beste)) { win = s; beste = es; }
    }
    *found_any = win >= 0;
    return win == q;
}

// ---- kernels -----------------------------------------------------------------------------------
__global__ void pnp_init_kernel(DevScalars* sc, int32_t* sets, const int32_t* d_n, PnpArgs a, RngTab rt, int chunk0_end)
{
    // one warp
    const int n = d_n ? min(*d_n, a.n_upper) : a.n_upper;
    __syncwarp();
    if (threadIdx.x == 0) {
        sc->n_tracked = n;
        sc->niters = a.max_iters; sc->best_count = 0; sc->best_hyp = -1; sc->iters_run = 0; sc->done = 0;
        sc->sets_ready = 0; sc->rng = 0; sc->ok = 0; sc->n_inliers = 0; sc->lm_iters = 0; sc->rng_exhausted = 0;
        if (n < kMinimal || a.max_iters <= 0) { sc->niters = 0; sc->done = 1; }
        else if (n == kMinimal) {   // cv2: count == modelPoints -> one direct solve on all points
            reinterpret_cast<int4*>(sets)[0] = make_int4(0, 1, 2, 3);
            sc->niters = 1; sc->sets_ready = a.max_iters;
        }
    }
    __syncwarp();
    if (n > kMinimal && a.max_iters > 0 && !a.explicit_sets) gen_sets_warp(sc, sets, rt, n, chunk0_end);
}

// one CTA per hypothesis: lanes 0-3 solve P3P, the CTA scores all points
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
    if (threadIdx.x < 4) {   // a quad of lanes solves the minimal problem
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
        bool found = false;
        const bool winner = ok && p3p4_quad(X, p, a.K, Rt, threadIdx.x, 0xFu, &found);
        if (threadIdx.x == 0) sValid = ok && found;
        if (winner) {
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
                                  RngTab rt, int h0, int h1, int next_end, int last)
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
    }
    __syncwarp();
    if (!done && !a.explicit_sets && next_end > sc->sets_ready && n > kMinimal) gen_sets_warp(sc, sets, rt, n, next_end);
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

// solves (L L^T) x = b for the SPD 6x6 A (lower triangle used), x returned in b.  Fully unrolled
// (registers only); one rsqrt per column is the only special op on the dependent chain.
__device__ __forceinline__ bool chol6_solve(double* A, double* b)
{
    double inv[6];
#pragma unroll
    for (int j = 0; j < 6; j++) {
        double s = A[6 * j + j];
#pragma unroll
        for (int k = 0; k < j; k++) s -= A[6 * j + k] * A[6 * j + k];
        if (!(s > 0.)) return false;
        inv[j] = rsqrt(s);
#pragma unroll
        for (int i = j + 1; i < 6; i++) {
            double t = A[6 * i + j];
#pragma unroll
            for (int k = 0; k < j; k++) t -= A[6 * i + k] * A[6 * j + k];
            A[6 * i + j] = t * inv[j];
        }
    }
#pragma unroll
    for (int i = 0; i < 6; i++) {
        double s = b[i];
#pragma unroll
        for (int k = 0; k < i; k++) s -= A[6 * i + k] * b[k];
        b[i] = s * inv[i];
    }
#pragma unroll
    for (int i = 5; i >= 0; i--) {
        double s = b[i];
#pragma unroll
        for (int k = i + 1; k < 6; k++) s -= A[6 * k + i] * b[k];
        b[i] = s * inv[i];
    }
    return true;
}

constexpr int kAcc = 28;  // 21 (lower triangle of JtJ) + 6 (Jt r) + 1 (cost)

// accumulate JtJ, Jt r and the cost at pose Rt over the masked points; result in sAcc (all threads
// see it after the trailing __syncthreads)
__device__ void lm_accumulate(const double* Rt, const double* K, const float4* __restrict__ q4,
                              const float* __restrict__ qv, const uint8_t* __restrict__ mask, int n,
                              double (*sPart)[kAcc], double* sAcc)
{
    double acc[32];   // kAcc used, padded to the warp width for the halving reduction
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = 0.;
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
    // warp reduction by recursive halving: after the 5 exchange steps lane L holds the warp total
    // of element L (31 shuffles instead of 28 x 5)
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int k = 0; k < off; k++) {
            const double send = upper ? acc[k] : acc[k + off];
            const double keep = upper ? acc[k + off] : acc[k];
            acc[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    __syncthreads();   // previous readers of sPart/sAcc are done
    if (lane < kAcc) sPart[threadIdx.x >> 5][lane] = acc[0];
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
                    // a step below 1e-11 (rad / scene units) cannot move the pose at the 1e-9 level:
                    // converged, no trial evaluation needed
                    sState = step < 1e-22 ? 1 : 0;
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
                    sState = (sStep < 1e-22 || dec <= 1e-13 * cn) ? 1 : 0;   // 1 = converged
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

// raw outputs of cv::RNG(2^64-1) in draw order; 16 draws per iteration + slack covers the re-draws
// of getSubset with overwhelming margin (4 + O(1/n) are used); if it ever runs out the device sets
// rng_exhausted and the host reports VO_ERR_CAPACITY
static int ensure_rng_table(Ctx* c, int H)
{
    Pipeline* p = c->pipe;
    const size_t need = (size_t)(H > 0 ? H : 1) * 16 + 1024;
    if ((size_t)p->rng_tab_len >= need) return VO_OK;
    size_t len = 1 << 16;
    while (len < need) len *= 2;
    std::vector<uint32_t> tab(len);
    uint64_t s = 0xFFFFFFFFFFFFFFFFULL;
    for (size_t i = 0; i < len; i++) {
        s = (uint64_t)(unsigned)s * 4164903690ULL + (unsigned)(s >> 32);
        tab[i] = (uint32_t)s;
    }
    VO_TRY(ensure(c, p->rng_tab, len));
    VO_CUDA(c, cudaMemcpyAsync(p->rng_tab.p, tab.data(), len * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    VO_CUDA(c, cudaStreamSynchronize(c->stream));   // `tab` is a local
    p->rng_tab_len = (int)len;
    return VO_OK;
}

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
    RngTab rt{ nullptr, 0 };
    if (!L.d_sets_in) {
        VO_TRY(ensure_rng_table(c, H));
        rt.tab = p->rng_tab.p; rt.len = p->rng_tab_len;
    }
    // chunk boundaries: small first chunk (the adaptive stop usually ends inside it)
    int bounds[kPnpChunks + 1] = { 0, H < 128 ? H : 128, H < 512 ? H : 512, H };
    pnp_init_kernel<<<1, 32, 0, c->stream>>>(p->d_sc, sets, L.d_n, a, rt, bounds[1]);
    c->launches++;
    for (int k = 0; k < kPnpChunks; k++) {
        const int h0 = bounds[k], h1 = bounds[k + 1];
        if (h1 <= h0) continue;
        const bool last = (h1 == H);
        pnp_hyp_kernel<<<h1 - h0, kHypThreads, 0, c->stream>>>(p->d_sc, sets, (const float4*)L.q4, L.qv, p->hyp.p,
                                                               p->counts.p, a, h0);
        pnp_select_kernel<<<1, 32, 0, c->stream>>>(p->d_sc, sets, p->counts.p, a, rt, h0, h1,
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
