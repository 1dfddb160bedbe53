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
#include <stdlib.h>
#include "ctx.cuh"
#include "pnp_sets.cuh"

namespace vo {

namespace {

constexpr int kHypThreads = 256;      // threads of a RANSAC CTA (hypothesis scoring, and the refit by the last CTA)
constexpr int kFinThreads = 512;            // threads of the refit kernel

struct PnpArgs {
    double K[9];
    double confidence;
    float thr2;          // (float)(thr*thr)
    int max_iters;
    int refine_iters;
    int n_upper;
    int explicit_sets;
    int compact;
    int f32_filter;      // score through the float32 filter (exact float64 only near the threshold)
};

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

// The inlier test has to be cv2's, bit for bit (float64 projection rounded to float32, float32 compare), but almost
// no (hypothesis, point) pair is anywhere near the threshold.  A float32 evaluation with a rigorous-enough error bound
// decides those: the same projection in FMA float32 (R, t rounded to float32; the points ARE float32), and a band
// around the threshold that covers its error - input rounding and three accumulations, ~2.5e-7 x (3 max|X| + max|t|)
// in camera space, times the projection's gain (fx + |u - cx|) / |z| - four times over, plus 0.02 px.  Pairs inside
// the band, with |z| < 0.5 or a non-finite result are "uncertain" and go to the exact float64 path.
struct HypF32 { float R[9], t[3], mt; };   // mt = max |t|

__device__ __forceinline__ int filter_err2(const HypF32& hf, const float* Kf, float thr, float X, float Y, float Z, float u, float v)
{
    const float x = fmaf(hf.R[0], X, fmaf(hf.R[1], Y, fmaf(hf.R[2], Z, hf.t[0])));
    const float y = fmaf(hf.R[3], X, fmaf(hf.R[4], Y, fmaf(hf.R[5], Z, hf.t[1])));
    const float z = fmaf(hf.R[6], X, fmaf(hf.R[7], Y, fmaf(hf.R[8], Z, hf.t[2])));
    const float iz = __frcp_rn(z);
    const float pu = x * iz * Kf[0], pv = y * iz * Kf[1];          // relative to the principal point
    const float du = (u - Kf[2]) - pu, dv = (v - Kf[3]) - pv;
    const float e = fmaf(du, du, dv * dv);
    const float mp = fmaxf(fmaxf(fabsf(X), fabsf(Y)), fabsf(Z));
    const float err_m = 1e-6f * fmaf(3.f, mp, hf.mt);               // 4 x 2.5e-7 x (3 max|X| + max|t|)
    const float band = fmaf(err_m * fabsf(iz), Kf[0] + Kf[1] + fabsf(pu) + fabsf(pv), 0.02f);
    const float lo = fmaxf(thr - band, 0.f), hi = thr + band;
    if (!(fabsf(z) > 0.5f) || !(e < 1e30f) || (e >= lo * lo && e <= hi * hi)) return -1;   // uncertain
    return e <= thr * thr ? 1 : 0;
}

// ---- P3P ---------------------------------------------------------------------------------------
// The serial chain of the minimal solver bounds a RANSAC chunk (every hypothesis CTA waits for
// its pose, and a lone warp retires one dependent FP64 instruction every ~14 cycles), so the
// solver is chosen for the shortest dependent chain and spread over a quad of lanes:
//   - depths l1,l2,l3 along the unit bearings satisfy  li^2 + bij li lj + lj^2 = aij.  The two
//     homogeneous conics D1 = a23 M12 - a12 M23, D2 = a23 M13 - a13 M23 share the solutions; one
//     real root g of det(D1 + g D2) = 0 (a cubic: Newton from the stationary-point expansion, no
//     bracketing) gives a degenerate conic = a pair of planes (e1 +- s e2).L = 0, each of which
//     turns one distance equation into a quadratic in tau = l3/l2 ("lambda twist", Persson &
//     Nordberg 2018): at most 4 depth triples with no quartic to isolate;
//   - lane q of the quad takes (plane q>>1, quadratic root q&1), polishes its triple by
//     Gauss-Newton on the three distance equations (so the result does not depend on how the
//     start was obtained: it agrees with the Grunert-quartic oracle to rounding), aligns the
//     triangles exactly and scores the 4th point; the quad keeps the smallest pixel error.

// one real root of x^3 + b x^2 + c x + d
__device__ __forceinline__ double cubic_root(double b, double c, double d)
{
    double r0;
    const double disc = b * b - 3. * c;
    if (disc >= 0.) {   // start beyond the outer root on the convex side of a stationary point
        const double v = sqrt(disc);
        const double t1 = (-b - v) * (1. / 3.);
        double k = ((t1 + b) * t1 + c) * t1 + d;
        if (k > 0.) r0 = t1 - sqrt(-k / (3. * t1 + b));
        else {
            const double t2 = (-b + v) * (1. / 3.);
            k = ((t2 + b) * t2 + c) * t2 + d;
            r0 = t2 + sqrt(-k / (3. * t2 + b));
        }
    } else {
        r0 = -b * (1. / 3.);
        if (fabs((3. * r0 + 2. * b) * r0 + c) < 1e-4) r0 += 1.;
    }
    for (int it = 0; it < 50; it++) {
        const double fx = ((r0 + b) * r0 + c) * r0 + d;
        const double fpx = (3. * r0 + 2. * b) * r0 + c;
        if (fpx == 0.) break;
        const double dx = fx / fpx;
        r0 -= dx;
        if (fabs(dx) <= 1e-15 * fabs(r0)) break;
    }
    return r0;
}

// unit vector spanning the null space of the symmetric rank-2 matrix M - s I (largest of the
// three row cross products)
__device__ __forceinline__ void null_vector(const double* M, double s, double* e)
{
    const double r0[3] = { M[0] - s, M[1], M[2] }, r1[3] = { M[1], M[3] - s, M[4] }, r2[3] = { M[2], M[4], M[5] - s };
    double c0[3], c1[3], c2[3];
    cross3(r0, r1, c0); cross3(r0, r2, c1); cross3(r1, r2, c2);
    const double n0 = dot3(c0, c0), n1 = dot3(c1, c1), n2 = dot3(c2, c2);
    const bool u1 = n1 > n0 && n1 >= n2, u2 = n2 > n0 && n2 > n1;
    const double n = u2 ? n2 : (u1 ? n1 : n0);
    const double inv = rsqrt(n);
#pragma unroll
    for (int i = 0; i < 3; i++) e[i] = (u2 ? c2[i] : (u1 ? c1[i] : c0[i])) * inv;
}

__device__ __forceinline__ double det3(const double* a, const double* b, const double* c)
{
    return a[0] * (b[1] * c[2] - b[2] * c[1]) - a[1] * (b[0] * c[2] - b[2] * c[0]) + a[2] * (b[0] * c[1] - b[1] * c[0]);
}

// X: 4 world points (rows), p: 4 pixel points.  Solutions come from the first three
// correspondences, the fourth picks the one with the smallest pixel error.  Called by the 4 lanes
// of a quad (q = lane in quad) with identical inputs; returns true on the ONE lane that holds the
// selected solution in Rt = [R|t] (12 doubles: R row-major then t); *found_any is set on all four.
__device__ bool p3p4_quad(const double* X, const double* p, const double* K, double* Rt, int q, unsigned qmask,
                          bool* found_any)
{
    *found_any = false;
    double j[9];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const double x = (p[2 * i] - K[2]) / K[0], y = (p[2 * i + 1] - K[5]) / K[4];
        const double n = rsqrt(x * x + y * y + 1.);
        j[3 * i] = x * n; j[3 * i + 1] = y * n; j[3 * i + 2] = n;
    }
    const double* X1 = X; const double* X2 = X + 3; const double* X3 = X + 6;
    double d12[3], d13[3], d23[3];
#pragma unroll
    for (int i = 0; i < 3; i++) { d12[i] = X1[i] - X2[i]; d13[i] = X1[i] - X3[i]; d23[i] = X2[i] - X3[i]; }
    const double a12 = dot3(d12, d12), a13 = dot3(d13, d13), a23 = dot3(d23, d23);
    if (!(a23 > 0. && a13 > 0. && a12 > 0.)) return false;
    const double c12 = dot3(j, j + 3), c13 = dot3(j, j + 6), c23 = dot3(j + 3, j + 6);
    const double b12 = -2. * c12, b13 = -2. * c13, b23 = -2. * c23;
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
    // rows of the symmetric conics
    const double A0[3] = { a23, a23 * b12 * .5, 0. }, A1[3] = { a23 * b12 * .5, a23 - a12, -a12 * b23 * .5 },
                 A2[3] = { 0., -a12 * b23 * .5, -a12 };
    const double B0[3] = { a23, 0., a23 * b13 * .5 }, B1[3] = { 0., -a13, -a13 * b23 * .5 },
                 B2[3] = { a23 * b13 * .5, -a13 * b23 * .5, a23 - a13 };
    const double k0 = det3(A0, A1, A2);
    const double k1 = det3(B0, A1, A2) + det3(A0, B1, A2) + det3(A0, A1, B2);
    const double k2 = det3(B0, B1, A2) + det3(B0, A1, B2) + det3(A0, B1, B2);
    const double k3 = det3(B0, B1, B2);
    if (!(fabs(k3) > 0.) || !isfinite(k3)) return false;
    const double ik3 = 1. / k3;
    const double g = cubic_root(k2 * ik3, k1 * ik3, k0 * ik3);
    // M = D1 + g D2, upper triangle (00 01 02 11 12 22)
    const double M[6] = { A0[0] + g * B0[0], A0[1] + g * B0[1], A0[2] + g * B0[2],
                          A1[1] + g * B1[1], A1[2] + g * B1[2], A2[2] + g * B2[2] };
    const double tr = M[0] + M[3] + M[5];
    const double m2 = M[0] * M[3] - M[1] * M[1] + M[0] * M[5] - M[2] * M[2] + M[3] * M[5] - M[4] * M[4];
    const double dsc = tr * tr - 4. * m2;
    if (!(dsc >= 0.)) return false;
    const double sq = sqrt(dsc);
    const double s1 = .5 * (tr + (tr >= 0. ? sq : -sq));   // larger-magnitude eigenvalue first
    if (s1 == 0.) return false;
    const double s2 = m2 / s1;
    if (!(s1 * s2 < 0.)) return false;   // a real plane pair needs eigenvalues of opposite sign
    double e1[3], e2[3];
    null_vector(M, s1, e1);
    null_vector(M, s2, e2);
    const double sv = sqrt(-s2 / s1);

    // lane q: plane (q >> 1), quadratic root (q & 1)
    bool mine = false;
    double e = 0.;
    do {
        const double ss = (q & 2) ? -sv : sv;
        const double u0 = e1[0] + ss * e2[0], u1 = e1[1] + ss * e2[1], u2 = e1[2] + ss * e2[2];
        if (!(fabs(u0) > 0.)) break;
        const double iu = -1. / u0;
        const double w0 = u1 * iu, w1 = u2 * iu;   // l1 = w0 l2 + w1 l3
        const double qa = (a13 - a12) * w1 * w1 - a12 * b13 * w1 - a12;
        const double qb = a13 * b12 * w1 - a12 * b13 * w0 - 2. * w0 * w1 * (a12 - a13);
        const double qc = (a13 - a12) * w0 * w0 + a13 * b12 * w0 + a13;
        double tau;
        if (fabs(qa) > 0.) {
            const double dd = qb * qb - 4. * qa * qc;
            if (!(dd >= 0.)) break;
            const double sd = sqrt(dd), q2 = -.5 * (qb + (qb >= 0. ? sd : -sd));
            if (q & 1) { if (q2 == 0.) break; tau = qc / q2; }
            else tau = q2 / qa;
        } else {
            if ((q & 1) || !(fabs(qb) > 0.)) break;
            tau = -qc / qb;
        }
        if (!(tau > 0.)) break;
        const double den = tau * (b23 + tau) + 1.;
        if (!(den > 0.)) break;
        double l2 = sqrt(a23 / den), l3 = tau * l2, l1 = w0 * l2 + w1 * l3;
        if (!(l1 > 0.)) break;
        for (int it = 0; it < 5; it++) {
            const double r12 = l1 * l1 + l2 * l2 - 2. * l1 * l2 * c12 - a12;
            const double r13 = l1 * l1 + l3 * l3 - 2. * l1 * l3 * c13 - a13;
            const double r23 = l2 * l2 + l3 * l3 - 2. * l2 * l3 * c23 - a23;
            const double J00 = 2. * (l1 - l2 * c12), J01 = 2. * (l2 - l1 * c12);
            const double J10 = 2. * (l1 - l3 * c13), J12 = 2. * (l3 - l1 * c13);
            const double J21 = 2. * (l2 - l3 * c23), J22 = 2. * (l3 - l2 * c23);
            const double det = -J00 * J12 * J21 - J01 * J10 * J22;
            if (!(fabs(det) > 1e-300)) break;
            const double id = 1. / det;
            const double dl1 = (-J12 * J21 * r12 - J01 * J22 * r13 + J01 * J12 * r23) * id;
            const double dl2 = (-J10 * J22 * r12 + J00 * J22 * r13 - J00 * J12 * r23) * id;
            const double dl3 = (J10 * J21 * r12 - J00 * J21 * r13 - J01 * J10 * r23) * id;
            l1 -= dl1; l2 -= dl2; l3 -= dl3;
            if (fabs(dl1) + fabs(dl2) + fabs(dl3) <= 1e-15 * (l1 + l2 + l3)) break;
        }
        if (!(l1 > 0. && l2 > 0. && l3 > 0.)) break;
        double e12[3], e13[3], nY[3], Y1[3];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            Y1[i] = l1 * j[i];
            e12[i] = Y1[i] - l2 * j[3 + i];
            e13[i] = Y1[i] - l3 * j[6 + i];
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

// ---- RANSAC bookkeeping ------------------------------------------------------------------------
// One warp: the sequential rule of the reference loop (first strictly-better model wins, the
// iteration budget shrinks with every new best) replayed over the counts of hypotheses [h0, h1).
// `first` initialises the run; when the budget is not exhausted the minimal sets of the next
// chunk are drawn.
__device__ void ransac_select(DevScalars* sc, int32_t* sets, const int32_t* counts, const PnpArgs& a, RngTab rt,
                              int n, int first, int h0, int h1, int next_end, int last)
{
    const int lane = threadIdx.x & 31;
    int niters, best, best_hyp, done;
    if (first) {
        niters = (n < kMinimal || a.max_iters <= 0) ? 0 : (n == kMinimal ? 1 : a.max_iters);
        best = 0; best_hyp = -1; done = niters == 0;
        if (lane == 0) sc->finalized = 0;
    } else {
        niters = sc->niters; best = sc->best_count; best_hyp = sc->best_hyp; done = sc->done;
    }
    if (done) {
        if (lane == 0 && first) { sc->niters = 0; sc->best_count = 0; sc->best_hyp = -1; sc->done = 1; sc->iters_run = 0; }
        return;
    }
    int h = h0;
    while (h < h1) {
        if (h >= niters) { done = 1; break; }
        const int lim = min(h1, niters);
        const int c = (h + lane < lim) ? __ldcg(counts + h + lane) : -1;
        const unsigned m = __ballot_sync(0xffffffffu, c > max(best, kMinimal - 1));
        if (!m) { h += 32; continue; }
        const int f = __ffs(m) - 1;
        best = __shfl_sync(0xffffffffu, c, f);
        best_hyp = h + f;
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
    next_end = min(next_end, niters);
    if (!done && !a.explicit_sets && next_end > sc->sets_ready && n > kMinimal) gen_sets_warp(sc, sets, rt, n, next_end);
}

// landmark buffers the pruning reads and writes (vo_pipeline.py:327-331)
struct FinalizeBufs {
    const float* P_in; const double* X_in; float* P_out; double* X_out;   // landmark rotation
    float4* q4i; float* qvi;                                              // the inliers, packed for the refit
    int32_t* src_out;   // per kept row: its row before the pruning (= its track when the next KLT ran ahead, Pipeline::Lsrc)
};

__device__ __forceinline__ void so3_exp(const double* w, double* R)
{
    const double th2 = dot3(w, w);
    double A, B;
    if (th2 < 1e-4) {   // LM steps are small: series to th^8 (error < 1e-21), no sqrt / sin / cos on the chain
        A = 1. + th2 * (-1. / 6. + th2 * (1. / 120. + th2 * (-1. / 5040. + th2 * (1. / 362880.))));
        B = 0.5 + th2 * (-1. / 24. + th2 * (1. / 720. + th2 * (-1. / 40320. + th2 * (1. / 3628800.))));
    } else {
        const double th = sqrt(th2);
        A = sin(th) / th; B = (1. - cos(th)) / th2;
    }
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

// accumulate JtJ, Jt r and the cost at pose Rt over the n packed inliers; result in sAcc (all threads
// see it after the trailing __syncthreads)
__device__ void lm_accumulate(const double* Rt, const double* K, const float4* __restrict__ q4,
                              const float* __restrict__ qv, int n, double (*sPart)[kAcc], double* sAcc)
{
    double acc[32];   // kAcc used, padded to the warp width for the halving reduction
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = 0.;
    for (int i = threadIdx.x; i < n; i += kFinThreads) {
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

__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ---- the refit kernel ------------------------------------------------------------------------------
// LM on SE(3) over the winner's inlier set, from the winning hypothesis to the converged optimum (what
// cv2.solvePnP(SOLVEPNP_ITERATIVE) returns on that set).  One CTA; separate from the RANSAC kernel because
// nothing on the landmark side waits for it: the pruned landmarks are final when the RANSAC kernel ends,
// only the pose (host read-back, DLT of the promoted candidates) comes from here.
__global__ void __launch_bounds__(kFinThreads)
pnp_refit_kernel(DevScalars* sc, const float4* __restrict__ q4i, const float* __restrict__ qvi, PnpArgs a)
{
    __shared__ double sPart[kFinThreads / 32][kAcc];
    __shared__ double sAcc[kAcc];
    __shared__ double sRt[12], sTrial[12], sH[21], sG[6];
    __shared__ double sCost, sLambda, sStep;
    __shared__ int sFails, sIters;
    __shared__ int sState;   // 0 = continue with trial, 1 = stop
    pdl_prologue();
    const int tid = threadIdx.x;
    if (tid == 0) sc->stamps[4] = gtime();
    if (!sc->ok) {
        if (tid < 16) sc->T[tid] = 0.;
        if (tid == 0) { sc->lm_iters = 0; sc->cost = 0.; sc->stamps[5] = gtime(); }
        return;
    }
    const int n = sc->n_inliers;
    if (tid < 12) sRt[tid] = sc->Th[tid];
    __syncthreads();
    // ---- evaluate at the current pose, then trial steps ---------------------------------------------
    lm_accumulate(sRt, a.K, q4i, qvi, n, sPart, sAcc);
    if (tid == 0) {
        for (int i = 0; i < 21; i++) sH[i] = sAcc[i];
        for (int i = 0; i < 6; i++) sG[i] = sAcc[21 + i];
        sCost = sAcc[27]; sLambda = 1e-6; sState = 0; sFails = 0; sIters = 0;
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
                // a step below 1e-9 (rad / scene units) cannot move the pose at the level any gate
                // looks at: converged, no trial evaluation needed
                sState = step < 1e-18 ? 1 : 0;
            }
        }
        __syncthreads();
        if (sState) break;
        lm_accumulate(sTrial, a.K, q4i, qvi, n, sPart, sAcc);
        if (tid == 0) {
            const double cn = sAcc[27];
            if (cn <= sCost) {
                const double dec = sCost - cn;
                for (int i = 0; i < 12; i++) sRt[i] = sTrial[i];
                for (int i = 0; i < 21; i++) sH[i] = sAcc[i];
                for (int i = 0; i < 6; i++) sG[i] = sAcc[21 + i];
                sCost = cn; sLambda = fmax(sLambda * 0.1, 1e-12);
                sFails = 0; sIters++;
                // Gauss-Newton contracts the error by orders of magnitude per step: after an accepted
                // step below 1e-7 the next one would be far below 1e-9
                sState = (sStep < 1e-14 || dec <= 1e-13 * cn) ? 1 : 0;   // 1 = converged
            } else if (sStep < 1e-14) {
                sState = 1;   // a rejected step this small: the cost is at its rounding floor
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
    if (tid == 0) {
        for (int r = 0; r < 3; r++) {
            for (int cc = 0; cc < 3; cc++) { sc->T[4 * r + cc] = sRt[3 * r + cc]; }
            sc->T[4 * r + 3] = sRt[9 + r];
        }
        sc->T[12] = 0.; sc->T[13] = 0.; sc->T[14] = 0.; sc->T[15] = 1.;
        sc->cost = sCost; sc->lm_iters = sIters;
        sc->stamps[5] = gtime();
    }
}

// Last CTA of the RANSAC kernel: inlier mask of the winner, pruning of the landmark buffers to the inliers
// (vo_pipeline.py:327-331; order preserved; no model: all rows kept) and the float32 inlier pack the refit
// reads - one pass over the points.
__device__ void pnp_mask_prune(DevScalars* sc, const double* hyp, const float4* __restrict__ q4,
                               const float* __restrict__ qv, uint8_t* mask, const PnpArgs& a, const FinalizeBufs& fb, int n)
{
    __shared__ double sRtF[12];
    __shared__ int sWarpKeep[kHypThreads / 32], sWarpInl[kHypThreads / 32];
    const int tid = threadIdx.x;
    const int best = sc->best_hyp;
    const bool ok = best >= 0;
    if (ok && tid < 12) { const double v = __ldcg(hyp + (size_t)best * 12 + tid); sRtF[tid] = v; sc->Th[tid] = v; }
    __syncthreads();
    double R[9], t[3];
#pragma unroll
    for (int i = 0; i < 9; i++) R[i] = ok ? sRtF[i] : 0.;
#pragma unroll
    for (int i = 0; i < 3; i++) t[i] = ok ? sRtF[9 + i] : 0.;
    int base_keep = 0, base_inl = 0;
    for (int i0 = 0; i0 < n; i0 += kHypThreads) {
        const int i = i0 + tid;
        const bool in = i < n;
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        float v = 0.f;
        if (in) { q = q4[i]; v = qv[i]; }
        const bool inl = in && ok && (n == kMinimal || reproj_err2(R, t, a.K, q.x, q.y, q.z, q.w, v) <= a.thr2);
        const bool keep = in && (!ok || inl);
        if (in) mask[i] = inl;
        double x0 = 0., x1 = 0., x2 = 0.;
        if (keep && a.compact) {   // issued before the barrier: the load latency overlaps the rank computation
            x0 = fb.X_in[3 * (size_t)i]; x1 = fb.X_in[3 * (size_t)i + 1]; x2 = fb.X_in[3 * (size_t)i + 2];
        }
        const unsigned bk = __ballot_sync(0xffffffffu, keep), bi = __ballot_sync(0xffffffffu, inl);
        const unsigned below = (1u << (tid & 31)) - 1;
        if ((tid & 31) == 0) { sWarpKeep[tid >> 5] = __popc(bk); sWarpInl[tid >> 5] = __popc(bi); }
        __syncthreads();
        int ok_ = base_keep, oi_ = base_inl, tk = 0, ti = 0;
#pragma unroll
        for (int w = 0; w < kHypThreads / 32; w++) {
            const int ck = sWarpKeep[w], ci = sWarpInl[w];
            if (w < (tid >> 5)) { ok_ += ck; oi_ += ci; }
            tk += ck; ti += ci;
        }
        if (keep && a.compact) {
            const int o = ok_ + __popc(bk & below);
            reinterpret_cast<float2*>(fb.P_out)[o] = make_float2(q.w, v);   // (u, v) as the tracking step left them
            fb.X_out[3 * (size_t)o] = x0; fb.X_out[3 * (size_t)o + 1] = x1; fb.X_out[3 * (size_t)o + 2] = x2;
            if (fb.src_out) fb.src_out[o] = i;
        }
        if (inl) {
            const int o = oi_ + __popc(bi & below);
            fb.q4i[o] = q; fb.qvi[o] = v;
        }
        base_keep += tk; base_inl += ti;
        __syncthreads();
    }
    if (tid == 0) {
        sc->ok = ok; sc->n_inliers = base_inl;
        if (a.compact) sc->n_land = base_keep;
    }
}

__device__ __forceinline__ int ld_acquire(const int* p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- the RANSAC kernel -----------------------------------------------------------------------------
// ONE persistent launch for the whole run (all CTAs resident: grid <= kPnpChunk0 <= number of SMs, one
// CTA per SM by registers).  The run proceeds in rounds; in a round every CTA takes B consecutive hypotheses:
// its warps solve the P3Ps (a quad of lanes each), then the whole CTA scores the points against them, up to
// kScoreTile hypotheses per pass over the points.  The last CTA to finish a round (ticket) replays the
// sequential bookkeeping of the reference loop (first strictly-better model wins, the budget shrinks) over
// the round's counts and opens the next round (epoch flag the others spin on) or ends the run: then it writes
// the inlier mask, prunes the landmarks and packs the inliers for the refit kernel.
//   round 0: the first kPnpChunk0 hypotheses, spread over the CTAs (the adaptive stop usually ends the run inside it);
//            one per CTA, or - PnpLaunch::small_grid - 4 per CTA on kPnpGridSmall CTAs: their P3Ps run side by side
//            on 4 warps and ONE pass over the points scores all 4 (16.5 us against 14.2 us for the round), and three
//            quarters of the SMs stay free for the next frame's KLT, which runs beside this kernel;
//   later rounds: B = ceil(remaining budget / CTAs), at most kBatch - a run with a pinned budget of 4096 is two rounds.
constexpr int kBatch = 32;       // hypotheses per CTA per round (P3P quads side by side: 128 threads)
constexpr int kScoreTile = 4;    // hypotheses scored per pass over the points
constexpr int kQueue = 512;      // per warp: (hypothesis, point) pairs waiting for the exact float64 test
constexpr int kPnpGridSmall = 32;   // CTAs of a run launched beside the next frame's KLT (4 hypotheses each in round 0)

__global__ void __launch_bounds__(kHypThreads)
pnp_ransac_kernel(DevScalars* sc, int32_t* sets, const float4* __restrict__ q4, const float* __restrict__ qv,
                  double* hyp, int32_t* counts, uint8_t* mask, PnpArgs a, RngTab rt, FinalizeBufs fb, int H, int epoch0,
                  const int32_t* __restrict__ gate, FusedCompact fc)
{
    __shared__ double sRt[kBatch][12];
    __shared__ int sValid[kBatch];
    __shared__ int sCount[kHypThreads / 32][kScoreTile];
    __shared__ int sLast;
    __shared__ HypF32 sHf[kScoreTile];
    __shared__ unsigned sQ[kHypThreads / 32][kQueue];
    __shared__ int sQn[kHypThreads / 32];
    pdl_prologue();
    if (gate && !*gate) return;   // enqueued ahead of a pose that did not come: every CTA leaves, nothing is touched
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int G = gridDim.x;
    const unsigned long long t_in = gtime();
    unsigned long long t_p3p = t_in, t_scored = t_in;
    int n_fused = -1;
    if (fc.on) {
        // ---- phase 0: the landmark compaction of the tracking step (what landmark_compact_kernel does with one CTA).
        // Every CTA scans ALL status bytes (a few KB out of L2), so each knows the kept count and the offset of every
        // 32-row-or-less chunk without talking to the others; the chunks are dealt round-robin to the CTAs, which write
        // their rows (order preserved); CTA 0 draws the first minimal sets meanwhile; a grid barrier (the ticket / epoch
        // pair of the rounds) ends the phase, and its last arriver publishes the counts - no CTA may have overwritten
        // sc->n_land before every CTA has read it.
        __shared__ unsigned sMask[kHypThreads];
        __shared__ int sOff[kHypThreads], sWarpTot[kHypThreads / 32], sWarpIt[kHypThreads / 32];
        const int n_in = min(sc->n_land, fc.n_upper);
        const int CH = (n_in + kHypThreads - 1) / kHypThreads;   // rows per chunk (<= 32: the host checked n_upper)
        RngPreload pre;
        if (blockIdx.x == 0 && warp == 0) pnp_preload_draws(rt, pre);
        unsigned m = 0;
        int it = 0;
        {
            // (fixed trip count, predicated: all loads of the chunk are in flight together - one L2 latency, not CH)
            const int r0 = tid * CH;
            uint8_t stv[32];
            int itv[32];
#pragma unroll
            for (int j = 0; j < 32; j++) {
                const int r = r0 + j;
                const bool in = j < CH && r < n_in;
                stv[j] = in ? fc.status[r] : (uint8_t)0;
                itv[j] = (in && fc.iters) ? fc.iters[r] : 0;
            }
#pragma unroll
            for (int j = 0; j < 32; j++) {
                if (stv[j] != 0) m |= 1u << j;
                it += itv[j];
            }
        }
        const int cnt = __popc(m);
        int inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        it = __reduce_add_sync(0xffffffffu, it);
        if (lane == 31) sWarpTot[warp] = inc;
        if (lane == 0) sWarpIt[warp] = it;
        __syncthreads();
        int base = 0, n_keep = 0, its = 0;
#pragma unroll
        for (int w = 0; w < kHypThreads / 32; w++) { if (w < warp) base += sWarpTot[w]; n_keep += sWarpTot[w]; its += sWarpIt[w]; }
        sMask[tid] = m; sOff[tid] = base + inc - cnt;
        __syncthreads();
        // this CTA's chunks: lane j of warp w writes row j of chunk blockIdx + (w + 8 i) G
        for (int c = (int)blockIdx.x + warp * G; c < kHypThreads; c += (kHypThreads / 32) * G) {
            const int r = c * CH + lane;
            if (lane < CH && r < n_in) {
                const unsigned mk = sMask[c];
                const unsigned below = (1u << lane) - 1;
                const float2 p = fc.klt_out[r];
                if ((mk >> lane) & 1u) {
                    const int o = sOff[c] + __popc(mk & below);
                    const double x = fc.X_in[3 * (size_t)r], y = fc.X_in[3 * (size_t)r + 1], z = fc.X_in[3 * (size_t)r + 2];
                    fc.P_out[o] = p;
                    fc.X_out[3 * (size_t)o] = x; fc.X_out[3 * (size_t)o + 1] = y; fc.X_out[3 * (size_t)o + 2] = z;
                    fc.q4[o] = make_float4((float)x, (float)y, (float)z, p.x);   // cv2 converts X to float32
                    fc.qv[o] = p.y;
                } else {
                    fc.lost[r - (sOff[c] + __popc(mk & below))] = p;   // (rows before r minus kept rows before r)
                }
            }
        }
        if (blockIdx.x == 0 && warp == 0) pnp_prepare_sets(sc, sets, rt, n_keep, pre);
        // grid barrier
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            sLast = atomicAdd(&sc->ticket, 1) == G - 1;
            if (sLast) {
                sc->klt_iters_cand = 0; sc->prev2_n_land = sc->prev_n_land; sc->prev_n_land = n_in; sc->n_land = n_keep;
                sc->n_tracked = n_keep; sc->n_lost = n_in - n_keep; sc->klt_iters = its;
                sc->ticket = 0;
                __threadfence();
                st_release(&sc->epoch, epoch0);
            } else {
                while (ld_acquire(&sc->epoch) != epoch0) __nanosleep(40);
            }
        }
        __syncthreads();
        n_fused = n_keep;
    }
    const int n = fc.on ? min(n_fused, a.n_upper) : min(sc->n_tracked, a.n_upper);
    const int niters0 = (n < kMinimal || a.max_iters <= 0) ? 0 : (n == kMinimal ? 1 : a.max_iters);
    int h0 = 0;
    for (int round = 0;; round++) {
        const int budget = round == 0 ? niters0 : sc->niters;
        const int remaining = budget - h0;
        // (round 0 covers the kPnpChunk0 sets drawn ahead by the compaction, however few CTAs share them)
        int B = round == 0 ? (min(remaining, kPnpChunk0) + G - 1) / G : (remaining + G - 1) / G;
        B = max(1, min(B, kBatch));
        const int h1 = round == 0 ? min(min(budget, kPnpChunk0), h0 + G * B) : min(budget, h0 + G * B);
        const int hb = h0 + (int)blockIdx.x * B;         // this CTA's hypotheses: [hb, hb + nb)
        const int nb = max(0, min(B, h1 - hb));
        // ---- P3P: one quad of lanes per warp at a time (the solver's Newton loops and early exits are data
        // dependent: quads sharing a warp would serialise each other), warp w takes hypotheses hb + w, hb + w + 8, ...
        for (int b = warp; b < nb; b += kHypThreads / 32) if (lane < 4) {
            const int h = hb + b;
            int4 s = n == kMinimal ? make_int4(0, 1, 2, 3) : reinterpret_cast<const int4*>(sets)[h];
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
            const bool winner = ok && p3p4_quad(X, p, a.K, Rt, lane, 0xFu, &found);
            if (lane == 0) sValid[b] = ok && found;
            if (winner) {
#pragma unroll
                for (int i = 0; i < 12; i++) { sRt[b][i] = Rt[i]; hyp[(size_t)h * 12 + i] = Rt[i]; }
            }
        }
        __syncthreads();
        if (round == 0) t_p3p = gtime();
        // ---- scoring: up to kScoreTile hypotheses per pass over the points ------------------------
        for (int b0 = 0; b0 < nb; b0 += kScoreTile) {
            int good[kScoreTile];
            bool valid[kScoreTile];
#pragma unroll
            for (int j = 0; j < kScoreTile; j++) { good[j] = 0; valid[j] = b0 + j < nb && sValid[b0 + j]; }
            if (a.f32_filter) {
                // float32 filter; the uncertain pairs are queued per warp and settled in float64 by all lanes together
                if (tid < kScoreTile && b0 + tid < nb && sValid[b0 + tid]) {
                    float mt = 0.f;
                    for (int i = 0; i < 9; i++) sHf[tid].R[i] = (float)sRt[b0 + tid][i];
                    for (int i = 0; i < 3; i++) { sHf[tid].t[i] = (float)sRt[b0 + tid][9 + i]; mt = fmaxf(mt, fabsf((float)sRt[b0 + tid][9 + i])); }
                    sHf[tid].mt = mt;
                }
                if (lane == 0) sQn[warp] = 0;
                __syncthreads();
                const float Kf[4] = { (float)a.K[0], (float)a.K[4], (float)a.K[2], (float)a.K[5] };
                const float thr = sqrtf(a.thr2);
                for (int i0 = 0; i0 < n; i0 += kHypThreads) {
                    const int i = i0 + tid;
                    if (i < n) {
                        const float4 q = q4[i];
                        const float v = qv[i];
#pragma unroll
                        for (int j = 0; j < kScoreTile; j++) {
                            if (!valid[j]) continue;
                            const int f = filter_err2(sHf[j], Kf, thr, q.x, q.y, q.z, q.w, v);
                            if (f >= 0) good[j] += f;
                            else {
                                const int pos = atomicAdd(&sQn[warp], 1);
                                if (pos < kQueue) sQ[warp][pos] = (unsigned)i | ((unsigned)j << 28);
                                else good[j] += reproj_err2(sRt[b0 + j], sRt[b0 + j] + 9, a.K, q.x, q.y, q.z, q.w, v) <= a.thr2;   // (queue full)
                            }
                        }
                    }
                    __syncwarp();
                    const int nq = min(sQn[warp], kQueue);
                    if (nq >= kQueue - 32 * kScoreTile || i0 + kHypThreads >= n) {   // drain: before it can overflow, and at the end
                        for (int k = lane; k < nq; k += 32) {
                            const unsigned e = sQ[warp][k];
                            const int pi = (int)(e & 0x0FFFFFFFu), pj = (int)(e >> 28);
                            const float4 q = q4[pi];
                            const int f = reproj_err2(sRt[b0 + pj], sRt[b0 + pj] + 9, a.K, q.x, q.y, q.z, q.w, qv[pi]) <= a.thr2;
#pragma unroll
                            for (int j = 0; j < kScoreTile; j++) good[j] += (pj == j) ? f : 0;
                        }
                        __syncwarp();
                        if (lane == 0) sQn[warp] = 0;
                        __syncwarp();
                    }
                }
            } else {
                for (int i = tid; i < n; i += kHypThreads) {
                    const float4 q = q4[i];
                    const float v = qv[i];
#pragma unroll
                    for (int j = 0; j < kScoreTile; j++)
                        if (valid[j]) good[j] += reproj_err2(sRt[b0 + j], sRt[b0 + j] + 9, a.K, q.x, q.y, q.z, q.w, v) <= a.thr2;
                }
            }
#pragma unroll
            for (int j = 0; j < kScoreTile; j++) {
                const int g = __reduce_add_sync(0xffffffffu, good[j]);
                if (lane == 0) sCount[warp][j] = g;
            }
            __syncthreads();
            if (tid < kScoreTile && b0 + tid < nb) {
                int sum = 0;
#pragma unroll
                for (int w = 0; w < kHypThreads / 32; w++) sum += sCount[w][tid];
                counts[hb + b0 + tid] = sValid[b0 + tid] ? sum : -1;
            }
            __syncthreads();
        }
        if (round == 0) t_scored = gtime();
        // ---- round barrier: the last CTA to arrive does the bookkeeping and opens the next round ----
        if (tid == 0) {
            __threadfence();
            sLast = atomicAdd(&sc->ticket, 1) == G - 1;
        }
        __syncthreads();
        if (sLast) {
            __threadfence();
            // (the minimal sets of the next round are drawn here, for what is left of the budget only)
            if (tid < 32) ransac_select(sc, sets, counts, a, rt, n, round == 0, h0, h1, h1 + G * kBatch, h1 >= H);
            __syncthreads();
            if (tid == 0) { sc->ticket = 0; __threadfence(); st_release(&sc->epoch, epoch0 + round + 1); }
        } else {
            if (tid == 0) while (ld_acquire(&sc->epoch) != epoch0 + round + 1) __nanosleep(40);
        }
        __syncthreads();
        if (*(volatile int32_t*)&sc->done) break;
        h0 = h1;
    }
    if (!sLast) return;
    // the run is over: this CTA finishes it
    if (tid == 0) { sc->stamps[0] = t_in; sc->stamps[1] = t_p3p; sc->stamps[2] = t_scored; sc->stamps[3] = gtime(); }
    pnp_mask_prune(sc, hyp, q4, qv, mask, a, fb, n);
    if (tid == 0) { sc->finalized = 1; sc->stamps[6] = gtime(); }
}

}  // namespace

// raw outputs of cv::RNG(2^64-1) in draw order; 16 draws per iteration + slack covers the re-draws
// of getSubset with overwhelming margin (4 + O(1/n) are used); if it ever runs out the device sets
// rng_exhausted and the host reports VO_ERR_CAPACITY
int ensure_rng_table(Ctx* c, int H)
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
    VO_NVTX("vo:pnp_ransac+refit");
    Pipeline* p = c->pipe;
    VO_TRY(join_pose_stream(c));   // (a refit still running there reads the inlier pack this run overwrites)
    PnpArgs a;
    memcpy(a.K, L.K, sizeof(a.K));
    a.confidence = L.prm.confidence;
    a.thr2 = (float)(L.prm.reproj_thresh * L.prm.reproj_thresh);
    a.max_iters = L.prm.max_iters < 0 ? 0 : L.prm.max_iters;
    a.refine_iters = L.prm.refine_iters < 0 ? 0 : L.prm.refine_iters;
    a.n_upper = L.n_upper;
    a.explicit_sets = L.d_sets_in != nullptr;
    a.compact = L.compact;
    {   // A/B and test switch, read per launch: VO_PNP_F32_FILTER=1 decides the pairs far from the threshold in float32.
        // Off by default: with 8 warps per SM the scoring is latency-, not throughput-bound, and the filter measured
        // SLOWER on the B200 (stress shape 0.507 ms against 0.432 ms, KITTI-shaped frames 4.6 us against 3.1 us of scoring)
        const char* e = getenv("VO_PNP_F32_FILTER");
        a.f32_filter = e && e[0] == '1';
    }
    const int H = a.max_iters;
    const size_t npts = (size_t)(L.n_upper > 0 ? L.n_upper : 1);
    VO_TRY(ensure(c, p->hyp, (size_t)(H > 0 ? H : 1) * 12));
    VO_TRY(ensure(c, p->counts, (size_t)(H > 0 ? H : 1)));
    VO_TRY(ensure(c, p->mask, npts));
    VO_TRY(ensure(c, p->q4i, npts * 4)); VO_TRY(ensure(c, p->qvi, npts));
    int32_t* sets = const_cast<int32_t*>(L.d_sets_in);
    RngTab rt{ nullptr, 0 };
    if (!sets) {
        // rows [0, kPnpChunk0) were drawn by the kernel that fixed the point count (compaction / packing)
        VO_TRY(ensure(c, p->sets, (size_t)(H > kPnpChunk0 ? H : kPnpChunk0) * kMinimal, true));
        sets = p->sets.p;
        VO_TRY(ensure_rng_table(c, H));
        rt.tab = p->rng_tab.p; rt.len = p->rng_tab_len;
    }
    FusedCompact fc{};
    const float* q4_arg = L.q4; const float* qv_arg = L.qv;
    if (L.fuse_compact) {
        // the tracking step's compaction, done by the RANSAC kernel: buffers and bookkeeping as launch_landmark_compact
        const size_t nn = npts;
        const int cnx = (p->cur + 1) % kLandBufs;
        VO_TRY(ensure(c, p->P[cnx], nn * 2)); VO_TRY(ensure(c, p->X[cnx], nn * 3));
        VO_TRY(ensure(c, p->q4, nn * 4)); VO_TRY(ensure(c, p->qv, nn)); VO_TRY(ensure(c, p->lost, nn * 2));
        fc.on = 1; fc.n_upper = L.n_upper;
        fc.klt_out = (const float2*)p->klt_out.p; fc.status = p->klt_status.p; fc.iters = p->klt_iters.p;
        fc.X_in = p->X[p->cur].p; fc.P_out = (float2*)p->P[cnx].p; fc.X_out = p->X[cnx].p; fc.lost = (float2*)p->lost.p;
        fc.q4 = (float4*)p->q4.p; fc.qv = p->qv.p;
        q4_arg = p->q4.p; qv_arg = p->qv.p;
        p->prev2 = p->prev == cnx ? -1 : p->prev; p->prev2_n_upper = p->prev_n_upper;
        p->prev = p->cur; p->prev_n_upper = L.n_upper;
        p->cur = cnx;
        p->b1_buf = cnx; p->b1_upper = L.n_upper;
    }
    FinalizeBufs fb{};
    fb.q4i = (float4*)p->q4i.p; fb.qvi = p->qvi.p;
    if (L.compact) {
        fb.P_in = p->P[p->cur].p; fb.X_in = p->X[p->cur].p;
        const int nxt = (p->cur + 1) % kLandBufs;
        VO_TRY(ensure(c, p->P[nxt], npts * 2));
        VO_TRY(ensure(c, p->X[nxt], npts * 3));
        fb.P_out = p->P[nxt].p; fb.X_out = p->X[nxt].p;
        VO_TRY(ensure(c, p->Lsrc[nxt], npts + (size_t)p->n_cand_upper + 8192));   // (room for the rows the frame tail appends)
        fb.src_out = p->Lsrc[nxt].p;
    }
    p->src_valid = false;
    // one persistent launch: every CTA must be resident (they meet at a flag between rounds)
    // one CTA per hypothesis of round 0 is the shortest way to the pose; when the next frame's KLT runs beside this
    // kernel (L.small_grid) and the run has an adaptive stop (it almost always ends in round 0), a quarter of the CTAs
    // does the same round 2 us slower and leaves the other SMs alone.  VO_PNP_GRID: A/B switch (same results)
    static const int grid_env = getenv("VO_PNP_GRID") ? atoi(getenv("VO_PNP_GRID")) : 0;
    int grid = (L.small_grid && a.confidence < 1.) ? kPnpGridSmall : kPnpChunk0;
    if (grid_env > 0) grid = grid_env;
    if (grid > kPnpChunk0) grid = kPnpChunk0;
    if (grid > H) grid = H > 0 ? H : 1;
    if (grid > c->sm_count) grid = c->sm_count;
    const int epoch0 = (int)((unsigned)(++p->pnp_run) << 12);
    VO_CUDA(c, launch_pdl(pnp_ransac_kernel, dim3(grid), dim3(kHypThreads), c->stream, p->d_sc, sets, (const float4*)q4_arg, qv_arg,
                          p->hyp.p, p->counts.p, p->mask.p, a, rt, fb, H, epoch0, L.gate, fc));
    c->launches++;
    // the refit: on the caller's stream, or on the pose stream when the landmark side goes on without it
    cudaStream_t rs = c->stream;
    if (L.refit_on_pose_stream) {
        VO_CUDA(c, cudaEventRecord(c->ev_ransac, c->stream));
        VO_CUDA(c, cudaStreamWaitEvent(c->stream4, c->ev_ransac, 0));
        rs = c->stream4;
    }
    VO_CUDA(c, launch_pdl(pnp_refit_kernel, dim3(1), dim3(kFinThreads), rs, p->d_sc, (const float4*)p->q4i.p,
                          (const float*)p->qvi.p, a));
    c->launches++;
    if (L.compact) {
        const int nxt = (p->cur + 1) % kLandBufs;
        if (nxt == p->prev) p->prev = -1;   // pruning without a preceding tracking step reuses that buffer
        if (nxt == p->prev2) p->prev2 = -1;
        p->cur = nxt;
    }
    VO_CUDA(c, cudaGetLastError());
    return VO_OK;
}

}  // namespace vo
