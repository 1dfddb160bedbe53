/*
 * ORACLE (test infrastructure only - never linked into or called by the product path).
 *
 * Scalar CPU restatement of the PnP-RANSAC the reference reaches through
 *   cv2.solvePnPRansac(X, P, K, zeros(4), flags=..., confidence=..., reprojectionError=2)
 *   /root/reference/src/vo_pipeline.py:258-264   (PoseEstimator.estimatePose)
 * followed by the rvec->R conversion of vo_pipeline.py:266 and the 4x4 assembly of :333.
 * The arithmetic lives in OpenCV (third-party, not vendored by the reference; pinned
 * opencv-python==4.8.1.78 in requirements.txt:2, 4.13.0.92 installed here).  This file restates
 * the published algorithm (behavioural spec in SURVEY.md Appendix A.2):
 *   - cv::RNG(2^64-1) multiply-with-carry generator, subset draw with duplicate re-draw
 *   - minimal solver: P3P on the first 3 points, 4th point selects the root (the variant
 *     BASELINE.json's north_star asks for and the only one an independent implementation can
 *     reproduce); here Grunert's quartic with bracketing root isolation,
 *     Gauss-Newton polishing of the three distances and exact 3-point alignment
 *   - scoring: float64 projection rounded to float32, float32 dx*dx+dy*dy <= (float)(thr*thr)
 *   - first strictly-better model wins, adaptive iteration count (RANSACUpdateNumIters)
 *   - final refit = converged reprojection-error minimum on the inlier set (what
 *     cv2.solvePnP(SOLVEPNP_ITERATIVE) returns), Levenberg-Marquardt on SE(3)
 * Pinned against cv2 4.13.0 outputs: tests/golden/pnp_*.npz (tests/golden/make_golden.py):
 * masks of cv2.solvePnPRansac(flags=SOLVEPNP_P3P), poses of cv2.solvePnP(ITERATIVE) on them.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---- cv::RNG ------------------------------------------------------------------------------- */
typedef struct { uint64_t s; } cvrng;

static unsigned rng_next(cvrng* r)
{
    r->s = (uint64_t)(unsigned)r->s * 4164903690U + (unsigned)(r->s >> 32);
    return (unsigned)r->s;
}

static int rng_uniform(cvrng* r, int a, int b) { return a == b ? a : (int)(rng_next(r) % (unsigned)(b - a)) + a; }

/* The H x m table of minimal-set indices cv2's RANSAC loop would draw for n points
 * (one row per iteration; duplicates inside a row are re-drawn, as getSubset does). */
void vo_oracle_ransac_sets(int n, int m, int H, int32_t* out)
{
    cvrng r = { 0xFFFFFFFFFFFFFFFFULL };
    for (int h = 0; h < H; h++) {
        int32_t* idx = out + (size_t)h * m;
        for (int i = 0; i < m; i++) {
            for (;;) {
                int v = rng_uniform(&r, 0, n), dup = 0;
                for (int j = 0; j < i; j++) dup |= idx[j] == v;
                if (!dup) { idx[i] = v; break; }
            }
        }
    }
}

/* RANSACUpdateNumIters */
static int update_num_iters(double p, double ep, int model_points, int max_iters)
{
    p = p < 0. ? 0. : (p > 1. ? 1. : p);
    ep = ep < 0. ? 0. : (ep > 1. ? 1. : ep);
    double num = 1. - p < DBL_MIN ? DBL_MIN : 1. - p;
    double denom = 1. - pow(1. - ep, model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)nearbyint(num / denom);
}

int vo_oracle_update_num_iters(double p, double ep, int m, int max_iters) { return update_num_iters(p, ep, m, max_iters); }

/* ---- polynomial real roots by recursive bracketing -------------------------------------------- */
static double peval(const double* c, int deg, double x)
{
    double v = c[deg];
    for (int i = deg - 1; i >= 0; i--) v = v * x + c[i];
    return v;
}

/* root of c (degree deg) inside [lo,hi] with f(lo)*f(hi) <= 0: Newton safeguarded by bisection */
static double bracket_root(const double* c, int deg, double lo, double hi, double flo, double fhi)
{
    if (flo == 0.) return lo;
    if (fhi == 0.) return hi;
    double d[5];
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

/* all real roots of c[0..deg] (deg <= 4) in (lo, hi], ascending; returns the count */
static int real_roots(const double* cin, int deg, double lo, double hi, double* roots)
{
    double c[5], big = 0.;
    for (int i = 0; i <= deg; i++) { c[i] = cin[i]; if (fabs(c[i]) > big) big = fabs(c[i]); }
    while (deg > 0 && fabs(c[deg]) <= 1e-14 * big) deg--;
    if (deg == 0) return 0;
    /* derivative chain: p[k] is the (deg-k)... keep every derivative down to degree 1 */
    double der[4][5];
    int degs[4];
    memcpy(der[0], c, sizeof(c));
    degs[0] = deg;
    int nd = 1;
    while (degs[nd - 1] > 1) {
        for (int i = 1; i <= degs[nd - 1]; i++) der[nd][i - 1] = der[nd - 1][i] * i;
        degs[nd] = degs[nd - 1] - 1;
        nd++;
    }
    /* roots of the linear one, then climb */
    double r[5], prev[5];
    int nprev = 0;
    {
        const double* l = der[nd - 1];
        const double x = -l[0] / l[1];
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
                if (fb == 0.) { r[nr++] = b; }
                else if (fa != 0. && (fa < 0.) != (fb < 0.)) r[nr++] = bracket_root(p, dg, a, b, fa, fb);
            }
            a = b; fa = fb;
        }
        nprev = nr;
        memcpy(prev, r, sizeof(double) * nr);
    }
    memcpy(roots, prev, sizeof(double) * nprev);
    return nprev;
}

/* ---- small linear algebra ---------------------------------------------------------------- */
static void cross3(const double* a, const double* b, double* c)
{
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}

static double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

/* inverse of the 3x3 matrix whose COLUMNS are a,b,c (row-major result); returns 0 if singular */
static int inv_cols(const double* a, const double* b, const double* c, double* inv)
{
    double bc[3], ca[3], ab[3];
    cross3(b, c, bc); cross3(c, a, ca); cross3(a, b, ab);
    const double det = dot3(a, bc);
    const double scale = sqrt(dot3(a, a) * dot3(b, b) * dot3(c, c));
    if (!(fabs(det) > 1e-12 * scale) || scale == 0.) return 0;
    const double id = 1. / det;
    for (int i = 0; i < 3; i++) { inv[i] = bc[i] * id; inv[3 + i] = ca[i] * id; inv[6 + i] = ab[i] * id; }
    return 1;
}

/* ---- P3P: all solutions for 3 world points X (3x3 row-major, one point per row) and unit
 * bearings j (3x3, one per row).  Rs: up to 4 rotations (row-major), ts: translations. ------- */
static int p3p_solve(const double* X, const double* j, double Rs[4][9], double ts[4][3])
{
    const double* X1 = X; const double* X2 = X + 3; const double* X3 = X + 6;
    double d12[3], d13[3], d23[3];
    for (int i = 0; i < 3; i++) { d12[i] = X1[i] - X2[i]; d13[i] = X1[i] - X3[i]; d23[i] = X2[i] - X3[i]; }
    const double a2 = dot3(d23, d23), b2 = dot3(d13, d13), c2 = dot3(d12, d12);
    if (!(a2 > 0. && b2 > 0. && c2 > 0.)) return 0;
    const double ca = dot3(j + 3, j + 6), cb = dot3(j, j + 6), cg = dot3(j, j + 3);
    /* world triangle frame */
    double nX[3], XmInv[9];
    cross3(d12, d13, nX);
    if (!inv_cols(d12, d13, nX, XmInv)) return 0;

    /* u = N(v)/D(v),  N = (q-1)v^2 - 2 q cb v + (1+q),  D = 2(cg - ca v),  q = (a2-c2)/b2.
     * Substituted into  u^2 - 2 cg u + 1 - (c2/b2)(1 + v^2 - 2 cb v) = 0  times D^2:
     *   N^2 - 2 cg N D + (1 - (c2/b2) W) D^2 = 0,  W = 1 - 2 cb v + v^2. */
    const double q = (a2 - c2) / b2, k = c2 / b2;
    const double N[3] = { 1. + q, -2. * q * cb, q - 1. };
    const double D[2] = { 2. * cg, -2. * ca };
    const double W[3] = { 1., -2. * cb, 1. };
    double NN[5] = { 0 }, ND[4] = { 0 }, DD[3] = { 0 }, G[3], GDD[5] = { 0 }, Q[5];
    for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) NN[a + b] += N[a] * N[b];
    for (int a = 0; a < 3; a++) for (int b = 0; b < 2; b++) ND[a + b] += N[a] * D[b];
    for (int a = 0; a < 2; a++) for (int b = 0; b < 2; b++) DD[a + b] += D[a] * D[b];
    for (int a = 0; a < 3; a++) G[a] = (a == 0 ? 1. : 0.) - k * W[a];
    for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) GDD[a + b] += G[a] * DD[b];
    for (int a = 0; a < 5; a++) Q[a] = NN[a] + GDD[a] - (a < 4 ? 2. * cg * ND[a] : 0.);

    double big = 0.;
    for (int a = 0; a < 5; a++) if (fabs(Q[a]) > big) big = fabs(Q[a]);
    if (!(big > 0.) || !isfinite(big)) return 0;
    double vr[4];
    const int nv = real_roots(Q, 4, 0., 1e6, vr);

    int ns = 0;
    for (int r = 0; r < nv && ns < 4; r++) {
        const double v = vr[r];
        const double Wv = 1. + v * (v - 2. * cb);
        if (!(Wv > 0.)) continue;
        const double Dv = D[0] + D[1] * v;
        double ucand[2];
        int nu = 0;
        if (fabs(Dv) > 1e-9) {
            ucand[nu++] = (N[0] + v * (N[1] + v * N[2])) / Dv;
        } else {
            /* D(v) ~ 0: take u from the quadratic directly, keep the one(s) consistent with eq. (1) */
            const double cc = 1. - k * Wv, disc = cg * cg - cc;
            if (disc >= 0.) {
                const double sq = sqrt(disc);
                const double u0 = cg + sq, u1 = cg - sq;
                const double e0 = fabs(u0 * u0 + v * v - 2. * u0 * v * ca - (a2 / b2) * Wv);
                const double e1 = fabs(u1 * u1 + v * v - 2. * u1 * v * ca - (a2 / b2) * Wv);
                ucand[nu++] = e0 <= e1 ? u0 : u1;
            }
        }
        for (int ui = 0; ui < nu && ns < 4; ui++) {
            const double u = ucand[ui];
            if (!(u > 0.)) continue;
            double s1 = sqrt(b2 / Wv), s2 = u * s1, s3 = v * s1;
            /* Gauss-Newton polish of the three distances on the law-of-cosines system */
            for (int it = 0; it < 5; it++) {
                const double r12 = s1 * s1 + s2 * s2 - 2. * s1 * s2 * cg - c2;
                const double r13 = s1 * s1 + s3 * s3 - 2. * s1 * s3 * cb - b2;
                const double r23 = s2 * s2 + s3 * s3 - 2. * s2 * s3 * ca - a2;
                /* J rows: d r12, d r13, d r23 w.r.t. (s1,s2,s3) */
                const double J00 = 2. * (s1 - s2 * cg), J01 = 2. * (s2 - s1 * cg);
                const double J10 = 2. * (s1 - s3 * cb), J12 = 2. * (s3 - s1 * cb);
                const double J21 = 2. * (s2 - s3 * ca), J22 = 2. * (s3 - s2 * ca);
                const double det = -J00 * J12 * J21 - J01 * J10 * J22;
                if (!(fabs(det) > 1e-300)) break;
                /* inverse via adjugate of [[J00,J01,0],[J10,0,J12],[0,J21,J22]] */
                const double i00 = -J12 * J21, i01 = -J01 * J22, i02 = J01 * J12;
                const double i10 = -J10 * J22, i11 = J00 * J22, i12 = -J00 * J12;
                const double i20 = J10 * J21, i21 = -J00 * J21, i22 = -J01 * J10;
                const double ds1 = (i00 * r12 + i01 * r13 + i02 * r23) / det;
                const double ds2 = (i10 * r12 + i11 * r13 + i12 * r23) / det;
                const double ds3 = (i20 * r12 + i21 * r13 + i22 * r23) / det;
                s1 -= ds1; s2 -= ds2; s3 -= ds3;
                if (fabs(ds1) + fabs(ds2) + fabs(ds3) <= 1e-15 * (s1 + s2 + s3)) break;
            }
            if (!(s1 > 0. && s2 > 0. && s3 > 0.)) continue;
            double Y1[3], Y2[3], Y3[3], e12[3], e13[3], nY[3];
            for (int i = 0; i < 3; i++) { Y1[i] = s1 * j[i]; Y2[i] = s2 * j[3 + i]; Y3[i] = s3 * j[6 + i]; }
            for (int i = 0; i < 3; i++) { e12[i] = Y1[i] - Y2[i]; e13[i] = Y1[i] - Y3[i]; }
            cross3(e12, e13, nY);
            /* R = [e12 e13 nY] * inv([d12 d13 nX]) */
            double* R = Rs[ns];
            for (int i = 0; i < 3; i++)
                for (int c = 0; c < 3; c++)
                    R[3 * i + c] = e12[i] * XmInv[c] + e13[i] * XmInv[3 + c] + nY[i] * XmInv[6 + c];
            for (int i = 0; i < 3; i++) ts[ns][i] = Y1[i] - (R[3 * i] * X1[0] + R[3 * i + 1] * X1[1] + R[3 * i + 2] * X1[2]);
            int fin = 1;
            for (int i = 0; i < 9; i++) fin &= isfinite(R[i]) != 0;
            if (fin) ns++;
        }
    }
    return ns;
}

/* float64 projection as cv::projectPoints evaluates it (no distortion) */
static void project(const double* R, const double* t, const double* K, const double* X, double* u, double* v)
{
    double x = R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + t[0];
    double y = R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + t[1];
    double z = R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + t[2];
    z = z != 0. ? 1. / z : 1.;
    x *= z; y *= z;
    *u = x * K[0] + K[2];
    *v = y * K[4] + K[5];
}

/* All P3P solutions of 3 correspondences: X 3x3 (float64), p 3x2 pixels.  For comparison with
 * cv2.solveP3P. */
int vo_oracle_p3p_all(const double* X, const double* p, const double* K, double* Rs, double* ts)
{
    double j[9], R4[4][9], t4[4][3];
    for (int i = 0; i < 3; i++) {
        const double x = (p[2 * i] - K[2]) / K[0], y = (p[2 * i + 1] - K[5]) / K[4];
        const double n = sqrt(x * x + y * y + 1.);
        j[3 * i] = x / n; j[3 * i + 1] = y / n; j[3 * i + 2] = 1. / n;
    }
    const int ns = p3p_solve(X, j, R4, t4);
    for (int s = 0; s < ns; s++) { memcpy(Rs + 9 * s, R4[s], sizeof(double) * 9); memcpy(ts + 3 * s, t4[s], sizeof(double) * 3); }
    return ns;
}

/* cv2.solvePnP(flags=SOLVEPNP_P3P) on exactly 4 correspondences: roots from the first three,
 * the fourth picks the root with the smallest pixel reprojection error.  Returns 0 if no root. */
int vo_oracle_p3p4(const double* X /*4x3*/, const double* p /*4x2*/, const double* K, double* R, double* t)
{
    double Rs[36], ts[12];
    const int ns = vo_oracle_p3p_all(X, p, K, Rs, ts);
    int best = -1;
    double beste = 0.;
    for (int s = 0; s < ns; s++) {
        double u, v;
        project(Rs + 9 * s, ts + 3 * s, K, X + 9, &u, &v);
        const double e = (u - p[6]) * (u - p[6]) + (v - p[7]) * (v - p[7]);
        if (best < 0 || e < beste) { best = s; beste = e; }
    }
    if (best < 0) return 0;
    memcpy(R, Rs + 9 * best, sizeof(double) * 9);
    memcpy(t, ts + 3 * best, sizeof(double) * 3);
    return 1;
}

/* Inlier mask of one pose over all points (X already float32-rounded, as doubles). */
int vo_oracle_score(const double* X, const float* P, int n, const double* K, const double* R, const double* t,
                    double thr, uint8_t* mask)
{
    const float t2 = (float)(thr * thr);
    int good = 0;
    for (int i = 0; i < n; i++) {
        double u, v;
        project(R, t, K, X + 3 * i, &u, &v);
        const float dx = P[2 * i] - (float)u, dy = P[2 * i + 1] - (float)v;
        const float e = dx * dx + dy * dy;
        const int f = e <= t2;
        if (mask) mask[i] = (uint8_t)f;
        good += f;
    }
    return good;
}

/* ---- final refit: Levenberg-Marquardt on the reprojection error over the masked points ------- */
static void so3_exp(const double* w, double* R)
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

static double reproj_cost(const double* X, const float* P, const uint8_t* mask, int n, const double* K,
                          const double* R, const double* t)
{
    double c = 0.;
    for (int i = 0; i < n; i++) {
        if (mask && !mask[i]) continue;
        double u, v;
        project(R, t, K, X + 3 * i, &u, &v);
        const double dx = u - (double)P[2 * i], dy = v - (double)P[2 * i + 1];
        c += dx * dx + dy * dy;
    }
    return c;
}

static int chol6_solve(double* A, double* b)
{
    for (int i = 0; i < 6; i++) {
        for (int j = 0; j <= i; j++) {
            double s = A[6 * i + j];
            for (int k = 0; k < j; k++) s -= A[6 * i + k] * A[6 * j + k];
            if (i == j) { if (!(s > 0.)) return 0; A[6 * i + i] = sqrt(s); }
            else A[6 * i + j] = s / A[6 * j + j];
        }
    }
    for (int i = 0; i < 6; i++) { double s = b[i]; for (int k = 0; k < i; k++) s -= A[6 * i + k] * b[k]; b[i] = s / A[6 * i + i]; }
    for (int i = 5; i >= 0; i--) { double s = b[i]; for (int k = i + 1; k < 6; k++) s -= A[6 * k + i] * b[k]; b[i] = s / A[6 * i + i]; }
    return 1;
}

int vo_oracle_refine_lm(const double* X, const float* P, const uint8_t* mask, int n, const double* K,
                        double* R, double* t, int max_iters)
{
    double lambda = 1e-3, cost = reproj_cost(X, P, mask, n, K, R, t);
    int it;
    for (it = 0; it < max_iters; it++) {
        double H[36] = { 0 }, g[6] = { 0 };
        for (int i = 0; i < n; i++) {
            if (mask && !mask[i]) continue;
            const double* Xi = X + 3 * i;
            const double rx = R[0] * Xi[0] + R[1] * Xi[1] + R[2] * Xi[2];
            const double ry = R[3] * Xi[0] + R[4] * Xi[1] + R[5] * Xi[2];
            const double rz = R[6] * Xi[0] + R[7] * Xi[1] + R[8] * Xi[2];
            const double x = rx + t[0], y = ry + t[1], z = rz + t[2];
            const double iz = 1. / z;
            const double ru = K[0] * x * iz + K[2] - (double)P[2 * i];
            const double rv = K[4] * y * iz + K[5] - (double)P[2 * i + 1];
            /* d(u,v)/d(xc) */
            const double ux = K[0] * iz, uz = -K[0] * x * iz * iz;
            const double vy = K[4] * iz, vz = -K[4] * y * iz * iz;
            /* xc' = xc + w x (R X) + dt  =>  d xc / d w = -[R X]x */
            double Ju[6], Jv[6];
            /* -[r]x = [[0, rz, -ry], [-rz, 0, rx], [ry, -rx, 0]] (columns = d xc / d w_k) */
            Ju[0] = ux * 0. + uz * ry;     Jv[0] = vy * (-rz) + vz * ry;
            Ju[1] = ux * rz + uz * (-rx);  Jv[1] = vy * 0. + vz * (-rx);
            Ju[2] = ux * (-ry) + uz * 0.;  Jv[2] = vy * rx + vz * 0.;
            Ju[3] = ux; Ju[4] = 0.; Ju[5] = uz;
            Jv[3] = 0.; Jv[4] = vy; Jv[5] = vz;
            for (int a = 0; a < 6; a++) {
                g[a] += Ju[a] * ru + Jv[a] * rv;
                for (int b = 0; b <= a; b++) H[6 * a + b] += Ju[a] * Ju[b] + Jv[a] * Jv[b];
            }
        }
        int improved = 0;
        for (int tries = 0; tries < 12; tries++) {
            double A[36], d[6];
            for (int a = 0; a < 6; a++) {
                for (int b = 0; b <= a; b++) A[6 * a + b] = H[6 * a + b];
                A[6 * a + a] *= 1. + lambda;
                d[a] = -g[a];
            }
            if (!chol6_solve(A, d)) { lambda *= 10.; continue; }
            double dR[9], Rn[9], tn[3];
            so3_exp(d, dR);
            for (int r = 0; r < 3; r++)
                for (int c = 0; c < 3; c++)
                    Rn[3 * r + c] = dR[3 * r] * R[c] + dR[3 * r + 1] * R[3 + c] + dR[3 * r + 2] * R[6 + c];
            for (int r = 0; r < 3; r++) tn[r] = t[r] + d[3 + r];
            const double cn = reproj_cost(X, P, mask, n, K, Rn, tn);
            if (cn <= cost) {
                double step = 0.;
                for (int a = 0; a < 6; a++) step += d[a] * d[a];
                memcpy(R, Rn, sizeof(Rn)); memcpy(t, tn, sizeof(tn));
                const double dec = cost - cn;
                cost = cn; lambda *= 0.1; if (lambda < 1e-12) lambda = 1e-12;
                improved = 1;
                if (step < 1e-24 || dec <= 1e-15 * cost) return it + 1;
                break;
            }
            lambda *= 10.;
        }
        if (!improved) break;
    }
    return it;
}

/* R (row-major) -> rotation vector, as cv2.Rodrigues would */
void vo_oracle_rotation_to_rvec(const double* R, double* rvec)
{
    double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
    const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
    double c = (R[0] + R[4] + R[8] - 1.) * 0.5;
    c = c > 1. ? 1. : (c < -1. ? -1. : c);
    const double theta = acos(c);
    if (s < 1e-5) {
        if (c > 0.) { rvec[0] = rvec[1] = rvec[2] = 0.; }
        else {
            double t;
            t = (R[0] + 1.) * 0.5; rx = sqrt(t > 0. ? t : 0.);
            t = (R[4] + 1.) * 0.5; ry = sqrt(t > 0. ? t : 0.) * (R[1] < 0. ? -1. : 1.);
            t = (R[8] + 1.) * 0.5; rz = sqrt(t > 0. ? t : 0.) * (R[2] < 0. ? -1. : 1.);
            if (fabs(rx) < fabs(ry) && fabs(rx) < fabs(rz) && (R[5] > 0.) != (ry * rz > 0.)) rz = -rz;
            const double nn = theta / sqrt(rx * rx + ry * ry + rz * rz);
            rvec[0] = rx * nn; rvec[1] = ry * nn; rvec[2] = rz * nn;
        }
    } else {
        const double vth = 0.5 / s * theta;
        rvec[0] = rx * vth; rvec[1] = ry * vth; rvec[2] = rz * vth;
    }
}

/* ---- the RANSAC loop -------------------------------------------------------------------------
 * X: n x 3 float64 (rounded to float32 internally, as cv2 does), P: n x 2 float32.
 * sets: H x 4 minimal-set table or NULL (= cv::RNG emulation).  Outputs: R,t = refit pose,
 * Rh,th = winning unrefined hypothesis, mask = its inlier mask, info = {iterations run,
 * winning iteration, inlier count, LM iterations}.  Returns 1 if a model was found. */
int vo_oracle_solve_pnp_ransac(const double* X64, const float* P, int n, const double* K, int max_iters,
                               double confidence, double thr, int refine_iters, const int32_t* sets_in,
                               double* R, double* t, double* Rh, double* th, uint8_t* mask, int32_t* info)
{
    const int m = 4;
    info[0] = info[2] = info[3] = 0; info[1] = -1;
    if (n < m) return 0;
    double* X = (double*)malloc(sizeof(double) * 3 * (size_t)n);
    for (int i = 0; i < 3 * n; i++) X[i] = (double)(float)X64[i];
    int32_t* sets = NULL;
    if (!sets_in) { sets = (int32_t*)malloc(sizeof(int32_t) * m * (size_t)(max_iters > 0 ? max_iters : 1)); vo_oracle_ransac_sets(n, m, max_iters, sets); sets_in = sets; }
    uint8_t* cur = (uint8_t*)malloc((size_t)n);
    int best = 0, niters = max_iters, iter;
    for (iter = 0; iter < niters; iter++) {
        const int32_t* idx = sets_in + (size_t)iter * m;
        double Xs[12], ps[8], Rc[9], tc[3];
        for (int k = 0; k < 4; k++) {
            for (int c = 0; c < 3; c++) Xs[3 * k + c] = X[3 * idx[k] + c];
            ps[2 * k] = (double)P[2 * idx[k]]; ps[2 * k + 1] = (double)P[2 * idx[k] + 1];
        }
        if (!vo_oracle_p3p4(Xs, ps, K, Rc, tc)) continue;
        const int good = vo_oracle_score(X, P, n, K, Rc, tc, thr, cur);
        if (good > (best > m - 1 ? best : m - 1)) {
            best = good;
            memcpy(mask, cur, (size_t)n);
            memcpy(Rh, Rc, sizeof(Rc)); memcpy(th, tc, sizeof(tc));
            info[1] = iter;
            niters = update_num_iters(confidence, (double)(n - good) / n, m, niters);
        }
    }
    info[0] = iter; info[2] = best;
    int ok = best > 0;
    if (ok) {
        memcpy(R, Rh, sizeof(double) * 9); memcpy(t, th, sizeof(double) * 3);
        info[3] = vo_oracle_refine_lm(X, P, mask, n, K, R, t, refine_iters);
    }
    free(cur); free(sets); free(X);
    return ok;
}
