/* ORACLE (test infrastructure only - never linked into the product).
 *
 * CPU restatement of the bootstrap's two OpenCV calls (reference: /root/reference/src/vo_pipeline.py:139-162,
 * VOInitializer.getPoseEstimate):
 *     E, mask = cv2.findEssentialMat(kps1, kps2, K, cv2.RANSAC, 0.999, 1.0, None)
 *     _, R, t, _ = cv2.recoverPose(E, kps1, kps2, K, mask=mask)          # ANDs the cheirality test into `mask` in place
 * The arithmetic lives in OpenCV (third party, pinned opencv-python==4.8.1.78 in the reference's requirements.txt,
 * installed here 4.13.0; modules/calib3d/src/five-point.cpp, ptsetreg.cpp): restated from its published algorithm -
 *   - points normalised with fx, fy, cx, cy; threshold / ((fx + fy) / 2);
 *   - RANSAC loop of cv::RANSACPointSetRegistrator: cv::RNG(2^64-1), 5-point subsets with duplicate re-draws, every model
 *     of a subset scored in turn, "first strictly better wins", adaptive iteration budget (model points = 5), no refit;
 *   - Nister's five-point solver: 4-dimensional null space of the 5x9 epipolar system, the ten cubic constraints
 *     (det E = 0, 2 E E^T E - tr(E E^T) E = 0) eliminated to a 3x3 polynomial matrix B(z), det B(z) = 0 (degree 10);
 *   - Sampson error, float32, <= (float)(thr^2);
 *   - recoverPose: decomposeEssentialMat, four (R, t) candidates, DLT triangulation, cheirality with distance 50.
 * Parity pin: tests/golden/bootstrap.npz (cv2 outputs made by tests/golden/make_golden.py bootstrap) - inlier masks,
 * E up to sign, R, t.  OpenCV finds the real roots with cv::solvePoly (Durand-Kerner, order arbitrary, |imag| <= 1e-10);
 * here they come from a Sturm sequence in ascending order: results differ only when two models of ONE subset tie. */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---- cv::RNG and the subset draw (as oracle/pnp.c) ------------------------------------------------------------ */
static uint32_t rng_next(uint64_t* s)
{
    *s = (uint64_t)(uint32_t)(*s) * 4164903690ULL + (uint32_t)(*s >> 32);
    return (uint32_t)*s;
}

static int update_num_iters(double p, double ep, int model_points, int max_iters)
{
    p = p < 0 ? 0 : (p > 1 ? 1 : p);
    ep = ep < 0 ? 0 : (ep > 1 ? 1 : ep);
    double num = 1. - p > DBL_MIN ? 1. - p : DBL_MIN;
    double denom = 1. - pow(1. - ep, model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)rint(num / denom);
}

/* ---- polynomials in (x, y, z) of total degree <= 3, dense over exponents ------------------------------------ */
typedef struct { double c[4][4][4]; } Poly;   /* c[i][j][k] * x^i y^j z^k */

static void p_zero(Poly* p) { memset(p, 0, sizeof(*p)); }
static void p_addmul(Poly* out, const Poly* a, const Poly* b, double s)   /* out += s * a * b (degrees must fit) */
{
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4 - i; j++) for (int k = 0; k < 4 - i - j; k++) {
        const double av = a->c[i][j][k];
        if (av == 0.) continue;
        for (int l = 0; l < 4 - i; l++) for (int m = 0; m < 4 - j - l + 0 && i + l + j + m < 4; m++)
            for (int n = 0; i + l + j + m + k + n < 4; n++) {
                const double bv = b->c[l][m][n];
                if (bv != 0.) out->c[i + l][j + m][k + n] += s * av * bv;
            }
    }
}
static void p_axpy(Poly* out, const Poly* a, double s)
{
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) for (int k = 0; k < 4; k++) out->c[i][j][k] += s * a->c[i][j][k];
}

/* Nister's monomial order: x^3 y^3 x^2y xy^2 x^2z x^2 y^2z y^2 xyz xy | xz^2 xz x yz^2 yz y z^3 z^2 z 1 */
static const int kMono[20][3] = { {3,0,0},{0,3,0},{2,1,0},{1,2,0},{2,0,1},{2,0,0},{0,2,1},{0,2,0},{1,1,1},{1,1,0},
                                  {1,0,2},{1,0,1},{1,0,0},{0,1,2},{0,1,1},{0,1,0},{0,0,3},{0,0,2},{0,0,1},{0,0,0} };

static void p_row(const Poly* p, double* row)
{
    for (int m = 0; m < 20; m++) row[m] = p->c[kMono[m][0]][kMono[m][1]][kMono[m][2]];
}

/* ---- null space of the 5 x 9 epipolar system: Gauss-Jordan with complete pivoting ---------------------------- */
static int null_space_5x9(double Q[5][9], double EE[4][9])
{
    int colperm[9];
    for (int j = 0; j < 9; j++) colperm[j] = j;
    for (int r = 0; r < 5; r++) {
        int pr = r, pc = r;
        double best = 0.;
        for (int i = r; i < 5; i++) for (int j = r; j < 9; j++) if (fabs(Q[i][j]) > best) { best = fabs(Q[i][j]); pr = i; pc = j; }
        if (!(best > 1e-14)) return 0;
        if (pr != r) for (int j = 0; j < 9; j++) { double t = Q[r][j]; Q[r][j] = Q[pr][j]; Q[pr][j] = t; }
        if (pc != r) {
            for (int i = 0; i < 5; i++) { double t = Q[i][r]; Q[i][r] = Q[i][pc]; Q[i][pc] = t; }
            int t = colperm[r]; colperm[r] = colperm[pc]; colperm[pc] = t;
        }
        const double inv = 1. / Q[r][r];
        for (int j = 0; j < 9; j++) Q[r][j] *= inv;
        for (int i = 0; i < 5; i++) if (i != r) {
            const double f = Q[i][r];
            if (f != 0.) for (int j = 0; j < 9; j++) Q[i][j] -= f * Q[r][j];
        }
    }
    for (int f = 0; f < 4; f++) {   /* free column 5 + f (in permuted order) */
        double v[9];
        memset(v, 0, sizeof(v));
        v[5 + f] = 1.;
        for (int r = 0; r < 5; r++) v[r] = -Q[r][5 + f];
        double nn = 0.;
        for (int j = 0; j < 9; j++) nn += v[j] * v[j];
        nn = 1. / sqrt(nn);
        for (int j = 0; j < 9; j++) EE[f][colperm[j]] = v[j] * nn;
    }
    return 1;
}

/* ---- real roots of a polynomial (coefficients lowest degree first), ascending: Sturm sequence + bisection ---- */
#define MAXDEG 10
static double p_eval(const double* c, int deg, double x)
{
    double v = c[deg];
    for (int i = deg - 1; i >= 0; i--) v = v * x + c[i];
    return v;
}
static int sturm_changes(double seq[MAXDEG + 2][MAXDEG + 1], const int* degs, int nseq, double x)
{
    int changes = 0, last = 0;
    for (int s = 0; s < nseq; s++) {
        const double v = p_eval(seq[s], degs[s], x);
        const int sg = v > 0 ? 1 : (v < 0 ? -1 : 0);
        if (sg != 0) { if (last != 0 && sg != last) changes++; last = sg; }
    }
    return changes;
}
static int real_roots(const double* coef, int deg, double* roots)
{
    while (deg > 0 && coef[deg] == 0.) deg--;
    if (deg <= 0) return 0;
    double seq[MAXDEG + 2][MAXDEG + 1];
    int degs[MAXDEG + 2];
    memset(seq, 0, sizeof(seq));
    double scale = 0.;
    for (int i = 0; i <= deg; i++) if (fabs(coef[i]) > scale) scale = fabs(coef[i]);
    for (int i = 0; i <= deg; i++) seq[0][i] = coef[i] / scale;
    degs[0] = deg;
    for (int i = 1; i <= deg; i++) seq[1][i - 1] = i * seq[0][i];
    degs[1] = deg - 1;
    int nseq = 2;
    while (degs[nseq - 1] > 0) {   /* seq[k+1] = -rem(seq[k-1], seq[k]) */
        double r[MAXDEG + 1];
        const double* d = seq[nseq - 1];
        const int dd = degs[nseq - 1];
        int dr = degs[nseq - 2];
        memcpy(r, seq[nseq - 2], sizeof(r));
        while (dr >= dd) {
            const double f = r[dr] / d[dd];
            for (int i = 0; i <= dd; i++) r[dr - dd + i] -= f * d[i];
            r[dr] = 0.;
            dr--;
        }
        double mx = 0.;
        for (int i = 0; i <= dr; i++) if (fabs(r[i]) > mx) mx = fabs(r[i]);
        while (dr > 0 && fabs(r[dr]) <= 1e-14 * mx) dr--;
        if (dr < 0 || mx == 0. || (dr == 0 && fabs(r[0]) <= 1e-300)) break;   /* common factor: multiple root */
        for (int i = 0; i <= dr; i++) seq[nseq][i] = -r[i] / mx;
        degs[nseq] = dr;
        nseq++;
    }
    /* Cauchy bound */
    double bound = 0.;
    for (int i = 0; i < deg; i++) { const double v = fabs(seq[0][i] / seq[0][deg]); if (v > bound) bound = v; }
    bound += 1.;
    const int c_lo = sturm_changes(seq, degs, nseq, -bound), c_hi = sturm_changes(seq, degs, nseq, bound);
    int nroots = c_lo - c_hi, found = 0;
    if (nroots <= 0) return 0;
    /* isolate by recursive bisection (explicit stack) */
    struct { double lo, hi; int clo, chi; } st[64];
    int sp = 0;
    st[sp].lo = -bound; st[sp].hi = bound; st[sp].clo = c_lo; st[sp].chi = c_hi; sp++;
    while (sp > 0 && found < MAXDEG) {
        sp--;
        double lo = st[sp].lo, hi = st[sp].hi;
        int clo = st[sp].clo, chi = st[sp].chi;
        const int k = clo - chi;
        if (k <= 0) continue;
        if (k == 1 || hi - lo < 1e-13 * (1. + fabs(lo))) {
            /* one root (or a cluster not worth splitting): bisection on the sign of p, then Newton polish */
            double flo = p_eval(seq[0], deg, lo), fhi = p_eval(seq[0], deg, hi);
            double x = 0.5 * (lo + hi);
            if (flo * fhi < 0.) {
                for (int it = 0; it < 200 && hi - lo > 1e-15 * (1. + fabs(x)); it++) {
                    x = 0.5 * (lo + hi);
                    const double fx = p_eval(seq[0], deg, x);
                    if ((fx < 0.) == (flo < 0.)) { lo = x; flo = fx; } else { hi = x; }
                }
                x = 0.5 * (lo + hi);
            }
            for (int it = 0; it < 4; it++) {
                const double fx = p_eval(seq[0], deg, x), dx = p_eval(seq[1], deg - 1, x);
                if (dx == 0.) break;
                const double xn = x - fx / dx;
                if (!(xn > st[sp].lo && xn < st[sp].hi)) break;
                x = xn;
            }
            roots[found++] = x;
            continue;
        }
        const double mid = 0.5 * (lo + hi);
        const int cm = sturm_changes(seq, degs, nseq, mid);
        if (sp + 2 > 64) continue;
        st[sp].lo = mid; st[sp].hi = hi; st[sp].clo = cm; st[sp].chi = chi; sp++;   /* upper half pushed first: */
        st[sp].lo = lo; st[sp].hi = mid; st[sp].clo = clo; st[sp].chi = cm; sp++;   /* lower half is popped first */
    }
    return found;
}

/* ---- the five-point solver: up to 10 essential matrices (row-major, unit Frobenius norm) --------------------- */
static void poly_mul(const double* a, int da, const double* b, int db, double* out)   /* out has da + db + 1 entries */
{
    for (int i = 0; i <= da + db; i++) out[i] = 0.;
    for (int i = 0; i <= da; i++) for (int j = 0; j <= db; j++) out[i + j] += a[i] * b[j];
}

int vo_oracle_five_point(const double* q1, const double* q2, double* Es /* 10 x 9 */)
{
    double Q[5][9], EE[4][9];
    for (int i = 0; i < 5; i++) {
        const double x1 = q1[2 * i], y1 = q1[2 * i + 1], x2 = q2[2 * i], y2 = q2[2 * i + 1];
        /* x2^T E x1 = 0 with E = reshape(v, 3, 3) row-major */
        Q[i][0] = x2 * x1; Q[i][1] = x2 * y1; Q[i][2] = x2; Q[i][3] = y2 * x1; Q[i][4] = y2 * y1; Q[i][5] = y2;
        Q[i][6] = x1; Q[i][7] = y1; Q[i][8] = 1.;
    }
    if (!null_space_5x9(Q, EE)) return 0;
    /* E(x, y, z) = x EE0 + y EE1 + z EE2 + EE3, entries e[r][c] (OpenCV stores the 9-vector so that E = reshape(3, 3)) */
    Poly e[3][3];
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) {
        p_zero(&e[r][c]);
        e[r][c].c[1][0][0] = EE[0][3 * r + c]; e[r][c].c[0][1][0] = EE[1][3 * r + c];
        e[r][c].c[0][0][1] = EE[2][3 * r + c]; e[r][c].c[0][0][0] = EE[3][3 * r + c];
    }
    double A[10][20];
    {   /* det E */
        Poly m, d;
        p_zero(&d);
        p_zero(&m); p_addmul(&m, &e[1][1], &e[2][2], 1.); p_addmul(&m, &e[1][2], &e[2][1], -1.); p_addmul(&d, &e[0][0], &m, 1.);
        p_zero(&m); p_addmul(&m, &e[1][0], &e[2][2], 1.); p_addmul(&m, &e[1][2], &e[2][0], -1.); p_addmul(&d, &e[0][1], &m, -1.);
        p_zero(&m); p_addmul(&m, &e[1][0], &e[2][1], 1.); p_addmul(&m, &e[1][1], &e[2][0], -1.); p_addmul(&d, &e[0][2], &m, 1.);
        p_row(&d, A[0]);
    }
    {   /* 2 E E^T E - tr(E E^T) E */
        Poly M[3][3], tr;
        p_zero(&tr);
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) {
            p_zero(&M[r][c]);
            for (int k = 0; k < 3; k++) p_addmul(&M[r][c], &e[r][k], &e[c][k], 1.);
            if (r == c) p_axpy(&tr, &M[r][c], 1.);
        }
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) {
            Poly t;
            p_zero(&t);
            for (int k = 0; k < 3; k++) p_addmul(&t, &M[r][k], &e[k][c], 2.);
            p_addmul(&t, &tr, &e[r][c], -1.);
            p_row(&t, A[1 + 3 * r + c]);
        }
    }
    /* A := inv(A[:, :10]) A : Gauss-Jordan with partial pivoting; afterwards row i belongs to monomial i */
    for (int c = 0; c < 10; c++) {
        int pr = c;
        for (int i = c + 1; i < 10; i++) if (fabs(A[i][c]) > fabs(A[pr][c])) pr = i;
        if (!(fabs(A[pr][c]) > 1e-300)) return 0;
        if (pr != c) for (int j = 0; j < 20; j++) { double t = A[c][j]; A[c][j] = A[pr][j]; A[pr][j] = t; }
        const double inv = 1. / A[c][c];
        for (int j = 0; j < 20; j++) A[c][j] *= inv;
        for (int i = 0; i < 10; i++) if (i != c) {
            const double f = A[i][c];
            if (f != 0.) for (int j = 0; j < 20; j++) A[i][j] -= f * A[c][j];
        }
    }
    /* B(z) (3 x 3, entries cubic, cubic, quartic in z; stored highest degree first as OpenCV's b[3][13]) */
    double b[3][13];
    for (int i = 0; i < 3; i++) {
        const double* r1 = A[2 * i + 4] + 10; const double* r2 = A[2 * i + 5] + 10;
        double row1[13], row2[13];
        memset(row1, 0, sizeof(row1)); memset(row2, 0, sizeof(row2));
        for (int k = 0; k < 3; k++) { row1[1 + k] = r1[k]; row1[5 + k] = r1[3 + k]; row2[k] = r2[k]; row2[4 + k] = r2[3 + k]; }
        for (int k = 0; k < 4; k++) { row1[9 + k] = r1[6 + k]; row2[8 + k] = r2[6 + k]; }
        for (int k = 0; k < 13; k++) b[i][k] = row1[k] - row2[k];
    }
    /* det B(z), lowest degree first */
    double p0[3][4], p1[3][4], p2[3][5];
    for (int i = 0; i < 3; i++) {
        for (int k = 0; k < 4; k++) { p0[i][k] = b[i][3 - k]; p1[i][k] = b[i][7 - k]; }
        for (int k = 0; k < 5; k++) p2[i][k] = b[i][12 - k];
    }
    double det[11];
    memset(det, 0, sizeof(det));
    static const int perm[6][3] = { {0,1,2},{1,2,0},{2,0,1},{0,2,1},{2,1,0},{1,0,2} };
    static const double sgn[6] = { 1, 1, 1, -1, -1, -1 };
    for (int s = 0; s < 6; s++) {   /* rows perm[s][0..2] supply columns 0, 1, 2 */
        double t6[7], t10[11];
        poly_mul(p0[perm[s][0]], 3, p1[perm[s][1]], 3, t6);
        poly_mul(t6, 6, p2[perm[s][2]], 4, t10);
        for (int k = 0; k <= 10; k++) det[k] += sgn[s] * t10[k];
    }
    double zs[MAXDEG];
    const int nz = real_roots(det, 10, zs);
    int count = 0;
    for (int ri = 0; ri < nz && count < 10; ri++) {
        const double z = zs[ri], z2 = z * z, z3 = z2 * z, z4 = z3 * z;
        double bz[3][3];
        for (int j = 0; j < 3; j++) {
            bz[j][0] = b[j][0] * z3 + b[j][1] * z2 + b[j][2] * z + b[j][3];
            bz[j][1] = b[j][4] * z3 + b[j][5] * z2 + b[j][6] * z + b[j][7];
            bz[j][2] = b[j][8] * z4 + b[j][9] * z3 + b[j][10] * z2 + b[j][11] * z + b[j][12];
        }
        /* unit null vector of bz: the largest cross product of two rows */
        double best[3] = { 0, 0, 0 }, bn = -1.;
        for (int a = 0; a < 3; a++) for (int c = a + 1; c < 3; c++) {
            const double v[3] = { bz[a][1] * bz[c][2] - bz[a][2] * bz[c][1], bz[a][2] * bz[c][0] - bz[a][0] * bz[c][2],
                                  bz[a][0] * bz[c][1] - bz[a][1] * bz[c][0] };
            const double nn = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
            if (nn > bn) { bn = nn; best[0] = v[0]; best[1] = v[1]; best[2] = v[2]; }
        }
        if (!(bn > 0.)) continue;
        const double inv = 1. / sqrt(bn);
        if (fabs(best[2] * inv) < 1e-10) continue;
        const double x = best[0] / best[2], y = best[1] / best[2];
        double E[9], nn = 0.;
        for (int k = 0; k < 9; k++) { E[k] = x * EE[0][k] + y * EE[1][k] + z * EE[2][k] + EE[3][k]; nn += E[k] * E[k]; }
        nn = 1. / sqrt(nn);
        for (int k = 0; k < 9; k++) Es[9 * count + k] = E[k] * nn;
        count++;
    }
    return count;
}

/* Sampson error of OpenCV's EMEstimatorCallback::computeError, float32 */
static float sampson(const double* E, double x1, double y1, double x2, double y2)
{
    const double Ex1[3] = { E[0] * x1 + E[1] * y1 + E[2], E[3] * x1 + E[4] * y1 + E[5], E[6] * x1 + E[7] * y1 + E[8] };
    const double Etx2[2] = { E[0] * x2 + E[3] * y2 + E[6], E[1] * x2 + E[4] * y2 + E[7] };
    const double x2tEx1 = x2 * Ex1[0] + y2 * Ex1[1] + Ex1[2];
    const double a = Ex1[0] * Ex1[0], b = Ex1[1] * Ex1[1], c = Etx2[0] * Etx2[0], d = Etx2[1] * Etx2[1];
    return (float)(x2tEx1 * x2tEx1 / (a + b + c + d));
}

/* cv2.findEssentialMat(p1, p2, K, RANSAC, prob, thr, maxIters): E (3x3 row-major) and mask; info = {iterations run,
 * winning iteration, winning model of that iteration, inliers}.  Returns 1 if a model was found. */
int vo_oracle_find_essential_mat(const double* p1, const double* p2, int n, const double* K, double prob, double thr,
                                 int max_iters, double* E_out, uint8_t* mask, int32_t* info)
{
    const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    double* q1 = (double*)malloc(sizeof(double) * 2 * (n > 0 ? n : 1));
    double* q2 = (double*)malloc(sizeof(double) * 2 * (n > 0 ? n : 1));
    uint8_t* cur = (uint8_t*)malloc(n > 0 ? n : 1);
    for (int i = 0; i < n; i++) {
        q1[2 * i] = (p1[2 * i] - cx) / fx; q1[2 * i + 1] = (p1[2 * i + 1] - cy) / fy;
        q2[2 * i] = (p2[2 * i] - cx) / fx; q2[2 * i + 1] = (p2[2 * i + 1] - cy) / fy;
    }
    thr /= (fx + fy) / 2;
    const float t2 = (float)(thr * thr);
    int niters = max_iters > 1 ? max_iters : 1, best = 0, best_iter = -1, best_model = -1, iters_run = 0, ok = 0;
    memset(mask, 0, n);
    if (n < 5) goto done;
    if (n == 5) {
        double Es[90];
        if (vo_oracle_five_point(q1, q2, Es) > 0) { memcpy(E_out, Es, 72); memset(mask, 1, n); best = 5; ok = 1; best_iter = 0; best_model = 0; }
        goto done;
    }
    {
        uint64_t rng = 0xFFFFFFFFFFFFFFFFULL;
        for (int iter = 0; iter < niters; iter++) {
            iters_run = iter + 1;
            int idx[5];
            for (int i = 0; i < 5; i++) {
                for (;;) {
                    const int x = (int)(rng_next(&rng) % (unsigned)n);
                    int dup = 0;
                    for (int j = 0; j < i; j++) dup |= idx[j] == x;
                    if (!dup) { idx[i] = x; break; }
                }
            }
            double s1[10], s2[10], Es[90];
            for (int i = 0; i < 5; i++) { s1[2 * i] = q1[2 * idx[i]]; s1[2 * i + 1] = q1[2 * idx[i] + 1]; s2[2 * i] = q2[2 * idx[i]]; s2[2 * i + 1] = q2[2 * idx[i] + 1]; }
            const int nm = vo_oracle_five_point(s1, s2, Es);
            for (int m = 0; m < nm; m++) {
                int good = 0;
                for (int i = 0; i < n; i++) {
                    const int f = sampson(Es + 9 * m, q1[2 * i], q1[2 * i + 1], q2[2 * i], q2[2 * i + 1]) <= t2;
                    cur[i] = (uint8_t)f; good += f;
                }
                if (good > (best > 4 ? best : 4)) {
                    memcpy(mask, cur, n); memcpy(E_out, Es + 9 * m, 72);
                    best = good; best_iter = iter; best_model = m; ok = 1;
                    niters = update_num_iters(prob, (double)(n - good) / n, 5, niters);
                }
            }
        }
    }
done:
    if (info) { info[0] = iters_run; info[1] = best_iter; info[2] = best_model; info[3] = ok ? best : 0; }
    free(q1); free(q2); free(cur);
    return ok;
}

/* ---- recoverPose ----------------------------------------------------------------------------------------------- */
/* one-sided Jacobi SVD of a 3x3: A = U diag(w) V^T, singular values descending */
static void svd3(const double* A, double* U, double* w, double* V)
{
    double a[3][3], v[3][3];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { a[i][j] = A[3 * i + j]; v[i][j] = i == j; }
    for (int sweep = 0; sweep < 60; sweep++) {
        int rotated = 0;
        for (int p = 0; p < 2; p++) for (int q = p + 1; q < 3; q++) {
            double app = 0, aqq = 0, apq = 0;
            for (int r = 0; r < 3; r++) { app += a[r][p] * a[r][p]; aqq += a[r][q] * a[r][q]; apq += a[r][p] * a[r][q]; }
            if (apq * apq > 1e-32 * (app * aqq) && fabs(apq) > 1e-300) {
                rotated = 1;
                const double d = aqq - app;
                const double t = (d >= 0 ? 2. : -2.) * apq / (fabs(d) + sqrt(d * d + 4 * apq * apq));
                const double cs = 1. / sqrt(1 + t * t), sn = cs * t;
                for (int r = 0; r < 3; r++) {
                    const double ap = a[r][p], aq = a[r][q];
                    a[r][p] = cs * ap - sn * aq; a[r][q] = sn * ap + cs * aq;
                    const double vp = v[r][p], vq = v[r][q];
                    v[r][p] = cs * vp - sn * vq; v[r][q] = sn * vp + cs * vq;
                }
            }
        }
        if (!rotated) break;
    }
    double nn[3];
    int ord[3] = { 0, 1, 2 };
    for (int c = 0; c < 3; c++) nn[c] = sqrt(a[0][c] * a[0][c] + a[1][c] * a[1][c] + a[2][c] * a[2][c]);
    for (int i = 0; i < 2; i++) for (int j = i + 1; j < 3; j++) if (nn[ord[j]] > nn[ord[i]]) { int t = ord[i]; ord[i] = ord[j]; ord[j] = t; }
    for (int k = 0; k < 3; k++) {
        const int c = ord[k];
        w[k] = nn[c];
        for (int r = 0; r < 3; r++) { V[3 * r + k] = v[r][c]; U[3 * r + k] = nn[c] > 1e-300 ? a[r][c] / nn[c] : 0.; }
    }
    /* a vanishing singular value leaves its column of U undefined: complete the basis */
    if (w[2] <= 1e-12 * w[0]) {
        U[2] = U[3] * U[7] - U[6] * U[4]; U[5] = U[6] * U[1] - U[0] * U[7]; U[8] = U[0] * U[4] - U[3] * U[1];
    }
}
static double det3(const double* M)
{
    return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}
void vo_oracle_triangulate_points(const double* M1, const double* M2, const double* p1, const double* p2, int n, double* X4);

void vo_oracle_decompose_essential(const double* E, double* R1, double* R2, double* t)
{
    double U[9], w[3], V[9], Vt[9];
    svd3(E, U, w, V);
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) Vt[3 * i + j] = V[3 * j + i];
    if (det3(U) < 0) for (int k = 0; k < 9; k++) U[k] = -U[k];
    if (det3(Vt) < 0) for (int k = 0; k < 9; k++) Vt[k] = -Vt[k];
    static const double W[9] = { 0, 1, 0, -1, 0, 0, 0, 0, 1 };
    double UW[9], UWt[9];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
        double s = 0, st = 0;
        for (int k = 0; k < 3; k++) { s += U[3 * i + k] * W[3 * k + j]; st += U[3 * i + k] * W[3 * j + k]; }
        UW[3 * i + j] = s; UWt[3 * i + j] = st;
    }
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
        double s = 0, st = 0;
        for (int k = 0; k < 3; k++) { s += UW[3 * i + k] * Vt[3 * k + j]; st += UWt[3 * i + k] * Vt[3 * k + j]; }
        R1[3 * i + j] = s; R2[3 * i + j] = st;
    }
    t[0] = U[2]; t[1] = U[5]; t[2] = U[8];
}

/* cv2.recoverPose(E, p1, p2, K, mask=mask): R (3x3), t (3), mask ANDed with the winner's cheirality mask; returns the
 * winner's count.  which (optional) = index 0..3 of the winning (R, t) candidate in OpenCV's order. */
int vo_oracle_recover_pose(const double* E, const double* p1, const double* p2, int n, const double* K, uint8_t* mask,
                           double* R, double* t, int32_t* which)
{
    const double fx = K[0], fy = K[4], cx = K[2], cy = K[5], dist = 50.;
    double R1[9], R2[9], tt[3];
    vo_oracle_decompose_essential(E, R1, R2, tt);
    double* q1 = (double*)malloc(sizeof(double) * 2 * (n > 0 ? n : 1));
    double* q2 = (double*)malloc(sizeof(double) * 2 * (n > 0 ? n : 1));
    double* X4 = (double*)malloc(sizeof(double) * 4 * (n > 0 ? n : 1));
    uint8_t* m[4];
    for (int k = 0; k < 4; k++) m[k] = (uint8_t*)malloc(n > 0 ? n : 1);
    for (int i = 0; i < n; i++) {
        q1[2 * i] = (p1[2 * i] - cx) / fx; q1[2 * i + 1] = (p1[2 * i + 1] - cy) / fy;
        q2[2 * i] = (p2[2 * i] - cx) / fx; q2[2 * i + 1] = (p2[2 * i + 1] - cy) / fy;
    }
    const double P0[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };
    int good[4];
    for (int k = 0; k < 4; k++) {
        const double* Rk = (k & 1) ? R2 : R1;
        const double s = k >= 2 ? -1. : 1.;
        double P[12];
        for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) P[4 * r + c] = Rk[3 * r + c]; P[4 * r + 3] = s * tt[r]; }
        vo_oracle_triangulate_points(P0, P, q1, q2, n, X4);
        good[k] = 0;
        for (int i = 0; i < n; i++) {
            const double* Q = X4 + 4 * i;
            int f = Q[2] * Q[3] > 0;
            const double x = Q[0] / Q[3], y = Q[1] / Q[3], z = Q[2] / Q[3];
            f = f && z < dist;
            const double z2 = P[8] * x + P[9] * y + P[10] * z + P[11];
            f = f && z2 > 0 && z2 < dist;
            if (mask) f = f && mask[i];
            m[k][i] = (uint8_t)f;
            good[k] += f;
        }
    }
    int win;
    if (good[0] >= good[1] && good[0] >= good[2] && good[0] >= good[3]) win = 0;
    else if (good[1] >= good[0] && good[1] >= good[2] && good[1] >= good[3]) win = 1;
    else if (good[2] >= good[0] && good[2] >= good[1] && good[2] >= good[3]) win = 2;
    else win = 3;
    memcpy(R, (win & 1) ? R2 : R1, 72);
    for (int r = 0; r < 3; r++) t[r] = win >= 2 ? -tt[r] : tt[r];
    if (mask) memcpy(mask, m[win], n);
    if (which) *which = win;
    const int g = good[win];
    for (int k = 0; k < 4; k++) free(m[k]);
    free(q1); free(q2); free(X4);
    return g;
}
