// the bootstrap on the device.
//
// Replaces, in VOInitializer.getPoseEstimate (/root/reference/src/vo_pipeline.py:139-162) and in the "< 4 inliers"
// re-bootstrap of PoseEstimator.estimatePose (:274-324), the two OpenCV calls
//     E, mask = cv2.findEssentialMat(kps1, kps2, K, cv2.RANSAC, 0.999, 1.0, None)
//     _, R, t, _ = cv2.recoverPose(E, kps1, kps2, K, mask=mask)      (ANDs its cheirality test into `mask` in place)
// following OpenCV's behaviour (modules/calib3d/src/five-point.cpp, ptsetreg.cpp; restated and pinned to cv2 in
// oracle/fivept.c):
//   - minimal sets: 5 indices per iteration from cv::RNG(2^64-1) with duplicate re-draws - the match count is
//     known on the host here, so the table is drawn there (exactly, 20 KB) and uploaded;
//   - em_solve_kernel: one thread per hypothesis - null space of the 5x9 epipolar system (Gauss-Jordan, complete
//     pivoting), Nister's ten cubic constraints expanded over 20 monomials, eliminated to the 3x3 polynomial matrix
//     B(z), det B(z) = 0 (degree 10) solved for its real roots by a Sturm sequence + bisection + Newton polish,
//     up to 10 essential matrices per hypothesis;
//   - em_score_kernel: one CTA per hypothesis, Sampson error (float64 arithmetic, float32 result, <= (float)thr^2)
//     of every match against every model of the hypothesis;
//   - em_select_kernel: the sequential rule of the reference loop (first strictly better model wins, the
//     iteration budget shrinks with every new best; model points = 5) replayed over the counts, then the mask;
//   - recover_pose_kernel: SVD of E, the four (R, t) candidates, DLT of every match under each (dlt.cuh, the
//     promotion step's routine), cheirality with OpenCV's distance threshold 50, the winner's mask ANDed in.
// All hypotheses of the budget are solved up front (there is no point in an adaptive stop for a step that runs
// once per sequence): FP64 CUDA-core work, no tensor cores.
#include <float.h>
#include "ctx.cuh"
#include "dlt.cuh"

namespace vo {

namespace {

constexpr int kMaxModels = 10;

// ---- polynomials in (x, y, z) of total degree <= 3, dense over exponents ------------------------------------------
struct Poly { double c[4][4][4]; };

__device__ void p_zero(Poly* p)
{
    for (int i = 0; i < 64; i++) (&p->c[0][0][0])[i] = 0.;
}
__device__ void p_addmul(Poly* out, const Poly* a, const Poly* b, double s)   // out += s * a * b (degrees must fit)
{
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4 - i; j++) for (int k = 0; k < 4 - i - j; k++) {
        const double av = a->c[i][j][k];
        if (av == 0.) continue;
        for (int l = 0; i + j + k + l < 4; l++) for (int m = 0; i + j + k + l + m < 4; m++)
            for (int n = 0; i + j + k + l + m + n < 4; n++) {
                const double bv = b->c[l][m][n];
                if (bv != 0.) out->c[i + l][j + m][k + n] += s * av * bv;
            }
    }
}
__device__ void p_axpy(Poly* out, const Poly* a, double s)
{
    for (int i = 0; i < 64; i++) (&out->c[0][0][0])[i] += s * (&a->c[0][0][0])[i];
}

// Nister's monomial order: x^3 y^3 x^2y xy^2 x^2z x^2 y^2z y^2 xyz xy | xz^2 xz x yz^2 yz y z^3 z^2 z 1
__constant__ int kMono[20][3] = { {3,0,0},{0,3,0},{2,1,0},{1,2,0},{2,0,1},{2,0,0},{0,2,1},{0,2,0},{1,1,1},{1,1,0},
                                  {1,0,2},{1,0,1},{1,0,0},{0,1,2},{0,1,1},{0,1,0},{0,0,3},{0,0,2},{0,0,1},{0,0,0} };

__device__ void p_row(const Poly* p, double* row)
{
    for (int m = 0; m < 20; m++) row[m] = p->c[kMono[m][0]][kMono[m][1]][kMono[m][2]];
}

__device__ bool null_space_5x9(double Q[5][9], double EE[4][9])
{
    int colperm[9];
    for (int j = 0; j < 9; j++) colperm[j] = j;
    for (int r = 0; r < 5; r++) {
        int pr = r, pc = r;
        double best = 0.;
        for (int i = r; i < 5; i++) for (int j = r; j < 9; j++) if (fabs(Q[i][j]) > best) { best = fabs(Q[i][j]); pr = i; pc = j; }
        if (!(best > 1e-14)) return false;
        if (pr != r) for (int j = 0; j < 9; j++) { const double t = Q[r][j]; Q[r][j] = Q[pr][j]; Q[pr][j] = t; }
        if (pc != r) {
            for (int i = 0; i < 5; i++) { const double t = Q[i][r]; Q[i][r] = Q[i][pc]; Q[i][pc] = t; }
            const int t = colperm[r]; colperm[r] = colperm[pc]; colperm[pc] = t;
        }
        const double inv = 1. / Q[r][r];
        for (int j = 0; j < 9; j++) Q[r][j] *= inv;
        for (int i = 0; i < 5; i++) if (i != r) {
            const double f = Q[i][r];
            if (f != 0.) for (int j = 0; j < 9; j++) Q[i][j] -= f * Q[r][j];
        }
    }
    for (int f = 0; f < 4; f++) {
        double v[9];
        for (int j = 0; j < 9; j++) v[j] = 0.;
        v[5 + f] = 1.;
        for (int r = 0; r < 5; r++) v[r] = -Q[r][5 + f];
        double nn = 0.;
        for (int j = 0; j < 9; j++) nn += v[j] * v[j];
        nn = 1. / sqrt(nn);
        for (int j = 0; j < 9; j++) EE[f][colperm[j]] = v[j] * nn;
    }
    return true;
}

// ---- real roots (ascending) of a polynomial given lowest degree first: Sturm sequence + bisection + Newton ------
constexpr int kMaxDeg = 10;

__device__ double p_eval(const double* c, int deg, double x)
{
    double v = c[deg];
    for (int i = deg - 1; i >= 0; i--) v = v * x + c[i];
    return v;
}
__device__ int sturm_changes(double (*seq)[kMaxDeg + 1], const int* degs, int nseq, double x)
{
    int changes = 0, last = 0;
    for (int s = 0; s < nseq; s++) {
        const double v = p_eval(seq[s], degs[s], x);
        const int sg = v > 0 ? 1 : (v < 0 ? -1 : 0);
        if (sg != 0) { if (last != 0 && sg != last) changes++; last = sg; }
    }
    return changes;
}
__device__ int real_roots(const double* coef, int deg, double* roots)
{
    while (deg > 0 && coef[deg] == 0.) deg--;
    if (deg <= 0) return 0;
    double seq[kMaxDeg + 2][kMaxDeg + 1];
    int degs[kMaxDeg + 2];
    for (int s = 0; s < kMaxDeg + 2; s++) for (int i = 0; i <= kMaxDeg; i++) seq[s][i] = 0.;
    double scale = 0.;
    for (int i = 0; i <= deg; i++) if (fabs(coef[i]) > scale) scale = fabs(coef[i]);
    for (int i = 0; i <= deg; i++) seq[0][i] = coef[i] / scale;
    degs[0] = deg;
    for (int i = 1; i <= deg; i++) seq[1][i - 1] = i * seq[0][i];
    degs[1] = deg - 1;
    int nseq = 2;
    while (degs[nseq - 1] > 0) {   // seq[k+1] = -rem(seq[k-1], seq[k])
        double r[kMaxDeg + 1];
        const double* d = seq[nseq - 1];
        const int dd = degs[nseq - 1];
        int dr = degs[nseq - 2];
        for (int i = 0; i <= kMaxDeg; i++) r[i] = seq[nseq - 2][i];
        while (dr >= dd) {
            const double f = r[dr] / d[dd];
            for (int i = 0; i <= dd; i++) r[dr - dd + i] -= f * d[i];
            r[dr] = 0.;
            dr--;
        }
        double mx = 0.;
        for (int i = 0; i <= dr; i++) if (fabs(r[i]) > mx) mx = fabs(r[i]);
        while (dr > 0 && fabs(r[dr]) <= 1e-14 * mx) dr--;
        if (dr < 0 || mx == 0. || (dr == 0 && fabs(r[0]) <= 1e-300)) break;   // common factor: multiple root
        for (int i = 0; i <= dr; i++) seq[nseq][i] = -r[i] / mx;
        degs[nseq] = dr;
        nseq++;
    }
    double bound = 0.;   // Cauchy
    for (int i = 0; i < deg; i++) { const double v = fabs(seq[0][i] / seq[0][deg]); if (v > bound) bound = v; }
    bound += 1.;
    const int c_lo = sturm_changes(seq, degs, nseq, -bound), c_hi = sturm_changes(seq, degs, nseq, bound);
    if (c_lo - c_hi <= 0) return 0;
    struct Iv { double lo, hi; int clo, chi; } st[64];
    int sp = 0, found = 0;
    st[sp++] = Iv{ -bound, bound, c_lo, c_hi };
    while (sp > 0 && found < kMaxDeg) {
        const Iv iv = st[--sp];
        double lo = iv.lo, hi = iv.hi;
        const int k = iv.clo - iv.chi;
        if (k <= 0) continue;
        if (k == 1 || hi - lo < 1e-13 * (1. + fabs(lo))) {
            double flo = p_eval(seq[0], deg, lo);
            const double fhi = p_eval(seq[0], deg, hi);
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
                if (!(xn > iv.lo && xn < iv.hi)) break;
                x = xn;
            }
            roots[found++] = x;
            continue;
        }
        const double mid = 0.5 * (lo + hi);
        const int cm = sturm_changes(seq, degs, nseq, mid);
        if (sp + 2 > 64) continue;
        st[sp++] = Iv{ mid, hi, cm, iv.chi };   // the upper half is pushed first: the lower half is popped first
        st[sp++] = Iv{ lo, mid, iv.clo, cm };
    }
    return found;
}

__device__ void poly_mul(const double* a, int da, const double* b, int db, double* out)
{
    for (int i = 0; i <= da + db; i++) out[i] = 0.;
    for (int i = 0; i <= da; i++) for (int j = 0; j <= db; j++) out[i + j] += a[i] * b[j];
}

// up to 10 essential matrices (row-major, unit Frobenius norm) through 5 normalised correspondences
__device__ int five_point(const double* q1, const double* q2, double* Es)
{
    double Q[5][9], EE[4][9];
    for (int i = 0; i < 5; i++) {
        const double x1 = q1[2 * i], y1 = q1[2 * i + 1], x2 = q2[2 * i], y2 = q2[2 * i + 1];
        Q[i][0] = x2 * x1; Q[i][1] = x2 * y1; Q[i][2] = x2; Q[i][3] = y2 * x1; Q[i][4] = y2 * y1; Q[i][5] = y2;
        Q[i][6] = x1; Q[i][7] = y1; Q[i][8] = 1.;
    }
    if (!null_space_5x9(Q, EE)) return 0;
    Poly e[3][3];
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) {
        p_zero(&e[r][c]);
        e[r][c].c[1][0][0] = EE[0][3 * r + c]; e[r][c].c[0][1][0] = EE[1][3 * r + c];
        e[r][c].c[0][0][1] = EE[2][3 * r + c]; e[r][c].c[0][0][0] = EE[3][3 * r + c];
    }
    double A[10][20];
    {   // det E
        Poly m, d;
        p_zero(&d);
        p_zero(&m); p_addmul(&m, &e[1][1], &e[2][2], 1.); p_addmul(&m, &e[1][2], &e[2][1], -1.); p_addmul(&d, &e[0][0], &m, 1.);
        p_zero(&m); p_addmul(&m, &e[1][0], &e[2][2], 1.); p_addmul(&m, &e[1][2], &e[2][0], -1.); p_addmul(&d, &e[0][1], &m, -1.);
        p_zero(&m); p_addmul(&m, &e[1][0], &e[2][1], 1.); p_addmul(&m, &e[1][1], &e[2][0], -1.); p_addmul(&d, &e[0][2], &m, 1.);
        p_row(&d, A[0]);
    }
    {   // 2 E E^T E - tr(E E^T) E
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
    for (int c = 0; c < 10; c++) {   // A := inv(A[:, :10]) A
        int pr = c;
        for (int i = c + 1; i < 10; i++) if (fabs(A[i][c]) > fabs(A[pr][c])) pr = i;
        if (!(fabs(A[pr][c]) > 1e-300)) return 0;
        if (pr != c) for (int j = 0; j < 20; j++) { const double t = A[c][j]; A[c][j] = A[pr][j]; A[pr][j] = t; }
        const double inv = 1. / A[c][c];
        for (int j = 0; j < 20; j++) A[c][j] *= inv;
        for (int i = 0; i < 10; i++) if (i != c) {
            const double f = A[i][c];
            if (f != 0.) for (int j = 0; j < 20; j++) A[i][j] -= f * A[c][j];
        }
    }
    double b[3][13];
    for (int i = 0; i < 3; i++) {
        const double* r1 = A[2 * i + 4] + 10; const double* r2 = A[2 * i + 5] + 10;
        double row1[13], row2[13];
        for (int k = 0; k < 13; k++) { row1[k] = 0.; row2[k] = 0.; }
        for (int k = 0; k < 3; k++) { row1[1 + k] = r1[k]; row1[5 + k] = r1[3 + k]; row2[k] = r2[k]; row2[4 + k] = r2[3 + k]; }
        for (int k = 0; k < 4; k++) { row1[9 + k] = r1[6 + k]; row2[8 + k] = r2[6 + k]; }
        for (int k = 0; k < 13; k++) b[i][k] = row1[k] - row2[k];
    }
    double p0[3][4], p1[3][4], p2[3][5];
    for (int i = 0; i < 3; i++) {
        for (int k = 0; k < 4; k++) { p0[i][k] = b[i][3 - k]; p1[i][k] = b[i][7 - k]; }
        for (int k = 0; k < 5; k++) p2[i][k] = b[i][12 - k];
    }
    double det[11];
    for (int k = 0; k <= 10; k++) det[k] = 0.;
    const int perm[6][3] = { {0,1,2},{1,2,0},{2,0,1},{0,2,1},{2,1,0},{1,0,2} };
    const double sgn[6] = { 1, 1, 1, -1, -1, -1 };
    for (int s = 0; s < 6; s++) {
        double t6[7], t10[11];
        poly_mul(p0[perm[s][0]], 3, p1[perm[s][1]], 3, t6);
        poly_mul(t6, 6, p2[perm[s][2]], 4, t10);
        for (int k = 0; k <= 10; k++) det[k] += sgn[s] * t10[k];
    }
    double zs[kMaxDeg];
    const int nz = real_roots(det, 10, zs);
    int count = 0;
    for (int ri = 0; ri < nz && count < kMaxModels; ri++) {
        const double z = zs[ri], z2 = z * z, z3 = z2 * z, z4 = z3 * z;
        double bz[3][3];
        for (int j = 0; j < 3; j++) {
            bz[j][0] = b[j][0] * z3 + b[j][1] * z2 + b[j][2] * z + b[j][3];
            bz[j][1] = b[j][4] * z3 + b[j][5] * z2 + b[j][6] * z + b[j][7];
            bz[j][2] = b[j][8] * z4 + b[j][9] * z3 + b[j][10] * z2 + b[j][11] * z + b[j][12];
        }
        double best[3] = { 0, 0, 0 }, bn = -1.;
        for (int a = 0; a < 3; a++) for (int c = a + 1; c < 3; c++) {
            const double v[3] = { bz[a][1] * bz[c][2] - bz[a][2] * bz[c][1], bz[a][2] * bz[c][0] - bz[a][0] * bz[c][2],
                                  bz[a][0] * bz[c][1] - bz[a][1] * bz[c][0] };
            const double nn = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
            if (nn > bn) { bn = nn; best[0] = v[0]; best[1] = v[1]; best[2] = v[2]; }
        }
        if (!(bn > 0.)) continue;
        if (fabs(best[2] / sqrt(bn)) < 1e-10) continue;
        const double x = best[0] / best[2], y = best[1] / best[2];
        double E[9], nn = 0.;
        for (int k = 0; k < 9; k++) { E[k] = x * EE[0][k] + y * EE[1][k] + z * EE[2][k] + EE[3][k]; nn += E[k] * E[k]; }
        nn = 1. / sqrt(nn);
        for (int k = 0; k < 9; k++) Es[9 * count + k] = E[k] * nn;
        count++;
    }
    return count;
}

__device__ __forceinline__ float sampson(const double* E, double x1, double y1, double x2, double y2)
{
    const double a0 = E[0] * x1 + E[1] * y1 + E[2], a1 = E[3] * x1 + E[4] * y1 + E[5], a2 = E[6] * x1 + E[7] * y1 + E[8];
    const double b0 = E[0] * x2 + E[3] * y2 + E[6], b1 = E[1] * x2 + E[4] * y2 + E[7];
    const double e = x2 * a0 + y2 * a1 + a2;
    return (float)(e * e / (a0 * a0 + a1 * a1 + b0 * b0 + b1 * b1));
}

struct BootScalars {
    int32_t ok, iters_run, best_iter, best_model, n_inliers, n_good, which, pad;
    double E[9], R[9], t[3];
};

__global__ void em_normalise_kernel(const double2* __restrict__ p1, const double2* __restrict__ p2, int n, double fx, double fy,
                                    double cx, double cy, double2* __restrict__ q1, double2* __restrict__ q2)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    q1[i] = make_double2((p1[i].x - cx) / fx, (p1[i].y - cy) / fy);
    q2[i] = make_double2((p2[i].x - cx) / fx, (p2[i].y - cy) / fy);
}

__global__ void __launch_bounds__(32)
em_solve_kernel(const double2* __restrict__ q1, const double2* __restrict__ q2, int n, const int32_t* __restrict__ sets, int H,
                double* __restrict__ models, int32_t* __restrict__ nmodels)
{
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    double s1[10], s2[10], Es[9 * kMaxModels];
    bool ok = true;
    for (int i = 0; i < 5; i++) {
        const int idx = n == 5 ? i : sets[5 * h + i];
        ok &= idx >= 0 && idx < n;
        const double2 a = q1[ok ? idx : 0], b = q2[ok ? idx : 0];
        s1[2 * i] = a.x; s1[2 * i + 1] = a.y; s2[2 * i] = b.x; s2[2 * i + 1] = b.y;
    }
    const int m = ok ? five_point(s1, s2, Es) : 0;
    nmodels[h] = m;
    for (int k = 0; k < 9 * m; k++) models[(size_t)h * 9 * kMaxModels + k] = Es[k];
}

__global__ void __launch_bounds__(128)
em_score_kernel(const double2* __restrict__ q1, const double2* __restrict__ q2, int n, const double* __restrict__ models,
                const int32_t* __restrict__ nmodels, float t2, int32_t* __restrict__ counts)
{
    __shared__ double sE[9 * kMaxModels];
    __shared__ int sCnt[kMaxModels];
    const int h = blockIdx.x, tid = threadIdx.x;
    const int m = nmodels[h];
    for (int k = tid; k < 9 * m; k += blockDim.x) sE[k] = models[(size_t)h * 9 * kMaxModels + k];
    if (tid < kMaxModels) sCnt[tid] = 0;
    __syncthreads();
    int good[kMaxModels];
#pragma unroll
    for (int j = 0; j < kMaxModels; j++) good[j] = 0;
    for (int i = tid; i < n; i += blockDim.x) {
        const double2 a = q1[i], b = q2[i];
#pragma unroll
        for (int j = 0; j < kMaxModels; j++)
            if (j < m) good[j] += sampson(sE + 9 * j, a.x, a.y, b.x, b.y) <= t2;
    }
#pragma unroll
    for (int j = 0; j < kMaxModels; j++) {
        const int g = __reduce_add_sync(0xffffffffu, good[j]);
        if ((tid & 31) == 0 && g) atomicAdd(&sCnt[j], g);
    }
    __syncthreads();
    if (tid < kMaxModels) counts[h * kMaxModels + tid] = tid < m ? sCnt[tid] : -1;
}

__device__ int update_num_iters5(double p, double ep, int max_iters)
{
    p = fmin(fmax(p, 0.), 1.);
    ep = fmin(fmax(ep, 0.), 1.);
    double num = fmax(1. - p, DBL_MIN);
    double denom = 1. - pow(1. - ep, 5.);
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : (int)rint(num / denom);
}

// one CTA: thread 0 replays the sequential bookkeeping, then all threads write the winner's mask
__global__ void __launch_bounds__(256)
em_select_kernel(BootScalars* bs, const double2* __restrict__ q1, const double2* __restrict__ q2, int n, const double* __restrict__ models,
                 const int32_t* __restrict__ nmodels, const int32_t* __restrict__ counts, int H, double prob, float t2,
                 uint8_t* __restrict__ mask)
{
    __shared__ double sE[9];
    __shared__ int sOk;
    const int tid = threadIdx.x;
    if (tid == 0) {
        int niters = n < 5 ? 0 : (n == 5 ? 1 : H), best = 0, bi = -1, bm = -1, iter = 0;
        for (; iter < niters; iter++) {
            const int m = nmodels[iter];
            for (int j = 0; j < m; j++) {
                const int g = n == 5 ? 5 : counts[iter * kMaxModels + j];
                if (n == 5 ? bi < 0 : g > max(best, 4)) {
                    best = g; bi = iter; bm = j;
                    if (n > 5) niters = update_num_iters5(prob, (double)(n - g) / n, niters);
                }
            }
        }
        bs->ok = bi >= 0; bs->iters_run = iter; bs->best_iter = bi; bs->best_model = bm; bs->n_inliers = bi >= 0 ? best : 0;
        sOk = bi >= 0;
        for (int k = 0; k < 9; k++) {
            const double v = bi >= 0 ? models[(size_t)bi * 9 * kMaxModels + 9 * bm + k] : 0.;
            sE[k] = v; bs->E[k] = v;
        }
    }
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x)
        mask[i] = sOk && (n == 5 || sampson(sE, q1[i].x, q1[i].y, q2[i].x, q2[i].y) <= t2);
}

// ---- recoverPose ----------------------------------------------------------------------------------------------------
__device__ double det3(const double* M)
{
    return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// one-sided Jacobi SVD of a 3x3: A = U diag(w) V^T, singular values descending
__device__ void svd3(const double* A, double* U, double* w, double* V)
{
    double a[3][3], v[3][3];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { a[i][j] = A[3 * i + j]; v[i][j] = i == j ? 1. : 0.; }
    for (int sweep = 0; sweep < 60; sweep++) {
        bool rotated = false;
        for (int p = 0; p < 2; p++) for (int q = p + 1; q < 3; q++) {
            double app = 0, aqq = 0, apq = 0;
            for (int r = 0; r < 3; r++) { app += a[r][p] * a[r][p]; aqq += a[r][q] * a[r][q]; apq += a[r][p] * a[r][q]; }
            if (apq * apq > 1e-32 * (app * aqq) && fabs(apq) > 1e-300) {
                rotated = true;
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
    for (int i = 0; i < 2; i++) for (int j = i + 1; j < 3; j++) if (nn[ord[j]] > nn[ord[i]]) { const int t = ord[i]; ord[i] = ord[j]; ord[j] = t; }
    for (int k = 0; k < 3; k++) {
        const int c = ord[k];
        w[k] = nn[c];
        for (int r = 0; r < 3; r++) { V[3 * r + k] = v[r][c]; U[3 * r + k] = nn[c] > 1e-300 ? a[r][c] / nn[c] : 0.; }
    }
    if (w[2] <= 1e-12 * w[0]) {   // a vanishing singular value leaves its column of U undefined: complete the basis
        U[2] = U[3] * U[7] - U[6] * U[4]; U[5] = U[6] * U[1] - U[0] * U[7]; U[8] = U[0] * U[4] - U[3] * U[1];
    }
}

__global__ void __launch_bounds__(256)
recover_pose_kernel(BootScalars* bs, const double2* __restrict__ q1, const double2* __restrict__ q2, int n, uint8_t* __restrict__ mask,
                    uint8_t* __restrict__ flags /* n x 4 scratch */)
{
    __shared__ double sP[4][12];
    __shared__ int sGood[4], sWin;
    const int tid = threadIdx.x;
    if (tid == 0) {
        double U[9], w[3], V[9], Vt[9], R1[9], R2[9];
        svd3(bs->E, U, w, V);
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) Vt[3 * i + j] = V[3 * j + i];
        if (det3(U) < 0) for (int k = 0; k < 9; k++) U[k] = -U[k];
        if (det3(Vt) < 0) for (int k = 0; k < 9; k++) Vt[k] = -Vt[k];
        const double W[9] = { 0, 1, 0, -1, 0, 0, 0, 0, 1 };
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
        for (int k = 0; k < 4; k++) {
            const double* Rk = (k & 1) ? R2 : R1;
            const double s = k >= 2 ? -1. : 1.;
            for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) sP[k][4 * r + c] = Rk[3 * r + c]; sP[k][4 * r + 3] = s * U[3 * r + 2]; }
            sGood[k] = 0;
        }
    }
    __syncthreads();
    const double P0[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };
    int good[4] = { 0, 0, 0, 0 };
    for (int i = tid; i < n; i += blockDim.x) {
        const double2 a = q1[i], b = q2[i];
        const bool in = mask[i] != 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            double P[12], Q[4];
            for (int j = 0; j < 12; j++) P[j] = sP[k][j];
            dlt_point(P0, P, a.x, a.y, b.x, b.y, Q);
            bool f = Q[2] * Q[3] > 0;
            const double x = Q[0] / Q[3], y = Q[1] / Q[3], z = Q[2] / Q[3];
            f = f && z < 50.;
            const double z2 = P[8] * x + P[9] * y + P[10] * z + P[11];
            f = f && z2 > 0 && z2 < 50. && in;
            flags[4 * (size_t)i + k] = f;
            good[k] += f;
        }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int g = __reduce_add_sync(0xffffffffu, good[k]);
        if ((tid & 31) == 0 && g) atomicAdd(&sGood[k], g);
    }
    __syncthreads();
    if (tid == 0) {
        const int g0 = sGood[0], g1 = sGood[1], g2 = sGood[2], g3 = sGood[3];
        int win;
        if (g0 >= g1 && g0 >= g2 && g0 >= g3) win = 0;
        else if (g1 >= g0 && g1 >= g2 && g1 >= g3) win = 1;
        else if (g2 >= g0 && g2 >= g1 && g2 >= g3) win = 2;
        else win = 3;
        sWin = win;
        bs->which = win; bs->n_good = sGood[win];
        for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) bs->R[3 * r + c] = sP[win][4 * r + c]; bs->t[r] = sP[win][4 * r + 3]; }
    }
    __syncthreads();
    const int win = sWin;
    for (int i = tid; i < n; i += blockDim.x) mask[i] = flags[4 * (size_t)i + win];
}

// the index table cv2's RANSAC draws for n points, 5 per iteration (cv::RNG(2^64-1), duplicates re-drawn)
void host_sets5(int n, int H, std::vector<int32_t>& out)
{
    out.resize((size_t)H * 5);
    uint64_t s = 0xFFFFFFFFFFFFFFFFULL;
    for (int h = 0; h < H; h++) {
        int32_t* idx = &out[(size_t)h * 5];
        for (int i = 0; i < 5; i++) {
            for (;;) {
                s = (uint64_t)(unsigned)s * 4164903690ULL + (unsigned)(s >> 32);
                const int x = (int)((unsigned)s % (unsigned)n);
                bool dup = false;
                for (int j = 0; j < i; j++) dup |= idx[j] == x;
                if (!dup) { idx[i] = x; break; }
            }
        }
    }
}

}  // namespace

}  // namespace vo

using namespace vo;

extern "C" {

void vo_boot_params_default(vo_boot_params* p)
{
    if (!p) return;
    p->prob = 0.999; p->threshold = 1.0; p->max_iters = 1000;
}

int vo_bootstrap_pose(vo_ctx* c, const double* kps1, const double* kps2, int n, const double K[9], const vo_boot_params* params,
                      double T34[12], uint8_t* mask, double E[9], int32_t info[6])
{
    if (!c || !c->pipe) return VO_ERR_INVALID;
    if (n < 0 || !K || (n > 0 && (!kps1 || !kps2))) return VO_ERR_INVALID;
    VO_NVTX("vo:bootstrap(5-point RANSAC + recoverPose)");
    vo_boot_params prm;
    if (params) prm = *params; else vo_boot_params_default(&prm);
    if (!(prm.prob > 0. && prm.prob < 1.)) return set_error(c, VO_ERR_INVALID, "bootstrap: confidence must be in (0, 1)");
    const int H = n < 5 ? 0 : (n == 5 ? 1 : (prm.max_iters > 1 ? prm.max_iters : 1));
    Pipeline* p = c->pipe;
    BootScalars hs;
    memset(&hs, 0, sizeof(hs));
    if (H > 0) {
        const size_t nn = (size_t)n;
        // scratch: one double buffer (points, normalised points, models, scalars) and one int buffer (sets, counts)
        const size_t off_q = 4 * nn, off_models = off_q + 4 * nn, off_bs = off_models + (size_t)H * 9 * kMaxModels;
        VO_TRY(ensure(c, p->stage_d, off_bs + sizeof(BootScalars) / sizeof(double) + 1));
        const size_t off_nm = (size_t)H * 5, off_cnt = off_nm + H, off_flags = off_cnt + (size_t)H * kMaxModels;
        VO_TRY(ensure(c, p->stage_i, off_flags + (5 * nn + 3) / 4 + 1));
        double* d = p->stage_d.p;
        int32_t* di = p->stage_i.p;
        double2 *dp1 = (double2*)d, *dp2 = (double2*)(d + 2 * nn), *dq1 = (double2*)(d + off_q), *dq2 = (double2*)(d + off_q + 2 * nn);
        double* dmodels = d + off_models;
        BootScalars* dbs = (BootScalars*)(d + off_bs);
        uint8_t* dmask = (uint8_t*)(di + off_flags);
        uint8_t* dflags = dmask + nn;
        std::vector<int32_t> sets;
        if (n > 5) host_sets5(n, H, sets);
        cudaStream_t st = c->stream;
        VO_CUDA(c, cudaMemcpyAsync(dp1, kps1, sizeof(double) * 2 * nn, cudaMemcpyDefault, st));
        VO_CUDA(c, cudaMemcpyAsync(dp2, kps2, sizeof(double) * 2 * nn, cudaMemcpyDefault, st));
        if (n > 5) VO_CUDA(c, cudaMemcpyAsync(di, sets.data(), sizeof(int32_t) * sets.size(), cudaMemcpyHostToDevice, st));
        double thr = prm.threshold / ((K[0] + K[4]) / 2);
        const float t2 = (float)(thr * thr);
        em_normalise_kernel<<<(n + 255) / 256, 256, 0, st>>>(dp1, dp2, n, K[0], K[4], K[2], K[5], dq1, dq2);
        em_solve_kernel<<<(H + 31) / 32, 32, 0, st>>>(dq1, dq2, n, di, H, dmodels, di + off_nm);
        em_score_kernel<<<H, 128, 0, st>>>(dq1, dq2, n, dmodels, di + off_nm, t2, di + off_cnt);
        em_select_kernel<<<1, 256, 0, st>>>(dbs, dq1, dq2, n, dmodels, di + off_nm, di + off_cnt, H, prm.prob, t2, dmask);
        recover_pose_kernel<<<1, 256, 0, st>>>(dbs, dq1, dq2, n, dmask, dflags);
        c->launches += 5;
        VO_CUDA(c, cudaGetLastError());
        VO_CUDA(c, cudaMemcpyAsync(&hs, dbs, sizeof(hs), cudaMemcpyDeviceToHost, st));
        if (mask) VO_CUDA(c, cudaMemcpyAsync(mask, dmask, nn, cudaMemcpyDefault, st));
        VO_CUDA(c, cudaStreamSynchronize(st));   // (also: `sets` is a local)
    } else if (mask && n > 0) {
        std::vector<uint8_t> zeros((size_t)n, 0);
        VO_CUDA(c, cudaMemcpy(mask, zeros.data(), n, cudaMemcpyDefault));
    }
    if (!hs.ok) { memset(hs.R, 0, sizeof(hs.R)); memset(hs.t, 0, sizeof(hs.t)); memset(hs.E, 0, sizeof(hs.E)); hs.n_good = 0; }
    if (T34) for (int r = 0; r < 3; r++) { for (int k = 0; k < 3; k++) T34[4 * r + k] = hs.R[3 * r + k]; T34[4 * r + 3] = hs.t[r]; }
    if (E) memcpy(E, hs.E, sizeof(hs.E));
    if (info) { info[0] = hs.ok; info[1] = hs.iters_run; info[2] = hs.best_iter; info[3] = hs.best_model; info[4] = hs.n_inliers; info[5] = hs.n_good; }
    return VO_OK;
}

}  // extern "C"
