// DLT triangulation of one point from two views (cv2.triangulatePoints behaviour), shared by the
// promotion step (tri.cu) and the bootstrap's cheirality test (boot.cu).
#pragma once

namespace vo {

// ---- DLT ---------------------------------------------------------------------------------------
// right singular vector of the smallest singular value of the 4x4 A (rows x*M[2]-M[0], ...),
// one-sided Jacobi (Hestenes).  M1, M2: 3x4 row-major.
__device__ inline void dlt_point(const double* M1, const double* M2, double x1, double y1, double x2, double y2, double* Xh)
{
    double A[4][4], V[4][4];
#pragma unroll
    for (int c = 0; c < 4; c++) {
        A[0][c] = x1 * M1[8 + c] - M1[c];
        A[1][c] = y1 * M1[8 + c] - M1[4 + c];
        A[2][c] = x2 * M2[8 + c] - M2[c];
        A[3][c] = y2 * M2[8 + c] - M2[4 + c];
#pragma unroll
        for (int r = 0; r < 4; r++) V[r][c] = r == c ? 1. : 0.;
    }
    // One rotation of columns (p,q).  The rotation angle costs one sqrt, one division and one rsqrt
    // on the dependent chain; the two rotations of a step touch disjoint columns, so their chains
    // overlap (round-robin ordering (0,1)(2,3) / (0,2)(1,3) / (0,3)(1,2)).  Converged when every
    // pair is orthogonal to 1e-15 relative (a tighter bound sits below the rounding noise of the
    // dot products and would spin until the sweep limit).
#define VO_JACOBI_ROT(p, q)                                                                         \
    {                                                                                               \
        double app = 0., aqq = 0., apq = 0.;                                                        \
        _Pragma("unroll") for (int r = 0; r < 4; r++) {                                             \
            app += A[r][p] * A[r][p]; aqq += A[r][q] * A[r][q]; apq += A[r][p] * A[r][q];           \
        }                                                                                           \
        if (apq * apq > 1e-30 * (app * aqq)) {                                                      \
            rotated = true;                                                                         \
            const double d = aqq - app;                                                             \
            const double t = (d >= 0. ? 2. : -2.) * apq / (fabs(d) + sqrt(d * d + 4. * apq * apq)); \
            const double cs = rsqrt(1. + t * t), sn = cs * t;                                       \
            _Pragma("unroll") for (int r = 0; r < 4; r++) {                                         \
                const double ap = A[r][p], aq = A[r][q];                                            \
                A[r][p] = cs * ap - sn * aq; A[r][q] = sn * ap + cs * aq;                           \
                const double vp = V[r][p], vq = V[r][q];                                            \
                V[r][p] = cs * vp - sn * vq; V[r][q] = sn * vp + cs * vq;                           \
            }                                                                                       \
        }                                                                                           \
    }
    for (int sweep = 0; sweep < 30; sweep++) {
        bool rotated = false;
        VO_JACOBI_ROT(0, 1) VO_JACOBI_ROT(2, 3)
        VO_JACOBI_ROT(0, 2) VO_JACOBI_ROT(1, 3)
        VO_JACOBI_ROT(0, 3) VO_JACOBI_ROT(1, 2)
        if (!rotated) break;
    }
#undef VO_JACOBI_ROT
    double bn = 0.;
    int best = 0;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        double nn = 0.;
#pragma unroll
        for (int r = 0; r < 4; r++) nn += A[r][c] * A[r][c];
        if (c == 0 || nn < bn) { bn = nn; best = c; }
    }
#pragma unroll
    for (int r = 0; r < 4; r++) Xh[r] = best == 0 ? V[r][0] : (best == 1 ? V[r][1] : (best == 2 ? V[r][2] : V[r][3]));
}


}  // namespace vo
