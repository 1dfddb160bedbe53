/*
 * ORACLE (test infrastructure only - never linked into or called by the product path).
 *
 * Scalar CPU restatement of
 *   (a) cv2.triangulatePoints(M1, M2, p1, p2)   /root/reference/src/vo_pipeline.py:666 (one
 *       point per call inside the promotion loop) and :172 (batched, bootstrap), and
 *   (b) the bearing-angle baseline test of vo_pipeline.py:621-642 that decides which candidate
 *       keypoints (C, F, T) are promoted to landmarks (P, X).
 * The DLT arithmetic lives in OpenCV (third-party, not vendored by the reference; pinned
 * opencv-python==4.8.1.78, 4.13.0.92 installed here); restated per SURVEY.md App. A.3: rows
 * x*M[2]-M[0], y*M[2]-M[1] for both views (no normalisation), right singular vector of the
 * smallest singular value, here by one-sided Jacobi (Hestenes) in float64.  The sign of the
 * homogeneous vector is arbitrary in OpenCV; the dehomogenised point (vo_pipeline.py:669) is not.
 * Pinned against cv2 4.13.0 outputs: tests/golden/tri_*.npz (tests/golden/make_golden.py).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

/* Xh: 4 doubles (unit norm); returns the number of Jacobi sweeps used */
int vo_oracle_triangulate_one(const double* M1, const double* M2, const double* p1, const double* p2, double* Xh)
{
    double A[4][4], V[4][4];
    for (int c = 0; c < 4; c++) {
        A[0][c] = p1[0] * M1[8 + c] - M1[c];
        A[1][c] = p1[1] * M1[8 + c] - M1[4 + c];
        A[2][c] = p2[0] * M2[8 + c] - M2[c];
        A[3][c] = p2[1] * M2[8 + c] - M2[4 + c];
    }
    memset(V, 0, sizeof(V));
    for (int i = 0; i < 4; i++) V[i][i] = 1.;
    int sweep;
    for (sweep = 0; sweep < 30; sweep++) {
        int rotated = 0;
        for (int p = 0; p < 3; p++)
            for (int q = p + 1; q < 4; q++) {
                double app = 0., aqq = 0., apq = 0.;
                for (int r = 0; r < 4; r++) { app += A[r][p] * A[r][p]; aqq += A[r][q] * A[r][q]; apq += A[r][p] * A[r][q]; }
                if (fabs(apq) <= 1e-17 * sqrt(app * aqq) || apq == 0.) continue;
                rotated = 1;
                const double zeta = (aqq - app) / (2. * apq);
                const double t = (zeta >= 0. ? 1. : -1.) / (fabs(zeta) + sqrt(1. + zeta * zeta));
                const double c = 1. / sqrt(1. + t * t), s = c * t;
                for (int r = 0; r < 4; r++) {
                    const double ap = A[r][p], aq = A[r][q];
                    A[r][p] = c * ap - s * aq; A[r][q] = s * ap + c * aq;
                    const double vp = V[r][p], vq = V[r][q];
                    V[r][p] = c * vp - s * vq; V[r][q] = s * vp + c * vq;
                }
            }
        if (!rotated) break;
    }
    int best = 0;
    double bn = -1.;
    for (int c = 0; c < 4; c++) {
        double nn = 0.;
        for (int r = 0; r < 4; r++) nn += A[r][c] * A[r][c];
        if (bn < 0. || nn < bn) { bn = nn; best = c; }
    }
    for (int r = 0; r < 4; r++) Xh[r] = V[r][best];
    return sweep;
}

/* batched: M1, M2 shared (3x4 row-major), p1/p2 n x 2, Xh n x 4 */
void vo_oracle_triangulate_points(const double* M1, const double* M2, const double* p1, const double* p2, int n, double* Xh)
{
    for (int i = 0; i < n; i++) vo_oracle_triangulate_one(M1, M2, p1 + 2 * i, p2 + 2 * i, Xh + 4 * i);
}

/* vo_pipeline.py:621-642, literally: bearing vectors from RAW pixel coordinates, normalised,
 * cos clipped, arccos, degrees. */
double vo_oracle_bearing_angle_deg(const double* c, const double* f)
{
    const double wc = sqrt(c[0] * c[0] + c[1] * c[1] + 1.), wf = sqrt(f[0] * f[0] + f[1] * f[1] + 1.);
    const double a[3] = { c[0] / wc, c[1] / wc, 1. / wc }, b[3] = { f[0] / wf, f[1] / wf, 1. / wf };
    const double na = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]), nb = sqrt(b[0] * b[0] + b[1] * b[1] + b[2] * b[2]);
    double cs = (a[0] * b[0] + a[1] * b[1] + a[2] * b[2]) / (na * nb);
    cs = cs < -1. ? -1. : (cs > 1. ? 1. : cs);
    return acos(cs) * (180. / 3.14159265358979323846);
}

/* The promotion step of LandmarkTriangulator.triangulate_landmark (vo_pipeline.py:608-721) for
 * candidates already tracked and appended: C,F n x 2, T n x 16 (4x4 world->camera of the first
 * observation), Tnew 4x4, K 3x3.  promote[i] = alpha_i > alpha_deg; Xout n x 3 (rows of promoted
 * candidates filled).  Returns the number promoted. */
int vo_oracle_bearing_triangulate(const double* C, const double* F, const double* T, int n, const double* Tnew,
                                  const double* K, double alpha_deg, uint8_t* promote, double* Xout, double* alpha_out)
{
    double M2[12];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 4; c++)
            M2[4 * r + c] = K[3 * r] * Tnew[c] + K[3 * r + 1] * Tnew[4 + c] + K[3 * r + 2] * Tnew[8 + c];
    int np = 0;
    for (int i = 0; i < n; i++) {
        const double al = vo_oracle_bearing_angle_deg(C + 2 * i, F + 2 * i);
        if (alpha_out) alpha_out[i] = al;
        promote[i] = al > alpha_deg;
        if (!promote[i]) continue;
        double M1[12], Xh[4];
        const double* Ti = T + 16 * (size_t)i;
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 4; c++)
                M1[4 * r + c] = K[3 * r] * Ti[c] + K[3 * r + 1] * Ti[4 + c] + K[3 * r + 2] * Ti[8 + c];
        vo_oracle_triangulate_one(M1, M2, F + 2 * i, C + 2 * i, Xh);
        for (int c = 0; c < 3; c++) Xout[3 * i + c] = Xh[c] / Xh[3];
        np++;
    }
    return np;
}
