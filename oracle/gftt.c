/*
 * ORACLE (test infrastructure only - never linked into or called by the product path).
 *
 * Scalar CPU restatement of
 *   cv2.goodFeaturesToTrack(new_frame, maxCorners=500, qualityLevel=0.03, minDistance=10, mask=mask,
 *                           blockSize=4, gradientSize=3, useHarrisDetector=False)
 * as the reference calls it in LandmarkTriangulator.find_new_candidates_shi
 * (/root/reference/src/vo_pipeline.py:357-390; SURVEY.md 8f-1, the first "next" row).
 * The arithmetic lives in OpenCV (third party, not vendored by the reference; pinned
 * opencv-python==4.8.1.78, installed here 4.13.0.92, x86-64 AVX2 dispatch).  Behaviour restated
 * from black-box probes of that build (cv2.Sobel, cv2.boxFilter, cv2.cornerMinEigenVal and
 * cv2.goodFeaturesToTrack are reproduced bit for bit, tests/test_gftt_oracle.py):
 *   1. Sobel 3x3 of the u8 image to float32, REFLECT_101, scale s = 1/(4*blockSize*255); the
 *      smoothing taps carry the scale (k1 = (float)s, k0 = (float)2s):
 *        Dx: rows first (exact integer differences S), then columns fused:
 *            Dx = fmaf(S[y-1]+S[y+1], k1, S[y]*k0)
 *        Dy: rows first, R = a*k1 + b*k0 + c*k1 evaluated as fmaf(c,k1,fmaf(b,k0,a*k1)) in the
 *            vectorised columns x < (W & ~31) and as ((a*k1 + b*k0) + c*k1), unfused, in the scalar
 *            tail; then Dy = R[y+1] - R[y-1].
 *   2. cov = (Dx*Dx, Dx*Dy, Dy*Dy) in float32; unnormalised blockSize x blockSize box sum with the
 *      anchor at blockSize/2 (rows y-2..y+1, cols x-2..x+1 for 4), REFLECT_101 of the cov image,
 *      accumulated in float64 and rounded to float32.
 *   3. a = 0.5f*sxx, b = sxy, c = 0.5f*syy;  eig = (a + c) - sqrtf((a - c)*(a - c) + b*b), unfused.
 *   4. maxVal over the masked pixels; keep eig > (float)(maxVal*qualityLevel), else 0; a pixel of
 *      the interior (1..W-2, 1..H-2) is a corner candidate if it is non-zero, unmasked and equal
 *      to the 3x3 maximum of the thresholded map.
 *   5. candidates sorted by value descending, ties by address descending; greedy acceptance:
 *      a candidate is taken unless an accepted one lies closer than minDistance (dx*dx+dy*dy <
 *      minDistance^2, float32); stop at maxCorners.  Output (x, y) as float32 integers.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static int reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) {
        if (p < 0) p = -p;
        else p = 2 * n - 2 - p;
    }
    return p;
}

/* Dx, Dy: w*h float32 each.  vec_cols < 0 selects the probed default W & ~31. */
void vo_oracle_sobel_scaled(const uint8_t* img, int w, int h, int block, int vec_cols, float* Dx, float* Dy)
{
    const double scale = 1.0 / ((double)(1 << 2) * block) / 255.0;
    const float k1 = (float)scale, k0 = (float)(2.0 * scale);
    if (vec_cols < 0) vec_cols = w & ~31;
    float* S = (float*)malloc(sizeof(float) * (size_t)w * 3);   /* row differences of rows y-1, y, y+1 */
    float* R = (float*)malloc(sizeof(float) * (size_t)w * 3);   /* row smoothings */
    for (int y = 0; y < h; y++) {
        for (int k = 0; k < 3; k++) {
            const uint8_t* row = img + (size_t)reflect101(y - 1 + k, h) * w;
            for (int x = 0; x < w; x++) {
                const float a = (float)row[reflect101(x - 1, w)], b = (float)row[x], c = (float)row[reflect101(x + 1, w)];
                S[k * w + x] = c - a;
                if (x < vec_cols) R[k * w + x] = fmaf(c, k1, fmaf(b, k0, a * k1));
                else {
                    volatile float t = a * k1;
                    volatile float u = b * k0;
                    volatile float v = t + u;
                    volatile float q = c * k1;
                    R[k * w + x] = v + q;
                }
            }
        }
        for (int x = 0; x < w; x++) {
            volatile float mid = S[w + x] * k0;
            Dx[(size_t)y * w + x] = fmaf(S[x] + S[2 * w + x], k1, mid);
            Dy[(size_t)y * w + x] = R[2 * w + x] - R[x];
        }
    }
    free(S); free(R);
}

void vo_oracle_corner_min_eigen_val(const uint8_t* img, int w, int h, int block, int vec_cols, float* eig)
{
    float* Dx = (float*)malloc(sizeof(float) * (size_t)w * h);
    float* Dy = (float*)malloc(sizeof(float) * (size_t)w * h);
    vo_oracle_sobel_scaled(img, w, h, block, vec_cols, Dx, Dy);
    const int an = block / 2;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            double sxx = 0., sxy = 0., syy = 0.;
            for (int j = 0; j < block; j++) {
                const int yy = reflect101(y - an + j, h);
                double rxx = 0., rxy = 0., ryy = 0.;
                for (int i = 0; i < block; i++) {
                    const int xx = reflect101(x - an + i, w);
                    const float dx = Dx[(size_t)yy * w + xx], dy = Dy[(size_t)yy * w + xx];
                    volatile float pxx = dx * dx, pxy = dx * dy, pyy = dy * dy;
                    rxx += pxx; rxy += pxy; ryy += pyy;
                }
                sxx += rxx; sxy += rxy; syy += ryy;
            }
            const float a = (float)sxx * 0.5f, b = (float)sxy, c = (float)syy * 0.5f;
            volatile float t = a - c;
            volatile float tt = t * t;
            volatile float bb = b * b;
            volatile float s = tt + bb;
            volatile float ac = a + c;
            eig[(size_t)y * w + x] = ac - sqrtf(s);
        }
    free(Dx); free(Dy);
}

typedef struct { float v; int idx; } Cand;

static int cand_cmp(const void* pa, const void* pb)
{
    const Cand* a = (const Cand*)pa; const Cand* b = (const Cand*)pb;
    if (a->v > b->v) return -1;
    if (a->v < b->v) return 1;
    return a->idx > b->idx ? -1 : (a->idx < b->idx ? 1 : 0);
}

/* corners: cap x 2 float32 (x, y).  mask: w*h u8 or NULL.  eig_out: optional w*h float32 (the
 * unthresholded min-eigenvalue map).  Returns the number of corners. */
int vo_oracle_good_features_to_track(const uint8_t* img, int w, int h, const uint8_t* mask, int max_corners,
                                     double quality, double min_distance, int block, int vec_cols,
                                     float* corners, int cap, float* eig_out, int* n_candidates)
{
    float* eig = (float*)malloc(sizeof(float) * (size_t)w * h);
    vo_oracle_corner_min_eigen_val(img, w, h, block, vec_cols, eig);
    if (eig_out) memcpy(eig_out, eig, sizeof(float) * (size_t)w * h);
    int have = 0;
    float maxv = 0.f;
    for (size_t i = 0; i < (size_t)w * h; i++)
        if (!mask || mask[i]) { if (!have || eig[i] > maxv) { maxv = eig[i]; have = 1; } }
    const float thr = (float)((double)maxv * quality);
    for (size_t i = 0; i < (size_t)w * h; i++) if (!(eig[i] > thr)) eig[i] = 0.f;
    Cand* cand = (Cand*)malloc(sizeof(Cand) * (size_t)w * h);
    int nc = 0;
    for (int y = 1; y < h - 1; y++)
        for (int x = 1; x < w - 1; x++) {
            const float v = eig[(size_t)y * w + x];
            if (v == 0.f || (mask && !mask[(size_t)y * w + x])) continue;
            float m = v;
            for (int j = -1; j <= 1; j++)
                for (int i = -1; i <= 1; i++) {
                    const float u = eig[(size_t)(y + j) * w + x + i];
                    if (u > m) m = u;
                }
            if (v == m) { cand[nc].v = v; cand[nc].idx = y * w + x; nc++; }
        }
    if (n_candidates) *n_candidates = nc;
    qsort(cand, nc, sizeof(Cand), cand_cmp);
    int n = 0;
    const float md2 = (float)(min_distance * min_distance);
    for (int k = 0; k < nc && n < cap; k++) {
        const int y = cand[k].idx / w, x = cand[k].idx - y * w;
        int good = 1;
        if (min_distance >= 1.) {
            /* cv2 keeps the accepted corners in a grid of cvRound(minDistance) cells and only looks
             * into the 3x3 cells around the candidate (identical to a full scan for integral distances) */
            const int cell = (int)lrint(min_distance);
            for (int j = 0; j < n; j++) {
                const int xj = (int)corners[2 * j], yj = (int)corners[2 * j + 1];
                if (abs(xj / cell - x / cell) > 1 || abs(yj / cell - y / cell) > 1) continue;
                const float dx = (float)x - corners[2 * j], dy = (float)y - corners[2 * j + 1];
                if (dx * dx + dy * dy < md2) { good = 0; break; }
            }
        }
        if (!good) continue;
        corners[2 * n] = (float)x; corners[2 * n + 1] = (float)y;
        n++;
        if (max_corners > 0 && n == max_corners) break;
    }
    free(cand); free(eig);
    return n;
}
