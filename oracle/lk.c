/*
 * ORACLE (test infrastructure only - never linked into or called by the product path).
 *
 * Scalar CPU restatement of the pyramidal Lucas-Kanade tracker the reference reaches
 * through cv2.calcOpticalFlowPyrLK(prev, next, pts, None) with all defaults:
 *   /root/reference/src/vo_pipeline.py:215  (landmark tracks, state['P'])
 *   /root/reference/src/vo_pipeline.py:501  (candidate tracks, extended_state['C'])
 * The arithmetic lives in OpenCV (third-party, not vendored in the reference; pinned
 * opencv-python==4.8.1.78 in requirements.txt:2, 4.13.0.92 installed here).  This file
 * restates the published algorithm (Bouguet's pyramidal LK as implemented by OpenCV's
 * lkpyramid/pyramids modules; behavioural spec in SURVEY.md Appendix A.1):
 *   - 5-tap [1 4 6 4 1]/16 integer pyrDown, (s+128)>>8, REFLECT_101
 *   - Scharr derivatives (gain 32) on the image, ZERO outside it, intensities REFLECT_101
 *   - 14-bit fixed point bilinear weights, int16 window with 5 fractional bits
 *   - float32 accumulators and 2x2 solve, the two break rules, level-0-only status writes
 * Pinned against cv2 4.13.0 outputs: tests/golden/klt_*.npz (made by tests/golden/make_golden.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

static int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

/* cv2.pyrDown restated: dst is ((w+1)/2) x ((h+1)/2), tightly packed. */
void vo_oracle_pyr_down(const uint8_t* src, int w, int h, uint8_t* dst)
{
    const int dw = (w + 1) / 2, dh = (h + 1) / 2;
    int* rows = (int*)malloc(sizeof(int) * (size_t)dw * 5);
    for (int y = 0; y < dh; y++) {
        for (int k = 0; k < 5; k++) {
            const uint8_t* s = src + (size_t)reflect101(2 * y + k - 2, h) * w;
            int* r = rows + (size_t)k * dw;
            for (int x = 0; x < dw; x++) {
                int c = 2 * x;
                r[x] = s[reflect101(c - 2, w)] + 4 * s[reflect101(c - 1, w)] + 6 * s[c] +
                       4 * s[reflect101(c + 1, w)] + s[reflect101(c + 2, w)];
            }
        }
        for (int x = 0; x < dw; x++) {
            int v = rows[x] + 4 * rows[dw + x] + 6 * rows[2 * dw + x] + 4 * rows[3 * dw + x] + rows[4 * dw + x];
            dst[(size_t)y * dw + x] = (uint8_t)((v + 128) >> 8);
        }
    }
    free(rows);
}

typedef struct {
    int w, h, pad, stride;
    uint8_t* img;   /* padded, REFLECT_101 */
    int16_t* deriv; /* padded, 2 channels (Ix, Iy), zero outside the image */
} Level;

static void level_build(Level* L, const uint8_t* src, int w, int h, int pad, int want_deriv)
{
    L->w = w; L->h = h; L->pad = pad; L->stride = w + 2 * pad;
    const int H = h + 2 * pad, S = L->stride;
    L->img = (uint8_t*)malloc((size_t)S * H);
    for (int y = -pad; y < h + pad; y++) {
        const uint8_t* s = src + (size_t)reflect101(y, h) * w;
        uint8_t* d = L->img + (size_t)(y + pad) * S;
        for (int x = -pad; x < w + pad; x++) d[x + pad] = s[reflect101(x, w)];
    }
    L->deriv = NULL;
    if (!want_deriv) return;
    L->deriv = (int16_t*)calloc((size_t)S * H * 2, sizeof(int16_t));
    for (int y = 0; y < h; y++) {
        const uint8_t* r0 = src + (size_t)reflect101(y - 1, h) * w;
        const uint8_t* r1 = src + (size_t)y * w;
        const uint8_t* r2 = src + (size_t)reflect101(y + 1, h) * w;
        int16_t* d = L->deriv + ((size_t)(y + pad) * S + pad) * 2;
        for (int x = 0; x < w; x++) {
            int xm = reflect101(x - 1, w), xp = reflect101(x + 1, w);
            int t0m = 3 * (r0[xm] + r2[xm]) + 10 * r1[xm];
            int t0p = 3 * (r0[xp] + r2[xp]) + 10 * r1[xp];
            int t1m = r2[xm] - r0[xm], t1c = r2[x] - r0[x], t1p = r2[xp] - r0[xp];
            d[2 * x] = (int16_t)(t0p - t0m);
            d[2 * x + 1] = (int16_t)(3 * (t1m + t1p) + 10 * t1c);
        }
    }
}

static void level_free(Level* L) { free(L->img); free(L->deriv); }

static inline int cv_round(float v) { return (int)lrintf(v); } /* round-half-even */
static inline int cv_floor(float v) { return (int)floorf(v); }
#define DESCALE(x, n) (((x) + (1 << ((n)-1))) >> (n))

/*
 * Returns the number of pyramid levels actually used minus one (the effective maxLevel).
 * iters_out (optional, n ints): total Gauss-Newton iterations executed per track.
 */
int vo_oracle_calc_optical_flow_pyr_lk(const uint8_t* prev, const uint8_t* next, int w, int h,
                                       const float* pts, int n, float* next_pts, uint8_t* status,
                                       float* err, int win, int max_level, int max_iters, double eps,
                                       double min_eig_thr, int* iters_out)
{
    enum { MAXL = 16 };
    if (max_level >= MAXL) max_level = MAXL - 1;
    Level LI[MAXL], LJ[MAXL];
    uint8_t* own_prev[MAXL] = {0};
    uint8_t* own_next[MAXL] = {0};
    const uint8_t* cp = prev; const uint8_t* cn = next;
    int cw = w, ch = h, nlev = 0;
    for (int l = 0; l <= max_level; l++) {
        level_build(&LI[l], cp, cw, ch, win, 1);
        level_build(&LJ[l], cn, cw, ch, win, 0);
        nlev = l + 1;
        int nw = (cw + 1) / 2, nh = (ch + 1) / 2;
        if (l == max_level || nw <= win || nh <= win) break;
        own_prev[l + 1] = (uint8_t*)malloc((size_t)nw * nh);
        own_next[l + 1] = (uint8_t*)malloc((size_t)nw * nh);
        vo_oracle_pyr_down(cp, cw, ch, own_prev[l + 1]);
        vo_oracle_pyr_down(cn, cw, ch, own_next[l + 1]);
        cp = own_prev[l + 1]; cn = own_next[l + 1]; cw = nw; ch = nh;
    }
    const int top = nlev - 1;

    if (max_iters < 0) max_iters = 0;
    if (max_iters > 100) max_iters = 100;
    if (eps < 0) eps = 0;
    if (eps > 10) eps = 10;
    const double eps2 = eps * eps;
    const float half = (win - 1) * 0.5f;
    const float FLT_SCALE = 1.f / (1 << 20);
    int16_t* Iw = (int16_t*)malloc(sizeof(int16_t) * win * win * 3);
    int16_t* dIw = Iw + win * win;

    for (int i = 0; i < n; i++) { status[i] = 1; if (err) err[i] = 0; if (iters_out) iters_out[i] = 0; }

    for (int level = top; level >= 0; level--) {
        const Level* I = &LI[level]; const Level* J = &LJ[level];
        const int S = I->stride, pad = I->pad;
        for (int p = 0; p < n; p++) {
            float px = pts[2 * p] * (float)(1. / (1 << level));
            float py = pts[2 * p + 1] * (float)(1. / (1 << level));
            float nx, ny;
            if (level == top) { nx = px; ny = py; }
            else { nx = next_pts[2 * p] * 2.f; ny = next_pts[2 * p + 1] * 2.f; }
            next_pts[2 * p] = nx; next_pts[2 * p + 1] = ny;

            px -= half; py -= half;
            int ix = cv_floor(px), iy = cv_floor(py);
            if (ix < -win || ix >= I->w || iy < -win || iy >= I->h) {
                if (level == 0) { status[p] = 0; if (err) err[p] = 0; }
                continue;
            }
            float a = px - ix, b = py - iy;
            int iw00 = cv_round((1.f - a) * (1.f - b) * (1 << 14));
            int iw01 = cv_round(a * (1.f - b) * (1 << 14));
            int iw10 = cv_round((1.f - a) * b * (1 << 14));
            int iw11 = (1 << 14) - iw00 - iw01 - iw10;

            float iA11 = 0, iA12 = 0, iA22 = 0;
            for (int y = 0; y < win; y++) {
                const uint8_t* src = I->img + (size_t)(y + iy + pad) * S + ix + pad;
                const int16_t* ds = I->deriv + ((size_t)(y + iy + pad) * S + ix + pad) * 2;
                for (int x = 0; x < win; x++, ds += 2) {
                    int ival = DESCALE(src[x] * iw00 + src[x + 1] * iw01 + src[x + S] * iw10 + src[x + S + 1] * iw11, 14 - 5);
                    int ixval = DESCALE(ds[0] * iw00 + ds[2] * iw01 + ds[2 * S] * iw10 + ds[2 * S + 2] * iw11, 14);
                    int iyval = DESCALE(ds[1] * iw00 + ds[3] * iw01 + ds[2 * S + 1] * iw10 + ds[2 * S + 3] * iw11, 14);
                    Iw[y * win + x] = (int16_t)ival;
                    dIw[2 * (y * win + x)] = (int16_t)ixval;
                    dIw[2 * (y * win + x) + 1] = (int16_t)iyval;
                    iA11 += (float)(ixval * ixval);
                    iA12 += (float)(ixval * iyval);
                    iA22 += (float)(iyval * iyval);
                }
            }
            float A11 = iA11 * FLT_SCALE, A12 = iA12 * FLT_SCALE, A22 = iA22 * FLT_SCALE;
            float D = A11 * A22 - A12 * A12;
            float minEig = (A22 + A11 - sqrtf((A11 - A22) * (A11 - A22) + 4.f * A12 * A12)) / (2 * win * win);
            if (minEig < min_eig_thr || D < FLT_EPSILON) {
                if (level == 0) status[p] = 0;
                continue;
            }
            D = 1.f / D;
            nx -= half; ny -= half;
            float pdx = 0, pdy = 0;
            for (int j = 0; j < max_iters; j++) {
                int jx = cv_floor(nx), jy = cv_floor(ny);
                if (jx < -win || jx >= J->w || jy < -win || jy >= J->h) {
                    if (level == 0) status[p] = 0;
                    break;
                }
                if (iters_out) iters_out[p]++;
                a = nx - jx; b = ny - jy;
                iw00 = cv_round((1.f - a) * (1.f - b) * (1 << 14));
                iw01 = cv_round(a * (1.f - b) * (1 << 14));
                iw10 = cv_round((1.f - a) * b * (1 << 14));
                iw11 = (1 << 14) - iw00 - iw01 - iw10;
                float ib1 = 0, ib2 = 0;
                for (int y = 0; y < win; y++) {
                    const uint8_t* Jp = J->img + (size_t)(y + jy + pad) * S + jx + pad;
                    for (int x = 0; x < win; x++) {
                        int diff = DESCALE(Jp[x] * iw00 + Jp[x + 1] * iw01 + Jp[x + S] * iw10 + Jp[x + S + 1] * iw11, 14 - 5) - Iw[y * win + x];
                        ib1 += (float)(diff * dIw[2 * (y * win + x)]);
                        ib2 += (float)(diff * dIw[2 * (y * win + x) + 1]);
                    }
                }
                float b1 = ib1 * FLT_SCALE, b2 = ib2 * FLT_SCALE;
                float dx = (float)((A12 * b2 - A22 * b1) * D);
                float dy = (float)((A12 * b1 - A11 * b2) * D);
                nx += dx; ny += dy;
                next_pts[2 * p] = nx + half; next_pts[2 * p + 1] = ny + half;
                if ((double)dx * dx + (double)dy * dy <= eps2) break;
                if (j > 0 && fabsf(dx + pdx) < 0.01 && fabsf(dy + pdy) < 0.01) {
                    next_pts[2 * p] -= dx * 0.5f; next_pts[2 * p + 1] -= dy * 0.5f;
                    break;
                }
                pdx = dx; pdy = dy;
            }
            if (status[p] && err && level == 0) {
                float qx = next_pts[2 * p] - half, qy = next_pts[2 * p + 1] - half;
                int jx = cv_floor(qx), jy = cv_floor(qy);
                if (jx < -win || jx >= J->w || jy < -win || jy >= J->h) { status[p] = 0; continue; }
                a = qx - jx; b = qy - jy;
                iw00 = cv_round((1.f - a) * (1.f - b) * (1 << 14));
                iw01 = cv_round(a * (1.f - b) * (1 << 14));
                iw10 = cv_round((1.f - a) * b * (1 << 14));
                iw11 = (1 << 14) - iw00 - iw01 - iw10;
                float e = 0.f;
                for (int y = 0; y < win; y++) {
                    const uint8_t* Jp = J->img + (size_t)(y + jy + pad) * S + jx + pad;
                    for (int x = 0; x < win; x++) {
                        int diff = DESCALE(Jp[x] * iw00 + Jp[x + 1] * iw01 + Jp[x + S] * iw10 + Jp[x + S + 1] * iw11, 14 - 5) - Iw[y * win + x];
                        e += fabsf((float)diff);
                    }
                }
                err[p] = e * 1.f / (32 * win * win);
            }
        }
    }
    free(Iw);
    for (int l = 0; l < nlev; l++) { level_free(&LI[l]); level_free(&LJ[l]); free(own_prev[l]); free(own_next[l]); }
    return top;
}

/* Scharr (Ix,Iy) plane of an image, int16 interleaved, REFLECT_101 at the image edge. */
void vo_oracle_scharr(const uint8_t* src, int w, int h, int16_t* out)
{
    Level L;
    level_build(&L, src, w, h, 1, 1);
    for (int y = 0; y < h; y++)
        memcpy(out + (size_t)y * w * 2, L.deriv + ((size_t)(y + 1) * L.stride + 1) * 2, sizeof(int16_t) * 2 * w);
    level_free(&L);
}
