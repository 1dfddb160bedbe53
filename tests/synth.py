"""Seeded synthetic inputs for the parity tests.

Everything here is integer / numpy-only so the same seeds give bit-identical inputs in the build
container (where the golden vectors were made with cv2 + the reference classes) and on the GPU
box.  No cv2, no oracle and no product imports.
"""
import numpy as np


def _box_blur_int(a, r):
    """Exact integer box blur (radius r) with edge replication, via cumulative sums."""
    for axis in (0, 1):
        pad = [(0, 0), (0, 0)]
        pad[axis] = (r + 1, r)
        p = np.pad(a, pad, mode="edge")
        c = np.cumsum(p, axis=axis, dtype=np.int64)
        n = a.shape[axis]
        hi = np.take(c, np.arange(2 * r + 1, 2 * r + 1 + n), axis=axis)
        lo = np.take(c, np.arange(0, n), axis=axis)
        a = (hi - lo) // (2 * r + 1)
    return a


def texture(h, w, seed, radius=2, passes=3):
    """Band-limited noise texture, u8: uniform noise, `passes` box blurs (~Gaussian sigma 2.4),
    contrast stretched to 0..255.  Non-periodic."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, size=(h, w), dtype=np.int64) * 256
    for _ in range(passes):
        a = _box_blur_int(a, radius)
    lo, hi = int(a.min()), int(a.max())
    return ((a - lo) * 255 // max(hi - lo, 1)).astype(np.uint8)


def warp_affine_fixed(src, M, out_h, out_w):
    """dst(x,y) = bilinear(src, M @ (x,y,1)) with 10-bit fixed-point weights; exact integers."""
    ys, xs = np.mgrid[0:out_h, 0:out_w]
    fx = np.round((M[0, 0] * xs + M[0, 1] * ys + M[0, 2]) * 1024).astype(np.int64)
    fy = np.round((M[1, 0] * xs + M[1, 1] * ys + M[1, 2]) * 1024).astype(np.int64)
    x0 = np.clip(fx >> 10, 0, src.shape[1] - 2)
    y0 = np.clip(fy >> 10, 0, src.shape[0] - 2)
    ax = np.clip(fx - (x0 << 10), 0, 1024)
    ay = np.clip(fy - (y0 << 10), 0, 1024)
    s = src.astype(np.int64)
    v = (s[y0, x0] * (1024 - ax) * (1024 - ay) + s[y0, x0 + 1] * ax * (1024 - ay)
         + s[y0 + 1, x0] * (1024 - ax) * ay + s[y0 + 1, x0 + 1] * ax * ay + (1 << 19)) >> 20
    return v.astype(np.uint8)


def klt_pair(h, w, seed, n, shift=(-3.3, 1.7), affine=True, flat_patch=True, margin=5.0):
    """Image pair with a known sub-pixel motion and n points (some up to `margin` px outside the
    frame).  Returns prev, next, pts(float32 n x 2), true displacement of the centre."""
    big = texture(h + 64, w + 64, seed)
    prev = np.ascontiguousarray(big[32:32 + h, 32:32 + w])
    if affine:
        M = np.array([[1.002, 0.001, 32 + shift[0]], [-0.001, 0.998, 32 + shift[1]]])
    else:
        M = np.array([[1.0, 0.0, 32 + shift[0]], [0.0, 1.0, 32 + shift[1]]])
    nxt = warp_affine_fixed(big, M, h, w)
    if flat_patch and h > 80 and w > 100:
        prev = prev.copy()
        nxt = nxt.copy()
        prev[10:60, 10:80] = 128   # flat patch: exercises the minEig rule
        nxt[10:60, 10:80] = 128
        band = slice(h - 40, h - 20)
        prev[band] = (prev[band] // 32) * 4 + 100   # quantised low-contrast band
        nxt[band] = (nxt[band] // 32) * 4 + 100
    rng = np.random.default_rng(seed + 1000)
    pts = np.stack([rng.uniform(-margin, w + margin, n), rng.uniform(-margin, h + margin, n)], 1)
    return prev, nxt, pts.astype(np.float32)


# ---- PnP scenes ------------------------------------------------------------------
K_KITTI = np.array([[718.856, 0, 607.1928], [0, 718.856, 185.2157], [0, 0, 1.0]])
K_PARKING = np.array([[331.37, 0, 320], [0, 369.568, 240], [0, 0, 1.0]])
K_MALAGA = np.array([[621.18428, 0, 404.0076], [0, 621.18428, 309.05989], [0, 0, 1.0]])


def rodrigues(rvec):
    rvec = np.asarray(rvec, np.float64).reshape(3)
    th = np.linalg.norm(rvec)
    if th < 1e-300:
        return np.eye(3)
    k = rvec / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.cos(th) * np.eye(3) + (1 - np.cos(th)) * np.outer(k, k) + np.sin(th) * Kx


def pnp_scene(n, outlier_frac, seed, K=K_KITTI, rvec=(0.01, -0.03, 0.005), tvec=(0.2, -0.05, 0.8),
              noise=0.3, w=1241, h=376):
    """X ~ U([-15,15]x[-3,3]x[5,50]) m, projected with (rvec,tvec), pixel noise N(0,noise),
    the first `outlier_frac` share (after shuffling) replaced by uniform random pixels."""
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(-15, 15, n), rng.uniform(-3, 3, n), rng.uniform(5, 50, n)], 1)
    R = rodrigues(rvec)
    Xc = X @ R.T + np.asarray(tvec)
    uv = Xc[:, :2] / Xc[:, 2:3]
    P = np.stack([K[0, 0] * uv[:, 0] + K[0, 2], K[1, 1] * uv[:, 1] + K[1, 2]], 1)
    P += rng.normal(0, noise, P.shape)
    n_out = int(round(n * outlier_frac))
    idx = rng.permutation(n)[:n_out]
    P[idx] = np.stack([rng.uniform(0, w, n_out), rng.uniform(0, h, n_out)], 1)
    is_outlier = np.zeros(n, bool)
    is_outlier[idx] = True
    return X, P.astype(np.float32), R, np.asarray(tvec, np.float64), is_outlier


def two_view_scene(n, seed, K=K_KITTI, noise=0.3, baseline=1.0):
    """Two poses ~`baseline` m apart observing X; returns X, T1(4x4), T2(4x4), p1, p2 (float64)."""
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(-15, 15, n), rng.uniform(-3, 3, n), rng.uniform(5, 50, n)], 1)
    T1 = np.eye(4)
    T1[:3, :3] = rodrigues((0.002, 0.01, -0.001))
    T1[:3, 3] = (0.05, 0.0, 0.1)
    T2 = np.eye(4)
    T2[:3, :3] = rodrigues((0.01, -0.03, 0.005))
    T2[:3, 3] = (0.05 - baseline * 0.6, -0.02, 0.1 - baseline * 0.8)

    def proj(T):
        Xc = X @ T[:3, :3].T + T[:3, 3]
        uv = Xc[:, :2] / Xc[:, 2:3]
        return np.stack([K[0, 0] * uv[:, 0] + K[0, 2], K[1, 1] * uv[:, 1] + K[1, 2]], 1)

    p1 = proj(T1) + rng.normal(0, noise, (n, 2))
    p2 = proj(T2) + rng.normal(0, noise, (n, 2))
    return X, T1, T2, p1, p2


def bearing_pairs(n, seed, w=1241, h=376, sigma=15.0):
    """Candidate current/first observations (float32-representable, as KLT output / detector
    output are) whose bearing angle straddles the reference's 1-1.5 degree thresholds."""
    rng = np.random.default_rng(seed)
    C = np.stack([rng.uniform(0, w, n), rng.uniform(0, h, n)], 1).astype(np.float32).astype(np.float64)
    F = (C + rng.normal(0, sigma, C.shape)).astype(np.float32).astype(np.float64)
    return C, F


# ---- two-view bootstrap scenes: matched pixel pairs with outliers ---------------------------------------
def bootstrap_scene(n, outlier_frac, seed, K=K_KITTI, w=1241, h=376, noise=0.3, t=(0.9, -0.05, 0.3), rvec=(0.01, -0.03, 0.005)):
    """n matches between two views ~1 m apart; outliers are uniform random pixels in the second view.
    Returns p1, p2 (n x 2 float64), R, t (the true relative pose, |t| not normalised), outlier flags."""
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(-15, 15, n), rng.uniform(-3, 3, n), rng.uniform(5, 50, n)], 1)
    R = rodrigues(rvec)
    t = np.asarray(t, np.float64)

    def proj(Xc):
        return np.stack([K[0, 0] * Xc[:, 0] / Xc[:, 2] + K[0, 2], K[1, 1] * Xc[:, 1] / Xc[:, 2] + K[1, 2]], 1)
    p1 = proj(X) + rng.normal(0, noise, (n, 2))
    p2 = proj(X @ R.T + t) + rng.normal(0, noise, (n, 2))
    no = int(outlier_frac * n)
    idx = rng.choice(n, no, replace=False)
    p2[idx] = np.stack([rng.uniform(0, w, no), rng.uniform(0, h, no)], 1)
    is_out = np.zeros(n, bool)
    is_out[idx] = True
    return p1, p2, R, t, is_out
