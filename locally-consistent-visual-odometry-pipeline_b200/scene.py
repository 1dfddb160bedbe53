"""Synthetic sequences with a known ground-truth trajectory.

The datasets the reference runs on (KITTI 05, Malaga 07, parking; main.py:14-68) are not available
offline, so throughput and trajectory parity are measured on ray-cast scenes of the same shape: a
corridor of textured planes (ground, two walls, a far wall) seen by a pinhole camera that drives
forward with a gentle yaw.  Per pixel: ray-plane intersection, nearest hit, bilinear texture fetch -
plain numpy, no graphics API.  Textures are multi-octave band-limited noise and never tile inside the
field of view (tiling made SIFT alias in the survey's probe).  Everything is seeded and integer /
float64 deterministic, so the build container and the GPU box render identical frames.

Conventions: camera x right, y down, z forward; T is the 4x4 world->camera matrix the reference
uses (vo_pipeline.py:333).
"""
import numpy as np

K_KITTI = np.array([[718.856, 0, 607.1928], [0, 718.856, 185.2157], [0, 0, 1.0]])      # main.py:23-25
K_MALAGA = np.array([[621.18428, 0, 404.0076], [0, 621.18428, 309.05989], [0, 0, 1.0]])  # main.py:41-43
K_PARKING = np.array([[331.37, 0, 320], [0, 369.568, 240], [0, 0, 1.0]])                # main.py:55-57

SHAPES = {  # name: (width, height, K)
    "kitti": (1241, 376, K_KITTI),
    "malaga": (1024, 768, K_MALAGA),
    "parking": (640, 480, K_PARKING),
}


def _box_blur(a, r):
    for axis in (0, 1):
        pad = [(0, 0), (0, 0)]
        pad[axis] = (r + 1, r)
        c = np.cumsum(np.pad(a, pad, mode="edge"), axis=axis, dtype=np.float64)
        n = a.shape[axis]
        a = (np.take(c, np.arange(2 * r + 1, 2 * r + 1 + n), axis=axis) - np.take(c, np.arange(0, n), axis=axis)) / (2 * r + 1)
    return a


def make_texture(h, w, seed):
    """Multi-octave band-limited noise in [0, 255] (float32)."""
    rng = np.random.default_rng(seed)
    out = np.zeros((h, w))
    for octave, (scale, gain) in enumerate([(1, 0.45), (4, 0.3), (16, 0.25)]):
        hh, ww = -(-h // scale) + 1, -(-w // scale) + 1
        a = rng.random((hh, ww))
        for _ in range(2):
            a = _box_blur(a, 1)
        a = (a - a.min()) / max(a.max() - a.min(), 1e-12)
        if scale > 1:   # bilinear upsample
            ys = np.arange(h) / scale
            xs = np.arange(w) / scale
            y0 = np.floor(ys).astype(int); x0 = np.floor(xs).astype(int)
            fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
            a = (a[y0][:, x0] * (1 - fy) * (1 - fx) + a[y0][:, x0 + 1] * (1 - fy) * fx
                 + a[y0 + 1][:, x0] * fy * (1 - fx) + a[y0 + 1][:, x0 + 1] * fy * fx)
        else:
            a = a[:h, :w]
        out += gain * a
    out = (out - out.min()) / (out.max() - out.min())
    return (out * 255.0).astype(np.float32)


class CorridorScene:
    """Ground plane y=cam_h, walls x=-half_w / x=+half_w, far wall z=length."""

    def __init__(self, shape="kitti", seed=0, length=260.0, half_w=7.0, cam_h=1.65, texels_per_m=24.0,
                 step=0.8, yaw_amp=0.04, yaw_period=90.0, scale=1.0, heading=0.0, z0=2.0):
        w, h, K = SHAPES[shape]
        self.w, self.h = int(round(w * scale)), int(round(h * scale))
        self.K = K.copy()
        self.K[:2] *= scale
        self.length, self.half_w, self.cam_h, self.tpm = length, half_w, cam_h, texels_per_m
        self.step, self.yaw_amp, self.yaw_period = step, yaw_amp, yaw_period
        self.z0 = z0
        self.heading = heading     # pi/2: the camera looks at the right wall and moves sideways (parking-shaped)
        self.wall_h = 12.0
        L = int(length * texels_per_m) + 8
        self.tex = {
            "ground": make_texture(int(2 * half_w * texels_per_m) + 8, L, seed * 10 + 1),
            "left": make_texture(int(self.wall_h * texels_per_m) + 8, L, seed * 10 + 2),
            "right": make_texture(int(self.wall_h * texels_per_m) + 8, L, seed * 10 + 3),
            "far": make_texture(int(self.wall_h * texels_per_m) + 8, int(2 * half_w * texels_per_m) + 8, seed * 10 + 4),
        }

    # ---- trajectory -------------------------------------------------------------------------------
    def pose(self, k):
        """World->camera 4x4 of frame k: forward motion `step` m/frame with a sinusoidal yaw/sway."""
        z = self.z0 + self.step * k
        yaw = self.heading + self.yaw_amp * np.sin(2 * np.pi * k / self.yaw_period)
        x = 0.8 * np.sin(2 * np.pi * k / (1.7 * self.yaw_period))
        c, s = np.cos(yaw), np.sin(yaw)
        Rwc = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])   # camera->world
        T = np.eye(4)
        T[:3, :3] = Rwc.T
        T[:3, 3] = -Rwc.T @ np.array([x, 0.0, z])
        return T

    # ---- ray casting ---------------------------------------------------------------------------------
    def _cast(self, T, u, v):
        """Nearest plane hit of the rays through pixels (u, v): returns depth along z_cam, world point,
        plane id (0 ground, 1 left, 2 right, 3 far) and texture coordinates."""
        K = self.K
        d_c = np.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], np.ones_like(u)], -1)
        Rcw = T[:3, :3]
        o = -Rcw.T @ T[:3, 3]
        d = d_c @ Rcw        # = Rcw.T applied to each row
        big = 1e30
        with np.errstate(divide="ignore", invalid="ignore"):
            tg = np.where(d[..., 1] > 1e-9, (self.cam_h - o[1]) / d[..., 1], big)
            tl = np.where(d[..., 0] < -1e-9, (-self.half_w - o[0]) / d[..., 0], big)
            tr = np.where(d[..., 0] > 1e-9, (self.half_w - o[0]) / d[..., 0], big)
            tf = np.where(d[..., 2] > 1e-9, (self.length - o[2]) / d[..., 2], big)
        ts = np.stack([tg, tl, tr, tf], -1)
        ts = np.where(ts > 1e-6, ts, big)
        pid = np.argmin(ts, -1)
        t = np.take_along_axis(ts, pid[..., None], -1)[..., 0]
        Xw = o + d * t[..., None]
        tpm = self.tpm
        tu = np.where(pid == 3, (Xw[..., 0] + self.half_w) * tpm, Xw[..., 2] * tpm)
        tv = np.where(pid == 0, (Xw[..., 0] + self.half_w) * tpm, (self.cam_h - Xw[..., 1]) * tpm)
        return t, Xw, pid, tu, tv

    def render(self, k):
        """Frame k as an (h, w) uint8 image."""
        T = self.pose(k)
        v, u = np.mgrid[0:self.h, 0:self.w].astype(np.float64)
        _, _, pid, tu, tv = self._cast(T, u, v)
        img = np.zeros((self.h, self.w), np.float32)
        for i, name in enumerate(["ground", "left", "right", "far"]):
            m = pid == i
            if not m.any():
                continue
            tex = self.tex[name]
            x = np.clip(tu[m], 0, tex.shape[1] - 1.001)
            per = tex.shape[0] - 1.001            # fold (mirror) instead of clamping: no streaks
            y = np.abs(np.mod(tv[m] + per, 2 * per) - per)
            y = np.clip(y, 0, per)
            x0 = x.astype(np.int64); y0 = y.astype(np.int64)
            fx = (x - x0).astype(np.float32); fy = (y - y0).astype(np.float32)
            img[m] = (tex[y0, x0] * (1 - fx) * (1 - fy) + tex[y0, x0 + 1] * fx * (1 - fy)
                      + tex[y0 + 1, x0] * (1 - fx) * fy + tex[y0 + 1, x0 + 1] * fx * fy)
        return np.clip(np.rint(img), 0, 255).astype(np.uint8)

    def backproject(self, k, pts):
        """World coordinates of the scene points seen at pixel positions pts (n x 2) in frame k."""
        pts = np.asarray(pts, np.float64).reshape(-1, 2)
        _, Xw, _, _, _ = self._cast(self.pose(k), pts[:, 0], pts[:, 1])
        return Xw

    def project(self, k, Xw):
        T = self.pose(k)
        Xc = Xw @ T[:3, :3].T + T[:3, 3]
        return np.stack([self.K[0, 0] * Xc[:, 0] / Xc[:, 2] + self.K[0, 2],
                         self.K[1, 1] * Xc[:, 1] / Xc[:, 2] + self.K[1, 2]], 1), Xc[:, 2]


def corner_points(img, n, seed, min_dist=6, border=12):
    """Up to n well-textured pixel positions (float32-exact) of img: local maxima of the Shi-Tomasi
    min-eigenvalue map (box 5x5 over central-difference gradients), greedily thinned on a grid of
    cell size min_dist.  Pure numpy stand-in for the reference's CPU detectors, which are excluded
    from the timed region."""
    f = img.astype(np.float64)
    gx = np.zeros_like(f); gy = np.zeros_like(f)
    gx[:, 1:-1] = f[:, 2:] - f[:, :-2]
    gy[1:-1] = f[2:] - f[:-2]
    a = _box_blur(gx * gx, 2); b = _box_blur(gx * gy, 2); c = _box_blur(gy * gy, 2)
    lam = 0.5 * (a + c - np.sqrt((a - c) ** 2 + 4 * b * b))
    lam[:border] = 0; lam[-border:] = 0; lam[:, :border] = 0; lam[:, -border:] = 0
    h, w = lam.shape
    gh, gw = h // min_dist, w // min_dist
    cells = lam[:gh * min_dist, :gw * min_dist].reshape(gh, min_dist, gw, min_dist).transpose(0, 2, 1, 3).reshape(gh, gw, -1)
    arg = cells.argmax(-1)
    val = cells.max(-1)
    ys = (np.arange(gh)[:, None] * min_dist + arg // min_dist).ravel()
    xs = (np.arange(gw)[None, :] * min_dist + arg % min_dist).ravel()
    val = val.ravel()
    keep = val > 0.02 * val.max()
    ys, xs, val = ys[keep], xs[keep], val[keep]
    rng = np.random.default_rng(seed)
    if len(val) > n:   # prefer strong corners, but spread: sample without replacement, weight = rank
        order = np.argsort(-val)[: max(n * 2, n + 50)]
        order = rng.permutation(order)[:n]
        ys, xs = ys[order], xs[order]
    return np.stack([xs, ys], 1).astype(np.float32)


_RENDER_SCENE = None


def _render_one(k):
    return _RENDER_SCENE.render(k)


def render_frames(scene, ks, workers=None):
    """Frames ks of `scene`, rendered by a fork pool when there are many full-size ones (the per-pixel numpy ray
    caster takes ~0.2 s per KITTI frame on one core); bit-identical to rendering them one by one."""
    import multiprocessing as mp
    import os
    global _RENDER_SCENE
    ks = list(ks)
    if workers is None:
        workers = min(os.cpu_count() or 1, 16) if len(ks) * scene.w * scene.h > 6e6 else 1
    if workers <= 1 or len(ks) < 4 or mp.current_process().daemon:
        return [scene.render(k) for k in ks]
    _RENDER_SCENE = scene
    try:
        with mp.get_context("fork").Pool(workers) as pool:
            return pool.map(_render_one, ks)
    finally:
        _RENDER_SCENE = None


class Sequence:
    """A rendered sequence plus everything a VO run needs: frames, K, ground-truth poses, the
    bootstrap state (true 3-D points of corner pixels of frame 0) and a per-frame candidate feed
    (corner pixels of each new frame - stands in for the reference's CPU detectors)."""

    def __init__(self, shape="kitti", n_frames=32, n_tracks=2000, n_feed=300, seed=0, scale=1.0, workers=None, **scene_kw):
        self.scene = CorridorScene(shape, seed=seed, scale=scale, **scene_kw)
        self.K = self.scene.K
        self.frames = render_frames(self.scene, range(n_frames), workers)
        self.poses = [self.scene.pose(k) for k in range(n_frames)]
        md = 6 if scale >= 0.75 else 4
        self.P0 = corner_points(self.frames[0], n_tracks, seed + 100, min_dist=md)
        self.X0 = self.scene.backproject(0, self.P0)
        self.feeds = [np.zeros((0, 2))] + [corner_points(self.frames[k], n_feed, seed + 200 + k, min_dist=md).astype(np.float64)
                                           for k in range(1, n_frames)]

    def checksum(self):
        return int(sum(int(f.astype(np.int64).sum()) * (i + 1) for i, f in enumerate(self.frames)))


def ate(est_T, gt_T):
    """RMS absolute trajectory error of camera centres after a Sim(3) (Umeyama) alignment; the
    reference defines no ATE and monocular scale is arbitrary."""
    a = np.array([(-T[:3, :3].T @ T[:3, 3]) for T in est_T])
    b = np.array([(-T[:3, :3].T @ T[:3, 3]) for T in gt_T])
    ma, mb = a.mean(0), b.mean(0)
    A, B = a - ma, b - mb
    U, S, Vt = np.linalg.svd(B.T @ A / len(a))
    D = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        D[2, 2] = -1
    R = U @ D @ Vt
    s = np.trace(np.diag(S) @ D) / max((A ** 2).sum() / len(a), 1e-300)
    al = (s * (R @ A.T)).T + mb
    return float(np.sqrt(((al - b) ** 2).sum(1).mean()))


# ---- steady-state streams for bench.py (SURVEY.md 8a sizing: |P| ~ n_tracks held stationary) ---------------
def bearing_deg(c, f):
    """The reference's bearing angle between two raw-pixel observations (vo_pipeline.py:621-642)."""
    a = np.concatenate([c, np.ones((len(c), 1))], 1)
    b = np.concatenate([f, np.ones((len(f), 1))], 1)
    a /= np.linalg.norm(a, axis=1, keepdims=True)
    b /= np.linalg.norm(b, axis=1, keepdims=True)
    return np.degrees(np.arccos(np.clip((a * b).sum(1), -1.0, 1.0)))


_POOL_SCENE = None


def _pool_frame(args):
    k, n_feed, seed, md, rate_lo, rate_hi, horizon, n_burst = args
    sc = _POOL_SCENE
    img = sc.render(k)
    out = {"k": k, "img": img}
    if n_feed > 0:
        # a detector that only offers what will be promotable: corners whose bearing (as the reference
        # measures it, on raw pixels) grows by >= alpha / horizon per frame and that stay inside the image
        # for `horizon` more frames; the rest of the corners (far wall, image centre) would sit in C forever
        pts = corner_points(img, 4 * n_feed, seed + 200 + k, min_dist=md)
        Xw = sc.backproject(k, pts)
        p1, z1 = sc.project(k + 1, Xw)
        ph, zh = sc.project(k + horizon, Xw)
        rate = bearing_deg(pts.astype(np.float64), p1)
        ok = (rate >= rate_lo) & (rate <= rate_hi) & (z1 > 0.5) & (zh > 0.5)
        ok &= (ph[:, 0] > 8) & (ph[:, 0] < sc.w - 8) & (ph[:, 1] > 8) & (ph[:, 1] < sc.h - 8)
        out["feed"] = pts[ok][:n_feed].astype(np.float64)
    if n_burst > 0:
        # a stale external map: well-textured pixels of this frame paired with the 3-D points of OTHER
        # pixels, so KLT keeps them and PnP has to reject them
        pts = corner_points(img, n_burst, seed + 7000 + k, min_dist=md)
        Xw = sc.backproject(k, pts)
        out["burst_P"] = pts
        out["burst_X"] = np.roll(Xw, len(Xw) // 2 + 1, axis=0)
    return out


class SteadyStream:
    """Forward drive down a long corridor, one rendered frame per step (no ping-pong inside `n_frames`),
    with what bench.py needs to hold the landmark count stationary:

      feeds[k]            promotable corner pixels of frame k (candidate feed)
      burst_P/burst_X[k]  an outlier landmark batch for frame k (relocalisation against a stale map): appended
                          to P/X after frame k on the frames bench.py chooses, so that PnP-RANSAC really iterates

    Frames are rendered by a fork pool (numpy ray casting, ~0.2 s per KITTI frame on one core)."""

    def __init__(self, shape="kitti", n_frames=64, n_tracks=2000, n_feed=600, n_burst=0, seed=0, rate_lo=0.08,
                 rate_hi=0.6, horizon=12, workers=None, scale=1.0, step=0.8, **scene_kw):
        import multiprocessing as mp
        import os
        global _POOL_SCENE
        self.scene = CorridorScene(shape, seed=seed, scale=scale, step=step,
                                   length=scene_kw.get("z0", 2.0) + step * (n_frames + horizon + 2) + 90.0, **scene_kw)
        self.K = self.scene.K
        md = 6 if scale >= 0.75 else 4
        jobs = [(k, n_feed if k > 0 else 0, seed, md, rate_lo, rate_hi, horizon, n_burst) for k in range(n_frames)]
        workers = workers or min(os.cpu_count() or 1, 16)
        _POOL_SCENE = self.scene
        if workers > 1 and n_frames > 4:
            with mp.get_context("fork").Pool(workers) as pool:
                res = pool.map(_pool_frame, jobs, chunksize=max(1, n_frames // (4 * workers)))
        else:
            res = [_pool_frame(j) for j in jobs]
        _POOL_SCENE = None
        self.frames = [r["img"] for r in res]
        self.poses = [self.scene.pose(k) for k in range(n_frames)]
        self.feeds = [r.get("feed", np.zeros((0, 2))) for r in res]
        self.burst_P = [r.get("burst_P", np.zeros((0, 2), np.float32)) for r in res]
        self.burst_X = [r.get("burst_X", np.zeros((0, 3))) for r in res]
        self.P0 = corner_points(self.frames[0], n_tracks, seed + 100, min_dist=md)
        self.X0 = self.scene.backproject(0, self.P0)

    def checksum(self):
        return int(sum(int(f.astype(np.int64).sum()) * (i + 1) for i, f in enumerate(self.frames)))


# ---- datasets on disk in the reference's layouts (main.py:14-68) -------------------------------------------------------
def write_dataset(seq, root, layout="parking", n_frames=None):
    """Write the frames and ground-truth poses of `seq` (Sequence / SteadyStream) under `root` the way the
    reference's main.py reads them: parking/images/img_%05d.png + parking/poses.txt (12 columns, row-major 3x4
    camera->world, main.py:48-64), kitti-dataset/05/image_0/%06d.png + kitti-dataset/poses/05.txt (main.py:14-31; main.py
    counts files ending in 'left.png' there, so as many placeholders are written), or the Malaga folder with
    *left.jpg frames and the GPS table whose columns 8, 9 it plots (main.py:33-46).  Returns the frame paths."""
    import os
    import cv2
    n = len(seq.frames) if n_frames is None else min(n_frames, len(seq.frames))
    rows = []
    for k in range(n):
        Twc = np.linalg.inv(seq.poses[k])
        rows.append(Twc[:3].ravel())
    rows = np.array(rows)
    paths = []
    if layout == "parking":
        d = os.path.join(root, "parking", "images")
        os.makedirs(d, exist_ok=True)
        for k in range(n):
            paths.append(os.path.join(d, "img_%05d.png" % k))
            cv2.imwrite(paths[-1], seq.frames[k])
        np.savetxt(os.path.join(root, "parking", "poses.txt"), rows)
    elif layout == "kitti":
        d = os.path.join(root, "kitti-dataset", "05", "image_0")
        os.makedirs(d, exist_ok=True)
        os.makedirs(os.path.join(root, "kitti-dataset", "poses"), exist_ok=True)
        for k in range(n):
            paths.append(os.path.join(d, "%06d.png" % k))
            cv2.imwrite(paths[-1], seq.frames[k])
            open(os.path.join(d, "%06d_left.png" % k), "wb").close()
        np.savetxt(os.path.join(root, "kitti-dataset", "poses", "05.txt"), rows)
    elif layout == "malaga":
        base = os.path.join(root, "malaga-urban-dataset-extract-07")
        d = os.path.join(base, "malaga-urban-dataset-extract-07_rectified_800x600_Images")
        os.makedirs(d, exist_ok=True)
        for k in range(n):
            paths.append(os.path.join(d, "img_CAMERA1_%010d_left.jpg" % k))
            cv2.imwrite(paths[-1], seq.frames[k], [cv2.IMWRITE_JPEG_QUALITY, 98])
        gps = np.zeros((n, 10))
        gps[:, 8], gps[:, 9] = rows[:, 3], rows[:, 11]
        np.savetxt(os.path.join(base, "malaga-urban-dataset-extract-07_all-sensors_GPS.txt"), gps)
    else:
        raise ValueError("layout must be parking, kitti or malaga")
    for sub in ("result-trajectory-plots", "initialization-plots-and-figures", "ground-truth-trajectories"):
        os.makedirs(os.path.join(root, sub), exist_ok=True)
    return paths
