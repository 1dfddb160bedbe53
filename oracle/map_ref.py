"""TEST INFRASTRUCTURE - CPU restatement of the reference's trajectory / map bookkeeping.

Follows /root/reference/src/main.py:239-255 (per-frame camera position, landmarks between 3 and 50
units from it appended to the plotting arrays) with the same numpy calls, and restates the raster
rules of csrc/map.cu (the reference draws with matplotlib, main.py:262-271 / :296-305, which is not
on this box: the plot itself is "parity unpinned" - only the DATA that is plotted, the axis limits
rule (matplotlib's default 5 % margins) and the colours follow the reference).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
import numpy as np


class MapRef:
    """poses_plotting / X_plotting of main.py:209-255."""

    def __init__(self, r_min=3.0, r_max=50.0):
        self.r_min, self.r_max = r_min, r_max
        self.X_plotting = np.zeros((0, 3))
        self.poses_plotting = np.zeros((0, 3))

    def update(self, new_pose, X):
        """main.py:239-255 for one frame: X is vision.state['X'] (the landmarks the frame started from)."""
        pos_h = np.zeros(4)
        pos_h[3] = 1
        pos = (np.linalg.inv(new_pose) @ pos_h)[0:3]                       # main.py:240-242
        self.poses_plotting = np.vstack((self.poses_plotting, pos))        # main.py:245
        X = np.asarray(X, np.float64).reshape(-1, 3)
        flt = np.linalg.norm(X - pos, axis=1) < self.r_max                 # main.py:249
        flt_add = np.linalg.norm(X - pos, axis=1) > self.r_min             # main.py:250
        flt = flt * flt_add                                                # main.py:251
        X_filtered = X[flt, :]
        self.X_plotting = np.vstack((self.X_plotting, X_filtered))         # main.py:255
        return pos, flt


def plot_limits(traj, pts):
    """Axis limits {xlo, xhi, zlo, zhi}: data range of landmarks and path + 5 % on each side."""
    xs = np.concatenate([traj[:, 0], pts[:, 0]])
    zs = np.concatenate([traj[:, 2], pts[:, 2]])
    out = []
    for v in (xs, zs):
        if len(v) == 0:
            out += [0.0, 1.0]
            continue
        lo, hi = float(v.min()), float(v.max())
        span = hi - lo
        if span == 0.0:
            out += [lo - 0.5, hi + 0.5]
        else:
            mg = 0.05 * span
            out += [lo - mg, hi + mg]
    return np.array(out)


def to_pixel(b, x, z, W, H, m):
    u = m + (x - b[0]) / (b[1] - b[0]) * float(W - 1 - 2 * m)
    v = float(H - 1 - m) - (z - b[2]) / (b[3] - b[2]) * float(H - 1 - 2 * m)
    return np.floor(u + 0.5).astype(np.int64), np.floor(v + 0.5).astype(np.int64)


def render(traj, pts, W, H, m, rad):
    """BGR canvas drawn by the rules in csrc/map.cu (white, black box, blue discs, red 3 px path on top)."""
    traj = np.asarray(traj, np.float64).reshape(-1, 3)
    pts = np.asarray(pts, np.float64).reshape(-1, 3)
    b = plot_limits(traj, pts)
    img = np.full((H, W, 3), 255, np.uint8)
    if m >= 1:
        img[m - 1, m - 1:W - m + 1] = 0
        img[H - m, m - 1:W - m + 1] = 0
        img[m - 1:H - m + 1, m - 1] = 0
        img[m - 1:H - m + 1, W - m] = 0

    def put(x, y, col):
        ok = (x >= 0) & (x < W) & (y >= 0) & (y < H)
        img[y[ok], x[ok]] = col

    if len(pts):
        cx, cy = to_pixel(b, pts[:, 0], pts[:, 2], W, H, m)
        for dy in range(-rad, rad + 1):
            for dx in range(-rad, rad + 1):
                if dx * dx + dy * dy <= rad * rad:
                    put(cx + dx, cy + dy, (255, 0, 0))
    if len(traj):
        ax, ay = to_pixel(b, traj[:, 0], traj[:, 2], W, H, m)
        segs = [(0, 0)] if len(traj) == 1 else [(k, k + 1) for k in range(len(traj) - 1)]
        for k0, k1 in segs:
            dx, dy = int(ax[k1] - ax[k0]), int(ay[k1] - ay[k0])
            n = max(abs(dx), abs(dy))
            i = np.arange(n + 1, dtype=np.int64)
            x = ax[k0] + ((2 * dx * i + n) // (2 * n) if n else 0 * i)
            y = ay[k0] + ((2 * dy * i + n) // (2 * n) if n else 0 * i)
            for oy in (-1, 0, 1):
                for ox in (-1, 0, 1):
                    put(x + ox, y + oy, (0, 0, 255))
    return img, b
