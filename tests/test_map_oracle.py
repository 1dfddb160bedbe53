"""CPU tests of the trajectory / map bookkeeping oracle (oracle/map_ref.py): it restates
main.py:239-255 with the reference's own numpy calls; here its output is checked against closed forms
and the raster rules against hand-computable cases."""
import numpy as np

from oracle import map_ref


def rigid(rng):
    q = rng.normal(size=4); q /= np.linalg.norm(q)
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    T = np.eye(4); T[:3, :3] = R; T[:3, 3] = rng.normal(size=3) * 5
    return T


def test_update_follows_main_py():
    rng = np.random.default_rng(3)
    mp = map_ref.MapRef()
    tot = 0
    for k in range(5):
        T = rigid(rng)
        X = rng.uniform(-60, 60, size=(400, 3))
        pos, flt = mp.update(T, X)
        assert np.allclose(pos, -T[:3, :3].T @ T[:3, 3], rtol=0, atol=1e-12)     # inv of a rigid pose
        d = np.sqrt(((X - pos) ** 2).sum(axis=1))
        assert np.array_equal(flt, (d < 50) & (d > 3))
        tot += int(flt.sum())
        assert mp.X_plotting.shape == (tot, 3) and mp.poses_plotting.shape == (k + 1, 3)
        assert np.array_equal(mp.X_plotting[tot - int(flt.sum()):], X[flt])        # order kept
    pos, flt = mp.update(np.eye(4), np.zeros((0, 3)))                              # empty state
    assert flt.shape == (0,) and mp.X_plotting.shape == (tot, 3)


def test_render_rules():
    traj = np.array([[0., 0, 0], [10, 0, 0], [10, 0, 10]])
    pts = np.array([[5., 1, 5], [0, 0, 10]])
    img, b = map_ref.render(traj, pts, 200, 100, 10, 2)
    assert np.allclose(b, [-0.5, 10.5, -0.5, 10.5])
    assert img.shape == (100, 200, 3) and (img[0, 0] == 255).all()
    assert (img[9, 9] == 0).all() and (img[90, 190] == 0).all() and (img[9, 100] == 0).all()   # box
    cx, cy = map_ref.to_pixel(b, np.array([5.]), np.array([5.]), 200, 100, 10)
    assert tuple(img[cy[0], cx[0]]) == (255, 0, 0)                                  # landmark: blue (BGR)
    assert tuple(img[cy[0] + 3, cx[0]]) == (255, 255, 255)                          # radius 2
    ax, ay = map_ref.to_pixel(b, traj[:, 0], traj[:, 2], 200, 100, 10)
    for x, y in zip(ax, ay):
        assert tuple(img[y, x]) == (0, 0, 255)                                      # path: red, on top
    mid = ((ax[0] + ax[1]) // 2, ay[0])
    assert tuple(img[mid[1], mid[0]]) == (0, 0, 255) and tuple(img[mid[1] - 2, mid[0]]) == (255, 255, 255)
    # degenerate inputs: nothing / one pose / identical points
    img0, b0 = map_ref.render(np.zeros((0, 3)), np.zeros((0, 3)), 64, 48, 4, 3)
    assert np.array_equal(b0, [0, 1, 0, 1]) and set(np.unique(img0)) <= {0, 255}
    img1, b1 = map_ref.render(np.array([[2., 0, 3]]), np.zeros((0, 3)), 64, 48, 4, 3)
    assert np.allclose(b1, [1.5, 2.5, 2.5, 3.5]) and (img1[..., 2] == 255).any() and (img1 == (0, 0, 255)).all(axis=2).sum() == 9
