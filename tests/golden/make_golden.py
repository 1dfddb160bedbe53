"""Generates tests/golden/*.npz in the BUILD container, where cv2 and /root/reference exist.

  python tests/golden/make_golden.py

The goldens pin the oracle (and through it the CUDA path) to outputs of the real thing:
cv2 (version recorded in each file) called exactly as the reference calls it, and the reference's
own classes imported from /root/reference/src (matplotlib stubbed; it is not installed).
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cv2  # noqa: E402
import synth  # noqa: E402


def klt():
    from test_klt_oracle import KLT_CASES
    for name, (h, w, seed, n) in KLT_CASES.items():
        prev, nxt, pts = synth.klt_pair(h, w, seed, n)
        # exactly the reference's call: vo_pipeline.py:215 / :501
        next_pts, status, err = cv2.calcOpticalFlowPyrLK(prev, nxt, pts, None)
        np.savez_compressed(os.path.join(HERE, "klt_%s.npz" % name), pts=pts, next_pts=next_pts,
                            status=status, err=err, prev_sum=prev.astype(np.int64).sum(),
                            next_sum=nxt.astype(np.int64).sum(), cv2_version=cv2.__version__)
    img = synth.texture(75, 101, 11)
    lv, out = img, {}
    for l in range(1, 4):
        lv = cv2.pyrDown(lv)
        out["l%d" % l] = lv
    np.savez_compressed(os.path.join(HERE, "pyrdown.npz"), cv2_version=cv2.__version__, **out)


if __name__ == "__main__":
    which = sys.argv[1:] or ["klt"]
    for w in which:
        globals()[w]()
    print("golden vectors written with cv2", cv2.__version__)
