"""Generates tests/golden/resize_9x.npz by RUNNING cv2.resize itself — the third-party call the reference issues at
utils/parse_skeletons.py:204 / :52: cv2.resize(image, (0, 0), fx=scale, fy=scale, interpolation=cv2.INTER_CUBIC) on a
uint8 BGR image — with IPP off (OpenCV's own portable code path).  The scales are the reference's
multipliers x * boxsize / img_h for its scale_search (utils/config:14) and a few awkward ones.  Inputs are regenerated
by the test from the seed; the output is stored whole when small, else as sha256 + a strided probe.  The number of
output bytes that differ with IPP ON (stock cv2) is stored as `ipp_mismatch` / `ipp_maxdiff`.
Re-run with `python tests/golden/make_golden_resize.py`."""
import hashlib
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [  # (id, seed, h, w, fx)
    (91, 9101, 480, 640, 0.5 * 512 / 480), (92, 9102, 480, 640, 1.0 * 512 / 480), (93, 9103, 240, 320, 1.5 * 512 / 240),
    (94, 9104, 97, 131, 2.0 * 512 / 480), (95, 9105, 64, 64, 0.7), (96, 9106, 50, 77, 3.0 * 512 / 480), (97, 9107, 33, 5, 1.37),
    (98, 9108, 426, 640, 512 / 426),
]


def image(seed, h, w):
    """A blocky random BGR image (low-frequency field + noise), uint8; numpy only, so it does not depend on cv2's state."""
    rng = np.random.default_rng(seed)
    low = rng.integers(0, 256, (h // 8 + 1, w // 8 + 1, 3)).astype(np.float32)
    img = np.kron(low, np.ones((8, 8, 1), dtype=np.float32))[:h, :w] + rng.normal(0, 12, (h, w, 3)).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def main():
    for cid, seed, h, w, fx in CASES:
        img = image(seed, h, w)
        cv2.ipp.setUseIPP(False)
        out = cv2.resize(img, (0, 0), fx=fx, fy=fx, interpolation=cv2.INTER_CUBIC)
        cv2.ipp.setUseIPP(True)
        stock = cv2.resize(img, (0, 0), fx=fx, fy=fx, interpolation=cv2.INTER_CUBIC)
        cv2.ipp.setUseIPP(False)
        diff = np.abs(out.astype(np.int32) - stock.astype(np.int32))
        d = dict(seed=np.int64(seed), shape=np.array([h, w]), fx=np.float64(fx), out_shape=np.array(out.shape),
                 img_sha256=np.array(hashlib.sha256(img.tobytes()).hexdigest()),
                 out_sha256=np.array(hashlib.sha256(out.tobytes()).hexdigest()), probe=out[::7, ::5].copy(),
                 ipp_mismatch=np.int64((diff > 0).sum()), ipp_maxdiff=np.int64(diff.max()),
                 versions=np.array("opencv %s (IPP off); numpy %s" % (cv2.__version__, np.__version__)))
        if out.size <= 30000:
            d["out"] = out
        np.savez_compressed(os.path.join(HERE, "resize_%d.npz" % cid), **d)
        print(cid, img.shape, "->", out.shape, "ipp mismatch", int(d["ipp_mismatch"]), "max", int(d["ipp_maxdiff"]))


if __name__ == "__main__":
    main()
