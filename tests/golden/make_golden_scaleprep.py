"""Golden vectors for the per-scale image preparation of the scale search (utils/parse_skeletons.py:196-224): what the
reference's predict() hands to the model at every scale — cv2.resize(image, fx=scale, INTER_CUBIC) on the uint8 image,
padRightDownCorner to a multiple of 64 with 128, np.float32(x / 255), the flipped twin.  Produced by RUNNING THE
REFERENCE'S predict() with the stand-ins of make_golden_ms.py (`.cuda()` = identity, the `multiplier = [1.]` pin removed
from an in-memory copy of the source) and a model stub that RECORDS its input and returns zeros.  IPP off.
The image comes from make_golden_resize.image(seed, h, w); the tensors are pinned by sha256 + a strided probe."""
import hashlib
import inspect
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import cv2  # noqa: E402
import torch  # noqa: E402

from make_golden_resize import image  # noqa: E402
from oracle import ref_python  # noqa: E402

CASES = [("scaleprep_101", 10101, 480, 640, (0.5, 1.0, 1.5, 2.0)), ("scaleprep_102", 10102, 427, 640, (0.5, 1.0, 1.5)),
         ("scaleprep_103", 10103, 375, 500, (1.0, 2.0, 3.0))]


def main():
    ref = ref_python.load()
    PS = ref.parse_skeletons
    src = inspect.getsource(PS.predict)
    pinned = [l for l in src.splitlines() if l.strip().startswith("multiplier = [1.]")]
    assert len(pinned) == 1
    ns = dict(PS.__dict__)
    exec(src.replace(pinned[0], ""), ns)
    predict = ns["predict"]
    for name, seed, h, w, ss in CASES:
        img = image(seed, h, w)
        seen = []

        def model(x):
            seen.append(x.numpy().copy())
            return [[torch.zeros((2, 50, x.shape[1] // 4, x.shape[2] // 4))]]

        cfg = dict(ref.test_cfg)
        cfg["scale_search"] = list(ss)
        old = torch.Tensor.cuda
        torch.Tensor.cuda = lambda self, *a, **k: self
        cv2.ipp.setUseIPP(False)
        try:
            predict(img, model, cfg, ref.model_cfg, "synthetic", flip_avg=True, config=ref.config)
        finally:
            torch.Tensor.cuda = old
        assert len(seen) == len(ss)
        d = dict(seed=np.int64(seed), shape=np.array([h, w]), scale_search=np.array(ss),
                 img_sha256=np.array(hashlib.sha256(img.tobytes()).hexdigest()),
                 versions=np.array("opencv %s (IPP off); numpy %s" % (cv2.__version__, np.__version__)))
        for k, x in enumerate(seen):
            d["shape_%d" % k] = np.array(x.shape)
            d["sha256_%d" % k] = np.array(hashlib.sha256(np.ascontiguousarray(x).tobytes()).hexdigest())
            d["probe_%d" % k] = x[:, ::23, ::19, :].copy()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **d)
        print(name, [x.shape for x in seen])


if __name__ == "__main__":
    main()
