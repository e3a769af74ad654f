"""Golden vectors for the rotation_search branch (utils/parse_skeletons.py:215-218, :265-267 predict; :63-66, :98-100
predict_refactor; knob utils/config:15), produced by RUNNING THE REFERENCE'S predict() / predict_refactor() with a
non-zero rotation_search and the stand-ins of make_golden_ms.py (`.cuda()` = identity; for predict the `multiplier = [1.]`
pin removed from an in-memory copy of the source when the case has a scale search; a model stub that RECORDS its input
and returns canned synthetic network outputs, one per call).  IPP off.  The image is make_golden_resize.image(seed);
model inputs and results are pinned by sha256 + strided probes."""
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

from improved_body_parts_b200 import synth  # noqa: E402
from make_golden_resize import image  # noqa: E402
from oracle import ref_python  # noqa: E402

CASES = [  # name, function, seed, people, h, w, scale_search (None = as shipped: 1.0), rotation_search
    ("rot_111", "predict", 11101, 4, 240, 320, (1.0, 1.5), (0.0, 30.0)),
    ("rot_112", "predict", 11102, 3, 200, 150, None, (-20.0,)),
    ("rot_113", "predict_refactor", 11103, 5, 240, 320, None, (6.0,)),
    ("rot_114", "predict_refactor", 11104, 2, 150, 210, None, (0.0, -4.0)),
]


def canned_outputs(seed, people, h, w, ss):
    """One (2,50,fh,fw) output per scale (the same maps serve every angle of that scale)."""
    if ss is None:
        ph, pw = -(-h // 64) * 64, -(-w // 64) * 64
        rng = np.random.default_rng(seed)
        clean = synth.render_maps(synth.random_people(rng, people, w, h), ph, pw, 4)
        return {(ph, pw): np.stack([synth.add_noise(clean, rng, 0.01), synth.add_noise(synth.flip_twin(clean), rng, 0.01)])}
    shapes = synth.multiscale_shapes(h, w, ss)
    outs, _ = synth.make_multiscale_output(seed, people, h, w, ss)
    return {tuple(s[2]): o for s, o in zip(shapes, outs)}


def main():
    ref = ref_python.load()
    PS = ref.parse_skeletons
    for name, fn, seed, people, h, w, ss, angles in CASES:
        func = getattr(PS, fn)
        if ss is not None:
            src = inspect.getsource(func)
            pinned = [l for l in src.splitlines() if l.strip().startswith("multiplier = [1.]")]
            assert len(pinned) == 1
            ns = dict(PS.__dict__)
            exec(src.replace(pinned[0], ""), ns)
            func = ns[fn]
        img = image(seed, h, w)
        canned = canned_outputs(seed, people, h, w, ss)
        seen = []

        def model(x):
            seen.append(x.numpy().copy())
            return [[torch.from_numpy(canned[(int(x.shape[1]), int(x.shape[2]))])]]

        cfg = dict(ref.test_cfg)
        cfg["rotation_search"] = list(angles)
        if ss is not None:
            cfg["scale_search"] = list(ss)
        old = torch.Tensor.cuda
        torch.Tensor.cuda = lambda self, *a, **k: self
        cv2.ipp.setUseIPP(False)
        try:
            heat, paf = func(img, model, cfg, ref.model_cfg, "synthetic", flip_avg=True, config=ref.config)
        finally:
            torch.Tensor.cuda = old
        d = dict(fn=np.array(fn), seed=np.int64(seed), people=np.int64(people), shape=np.array([h, w]),
                 scale_search=np.array(ss if ss is not None else [-1.0]), rotation_search=np.array(angles),
                 img_sha256=np.array(hashlib.sha256(img.tobytes()).hexdigest()), n_calls=np.int64(len(seen)),
                 heat_dtype=np.array(str(heat.dtype)),
                 heat_sha256=np.array(hashlib.sha256(np.ascontiguousarray(heat).tobytes()).hexdigest()),
                 paf_sha256=np.array(hashlib.sha256(np.ascontiguousarray(paf).tobytes()).hexdigest()),
                 heat_probe=heat[::11, ::13].copy(), paf_probe=paf[::11, ::13].copy(),
                 versions=np.array("opencv %s (IPP off); numpy %s" % (cv2.__version__, np.__version__)))
        for k, x in enumerate(seen):
            d["in_shape_%d" % k] = np.array(x.shape)
            d["in_sha256_%d" % k] = np.array(hashlib.sha256(np.ascontiguousarray(x).tobytes()).hexdigest())
            d["in_probe_%d" % k] = x[:, ::23, ::19, :].copy()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **d)
        print(name, fn, heat.shape, heat.dtype, len(seen), "calls", float(heat.max()))


if __name__ == "__main__":
    main()
