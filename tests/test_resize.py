"""Image resize of the scale search (utils/parse_skeletons.py:204, :52): cv2.resize(image, (0,0),
fx, fy, INTER_CUBIC) on uint8.  The oracle restatement against goldens produced by cv2 itself
(tests/golden/make_golden_resize.py) and against the live cv2; the CUDA entry point (ibp_resize_image) against the
oracle and the goldens; the predict mirror, which now prepares every scale's input on the device.  Bit-exact."""
import glob
import hashlib
import os

import numpy as np
import pytest

from oracle import np_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLD, "resize_*.npz")))


def _image(seed, h, w):
    from importlib import util
    spec = util.spec_from_file_location("make_golden_resize", os.path.join(GOLD, "make_golden_resize.py"))
    mod = util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.image(seed, h, w)


def _load(path):
    g = np.load(path)
    h, w = (int(v) for v in g["shape"])
    img = _image(int(g["seed"]), h, w)
    assert hashlib.sha256(img.tobytes()).hexdigest() == str(g["img_sha256"]), "input generator drifted"
    return img, float(g["fx"]), g


def _check(out, g):
    assert tuple(out.shape) == tuple(int(v) for v in g["out_shape"])
    assert np.array_equal(out[::7, ::5], g["probe"])
    assert hashlib.sha256(np.ascontiguousarray(out).tobytes()).hexdigest() == str(g["out_sha256"])
    if "out" in g.files:
        assert np.array_equal(out, g["out"])


def test_resize_goldens_present():
    assert len(CASES) >= 8


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_oracle_resize_matches_cv2_goldens(path):
    img, fx, g = _load(path)
    _check(O.cv_resize_cubic_u8(img, fx), g)
    assert int(g["ipp_maxdiff"]) <= 1  # stock cv2 (IPP's 8-bit cubic) is within one grey level of OpenCV's own path


def test_oracle_resize_matches_live_cv2():
    """Random sizes / factors against the installed cv2 (IPP off), both factors different, 1 and 4 channels too."""
    import cv2
    cv2.ipp.setUseIPP(False)
    rng = np.random.default_rng(7)
    for _ in range(25):
        h, w, cn = int(rng.integers(3, 90)), int(rng.integers(3, 120)), int(rng.choice([1, 3, 3, 4]))
        fx, fy = float(rng.uniform(0.3, 3.5)), float(rng.uniform(0.3, 3.5))
        img = rng.integers(0, 256, (h, w, cn), dtype=np.uint8)
        ref = cv2.resize(img, (0, 0), fx=fx, fy=fy, interpolation=cv2.INTER_CUBIC).reshape(-1, int(np.rint(w * fx)), cn)
        assert np.array_equal(O.cv_resize_cubic_u8(img, fx, fy), ref), (h, w, cn, fx, fy)
    one = rng.integers(0, 256, (40, 52, 3), dtype=np.uint8)
    assert np.array_equal(O.cv_resize_cubic_u8(one, 1.0), one)  # parse_skeletons.py:36 pins the scale to 1.0: an identity


@pytest.mark.gpu
@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_gpu_resize_matches_goldens(path):
    from improved_body_parts_b200.pipeline import PostProcessor
    img, fx, g = _load(path)
    pp = PostProcessor(device=0, max_batch=1)
    _check(pp.resize_images(img, fx)[0].cpu().numpy(), g)


@pytest.mark.gpu
def test_gpu_resize_batch_and_errors():
    import torch
    from improved_body_parts_b200 import _lib
    from improved_body_parts_b200.pipeline import PostProcessor
    pp = PostProcessor(device=0, max_batch=1)
    rng = np.random.default_rng(11)
    batch = rng.integers(0, 256, (5, 61, 83, 3), dtype=np.uint8)
    for fx, fy in ((1.7, 0.6), (0.45, 2.9), (1.0, 1.0)):
        got = pp.resize_images(batch, fx, fy).cpu().numpy()
        for b in range(5):
            assert np.array_equal(got[b], O.cv_resize_cubic_u8(batch[b], fx, fy)), (fx, fy, b)
    assert _lib.resized_size(480, 640, 512 / 480, 512 / 480) == (512, 683) and _lib.resized_size(5, 5, 0.5, 0.5) == (2, 2)
    t = torch.from_numpy(batch).cuda()
    out = torch.empty((5, 10, 10, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(_lib.IbpError) as e:
        pp.handle.resize_image(t.data_ptr(), 5, 61, 83, 2.0, 2.0, out.data_ptr(), 10, 10)
    assert e.value.code == -1


SCALEPREP = sorted(glob.glob(os.path.join(GOLD, "scaleprep_*.npz")))


def _scaleprep_case(path):
    g = np.load(path)
    h, w = (int(v) for v in g["shape"])
    img = _image(int(g["seed"]), h, w)
    assert hashlib.sha256(img.tobytes()).hexdigest() == str(g["img_sha256"]), "input generator drifted"
    return img, [float(v) for v in g["scale_search"]], g


def _check_model_input(x, g, k):
    assert tuple(x.shape) == tuple(int(v) for v in g["shape_%d" % k])
    assert np.array_equal(x[:, ::23, ::19, :], g["probe_%d" % k])
    assert hashlib.sha256(np.ascontiguousarray(x).tobytes()).hexdigest() == str(g["sha256_%d" % k])


@pytest.mark.parametrize("path", SCALEPREP, ids=[os.path.basename(p) for p in SCALEPREP])
def test_oracle_scale_search_inputs_match_reference_predict(path):
    """resize (uint8 cubic) -> pad -> /255 -> flip twin == the tensor the reference's predict() gave its model, per scale."""
    img, ss, g = _scaleprep_case(path)
    assert len(SCALEPREP) >= 3
    for k, x in enumerate(ss):
        m = x * 512 / img.shape[0]                      # multiplier, parse_skeletons.py:186 with boxsize 512
        _check_model_input(O.prepare_input(O.cv_resize_cubic_u8(img, m), 64, 128, flip=True), g, k)


@pytest.mark.gpu
@pytest.mark.parametrize("path", SCALEPREP, ids=[os.path.basename(p) for p in SCALEPREP])
def test_gpu_predict_mirror_prepares_every_scale_on_the_device(path):
    """The predict mirror with a recording model: every tensor it hands to the model (resized, padded, /255, flipped — all
    on the device) equals what the reference's predict() handed to its model."""
    import torch
    from improved_body_parts_b200 import parse_skeletons as PS
    img, ss, g = _scaleprep_case(path)
    seen = []

    def model(x):
        assert x.is_cuda and x.dtype == torch.float32
        seen.append(x.cpu().numpy())
        return [[torch.zeros((2, 50, x.shape[1] // 4, x.shape[2] // 4), device=x.device)]]

    test_cfg = dict(thre1=0.1, thre2=0.1, offset_radius=2, rotation_search=[0.0], scale_search=ss)
    model_cfg = dict(boxsize=512, padValue=128, stride=4, max_downsample=64)
    heat, paf = PS.predict(img, model, test_cfg, model_cfg, "synthetic", scale_search=ss)
    assert heat.shape == img.shape[:2] + (20,) and paf.shape == img.shape[:2] + (30,) and not heat.any()
    assert len(seen) == len(ss)
    for k, x in enumerate(seen):
        _check_model_input(x, g, k)
