"""The two ends of the path: image preparation (utils/parse_skeletons.py:52-72,
utils/util.py:44-65) and COCO result rows (evaluate.py:182-209).  Golden vectors come from the reference's own
predict_refactor / append_result (tests/golden/make_golden_prep.py); the oracle is checked against them on the CPU and
the CUDA entry points (ibp_prepare_image, ibp_fetch_coco) against both on the GPU.  Bit-exact."""
import glob
import hashlib
import os

import numpy as np
import pytest

from oracle import np_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PREP = sorted(glob.glob(os.path.join(GOLD, "prep_*.npz")))
COCO = sorted(glob.glob(os.path.join(GOLD, "coco_*.npz")))


def prep_image(g):
    return np.random.default_rng(int(g["seed"])).integers(0, 256, (int(g["img_h"]), int(g["img_w"]), 3), dtype=np.uint8)


def check_prep(x, g):
    assert x.dtype == np.float32 and tuple(x.shape) == tuple(int(v) for v in g["shape"])
    assert np.array_equal(x[:, ::37, ::41, :], g["probe"])
    assert hashlib.sha256(np.ascontiguousarray(x).tobytes()).hexdigest() == str(g["sha256"])


def test_golden_present():
    assert len(PREP) >= 4 and len(COCO) >= 3


@pytest.mark.parametrize("path", PREP, ids=[os.path.basename(p) for p in PREP])
def test_oracle_prepare_input_matches_reference(path):
    g = np.load(path)
    check_prep(O.prepare_input(prep_image(g), 64, 128, True), g)


@pytest.mark.parametrize("path", COCO, ids=[os.path.basename(p) for p in COCO])
def test_oracle_coco_rows_match_reference(path):
    g = np.load(path)
    humans = [(g["ids"][k], g["scores"][k]) for k in range(g["ids"].shape[0])]
    rows = O.coco_rows(g["xy"], humans)
    assert rows.shape == g["rows"].shape and np.array_equal(rows, g["rows"])


def test_oracle_vs_reference_pad(tmp_path):
    """padRightDownCorner itself (utils/util.py:44-65) on sizes around the multiples, when the reference is present."""
    from oracle import ref_python
    if not ref_python.available():
        pytest.skip("reference tree not present")
    ref = ref_python.load()
    rng = np.random.default_rng(5)
    for h, w in [(64, 64), (63, 65), (1, 70), (130, 2), (200, 1)]:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        padded, _ = ref.util.padRightDownCorner(img, 64, 128)
        want = np.float32(padded / 255)
        got = O.prepare_input(img, 64, 128, flip=False)[0]
        assert got.shape == want.shape and np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("path", PREP, ids=[os.path.basename(p) for p in PREP])
def test_gpu_prepare_image_matches_reference(path):
    from improved_body_parts_b200.pipeline import PostProcessor
    g = np.load(path)
    pp = PostProcessor(device=0, max_batch=1)
    check_prep(pp.prepare_images(prep_image(g), 64, 128, True).cpu().numpy(), g)


@pytest.mark.gpu
def test_gpu_prepare_image_batch_and_options():
    import torch
    from improved_body_parts_b200.pipeline import PostProcessor
    pp = PostProcessor(device=0, max_batch=1)
    rng = np.random.default_rng(9)
    imgs = rng.integers(0, 256, (3, 250, 333, 3), dtype=np.uint8)
    got = pp.prepare_images(torch.from_numpy(imgs).cuda(), 64, 128, True).cpu().numpy()
    want = np.concatenate([O.prepare_input(imgs[b], 64, 128, True) for b in range(3)], axis=0)
    assert got.shape == want.shape and np.array_equal(got, want)
    got = pp.prepare_images(imgs[1], 32, 7, flip=False).cpu().numpy()
    assert np.array_equal(got, O.prepare_input(imgs[1], 32, 7, flip=False))
    with pytest.raises(ValueError):
        pp.prepare_images(np.zeros((4, 4, 3), dtype=np.float32))


@pytest.mark.gpu
def test_gpu_predict_refactor_mirror_prepares_on_device():
    """The predict_refactor mirror hands the model the reference's tensor (golden) and returns the flip average."""
    import torch
    from improved_body_parts_b200 import parse_skeletons as PS
    from improved_body_parts_b200.config import GetConfig, config_reader
    test_cfg, model_cfg = config_reader()
    g = np.load(PREP[0])
    image = prep_image(g)
    seen = {}
    rng = np.random.default_rng(3)

    def model(x):
        seen["x"] = x.cpu().numpy()
        seen["out"] = rng.uniform(0, 1, (2, 50, x.shape[1] // 4, x.shape[2] // 4)).astype(np.float32)
        return [[torch.from_numpy(seen["out"]).to(x.device)]]

    heat, paf = PS.predict_refactor(image, model, test_cfg, model_cfg, "synthetic", True, GetConfig("Canonical"))
    check_prep(seen["x"], g)
    want_h, want_p = O.fuse_flip(seen["out"])
    assert np.array_equal(heat, want_h) and np.array_equal(paf, want_p)


@pytest.mark.gpu
def test_gpu_fetch_coco_matches_oracle():
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    batch = synth.make_batch(8100, 6, people=(0, 9))
    batch[2] = 0  # an image without people: no rows, offsets still advance
    pp = PostProcessor(device=0, max_batch=8)
    res = pp(batch, 512)
    rows = pp.fetch_coco(6)
    assert len(rows) == 6 and rows[2].shape == (0, 52)
    total = 0
    for b in range(6):
        xy = np.stack([res.peaks[b]["x"], res.peaks[b]["y"]], axis=1) if res.n_peaks[b] else np.zeros((0, 2), np.int32)
        humans = [(rec["peak_id"], rec["score"]) for rec in res.humans[b]]
        want = O.coco_rows(xy, humans)
        assert rows[b].shape == want.shape and np.array_equal(rows[b], want)
        total += len(humans)
    assert total > 0
