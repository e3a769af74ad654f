"""The rotation_search branch (utils/parse_skeletons.py:215-218, :265-267, :63-66, :98-100; knob
utils/config:15).  Goldens come from the reference's own predict() / predict_refactor() run with a non-zero rotation
search (tests/golden/make_golden_rot.py): every tensor handed to the model (forward warp of the padded input + flipped twin)
and the returned maps (backward warps, fusion over scales x angles).  The oracle is held to them and to the live cv2 on the
CPU; the CUDA entry points (ibp_rotation_matrix, ibp_warp_affine, ibp_crop_resize_accumulate) and the mirrors on the GPU.
Bit-exact."""
import glob
import hashlib
import os

import numpy as np
import pytest

from oracle import np_oracle as O
from test_resize import _image

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLD, "rot_*.npz")))
MODEL_CFG = dict(boxsize=512, padValue=128, stride=4, max_downsample=64)


def _case(path):
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_rot", os.path.join(GOLD, "make_golden_rot.py"))
    g = np.load(path)
    h, w = (int(v) for v in g["shape"])
    img = _image(int(g["seed"]), h, w)
    assert hashlib.sha256(img.tobytes()).hexdigest() == str(g["img_sha256"]), "input generator drifted"
    ss = None if g["scale_search"][0] < 0 else tuple(float(v) for v in g["scale_search"])
    angles = [float(v) for v in g["rotation_search"]]
    # canned network outputs, regenerated exactly like the generator does (not imported: it needs /root/reference)
    from improved_body_parts_b200 import synth
    seed, people = int(g["seed"]), int(g["people"])
    if ss is None:
        ph, pw = -(-h // 64) * 64, -(-w // 64) * 64
        rng = np.random.default_rng(seed)
        clean = synth.render_maps(synth.random_people(rng, people, w, h), ph, pw, 4)
        canned = {(ph, pw): np.stack([synth.add_noise(clean, rng, 0.01), synth.add_noise(synth.flip_twin(clean), rng, 0.01)])}
    else:
        shapes = synth.multiscale_shapes(h, w, ss)
        outs, _ = synth.make_multiscale_output(seed, people, h, w, ss)
        canned = {tuple(s[2]): o for s, o in zip(shapes, outs)}
    return str(g["fn"]), img, ss, angles, canned, g


def _check_input(x, g, k):
    assert tuple(x.shape) == tuple(int(v) for v in g["in_shape_%d" % k])
    assert np.array_equal(x[:, ::23, ::19, :], g["in_probe_%d" % k])
    assert hashlib.sha256(np.ascontiguousarray(x).tobytes()).hexdigest() == str(g["in_sha256_%d" % k])


def _check_result(heat, paf, g):
    assert str(heat.dtype) == str(g["heat_dtype"])
    assert np.array_equal(heat[::11, ::13], g["heat_probe"]) and np.array_equal(paf[::11, ::13], g["paf_probe"])
    assert hashlib.sha256(np.ascontiguousarray(heat).tobytes()).hexdigest() == str(g["heat_sha256"])
    assert hashlib.sha256(np.ascontiguousarray(paf).tobytes()).hexdigest() == str(g["paf_sha256"])


def test_rotation_goldens_present():
    assert len(CASES) >= 4


def test_oracle_warp_matches_live_cv2():
    import cv2
    cv2.ipp.setUseIPP(False)
    rng = np.random.default_rng(3)
    for cn, (h, w), ang in [(3, (64, 128), 25.), (20, (96, 64), -33.), (30, (40, 72), 90.), (1, (33, 47), 7.3), (30, (128, 192), 180.),
                            (4, (17, 90), -2.5)]:
        s = rng.random((h, w, cn), dtype=np.float32)
        M = cv2.getRotationMatrix2D((h / 2, w / 2), ang, 1)
        assert np.array_equal(M, O.rotation_matrix(h / 2, w / 2, ang, 1))
        assert np.array_equal(O.warp_affine(s, M), cv2.warpAffine(s, M, (0, 0)).reshape(h, w, cn)), (cn, h, w, ang)


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_oracle_rotation_matches_reference(path):
    fn, img, ss, angles, canned, g = _case(path)
    h, w = img.shape[:2]
    mult = [1.0] if ss is None else [x * 512 / h for x in ss]
    k = 0
    if fn == "predict":
        items = []
        for m in mult:
            for a in angles:
                scaled = O.cv_resize_cubic_u8(img, m) if m != 1.0 else img
                x, _ = O.rotated_input(scaled, a)
                _check_input(x, g, k)
                k += 1
                items.append((canned[x.shape[1:3]], scaled.shape[:2], a))
        heat, paf = O.predict_fuse_rot(items, h, w)
    else:
        for a in angles:  # predict_refactor keeps the LAST item (l.101-102)
            x, m_rev = O.rotated_input(img, a)
            _check_input(x, g, k)
            k += 1
            heat, paf = O.fuse_flip(canned[x.shape[1:3]])
            if a != 0:
                heat, paf = O.warp_affine(heat, m_rev), O.warp_affine(paf, m_rev)
    assert k == int(g["n_calls"])
    _check_result(heat, paf, g)


@pytest.mark.gpu
def test_gpu_warp_affine_matches_oracle():
    import torch
    from improved_body_parts_b200 import _lib
    from improved_body_parts_b200.pipeline import PostProcessor
    pp = PostProcessor(device=0, max_batch=1)
    rng = np.random.default_rng(5)
    for cn, (h, w), ang in [(3, (64, 128), 25.), (20, (96, 64), -33.), (30, (40, 72), 90.), (50, (128, 176), 11.), (1, (33, 47), 7.3)]:
        s = rng.random((2, h, w, cn), dtype=np.float32)
        M = _lib.rotation_matrix(h / 2, w / 2, ang, 1)
        assert np.array_equal(M, O.rotation_matrix(h / 2, w / 2, ang, 1))
        want = np.stack([O.warp_affine(s[b], M) for b in range(2)])
        got = pp.warp_affine(torch.from_numpy(s).cuda(), M, "hwc").cpu().numpy()
        assert np.array_equal(got, want), (cn, h, w, ang)
        planar = np.ascontiguousarray(s.transpose(0, 3, 1, 2))
        got = pp.warp_affine(torch.from_numpy(planar).cuda(), M, "planar").cpu().numpy()
        assert np.array_equal(got, want.transpose(0, 3, 1, 2))


@pytest.mark.gpu
@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_gpu_mirrors_with_rotation_match_reference(path):
    import torch
    from improved_body_parts_b200 import parse_skeletons as PS
    fn, img, ss, angles, canned, g = _case(path)
    seen = []

    def model(x):
        seen.append(x.cpu().numpy())
        return [[torch.from_numpy(canned[(int(x.shape[1]), int(x.shape[2]))]).to(x.device)]]

    test_cfg = dict(thre1=0.1, thre2=0.1, offset_radius=2, rotation_search=angles, scale_search=list(ss) if ss else [1.0])
    if fn == "predict":
        heat, paf = PS.predict(img, model, test_cfg, MODEL_CFG, "synthetic", scale_search=ss)
    else:
        heat, paf = PS.predict_refactor(img, model, test_cfg, MODEL_CFG, "synthetic")
    assert len(seen) == int(g["n_calls"])
    for k, x in enumerate(seen):
        _check_input(x, g, k)
    _check_result(heat, paf, g)
