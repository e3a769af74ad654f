"""The ORIGINAL route's peak step (utils/parse_skeletons.py:286-321 find_peaks: 3x3 NMS utils/util.py:177-185 +
centroid refinement utils/util.py:188-213): oracle against the golden vectors made by the reference's own functions
(tests/golden/make_golden_orig.py), and the CUDA path (peak kernel flavour IBP_PEAKS_BOX_CENTROID through the C ABI)
against both.  Positions are float64 and must be bit-exact; scores float32, bit-exact."""
import glob
import hashlib
import os

import numpy as np
import pytest

from oracle import np_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLD, "orig_*.npz")))
TEST_CFG = dict(thre1=0.1, offset_radius=2)  # utils/config:16,25


def _heatmaps(g):
    from improved_body_parts_b200 import synth
    out, _ = synth.make_network_output(int(g["seed"]), int(g["people"]), height=int(g["height"]), width=int(g["width"]))
    heat, _ = O.fuse_flip(out)
    hm = O.cubic_upsample(heat, 4)
    assert hashlib.sha256(hm.tobytes()).hexdigest() == str(g["heatmap_sha256"]), "input generator drifted"
    return hm


def test_golden_present():
    assert len(CASES) >= 5


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_oracle_find_peaks_matches_reference(path):
    g = np.load(path)
    table, counts = O.peaks_table(O.find_peaks(_heatmaps(g), TEST_CFG["thre1"], TEST_CFG["offset_radius"]))
    assert np.array_equal(counts, g["part_count"])
    assert np.array_equal(table, g["peaks"])


def test_box_nms_border_rule():
    """reflect padding never lets an out-of-image value win: a maximum on the edge or in a corner is kept."""
    a = np.zeros((1, 6, 7), dtype=np.float32)
    a[0, 0, 0], a[0, 5, 3], a[0, 2, 6], a[0, 3, 3] = 0.9, 0.8, 0.7, 0.1
    k = O.box_nms(a, 0.1)
    assert sorted(zip(*np.nonzero(k[0]))) == [(0, 0), (2, 6), (3, 3), (5, 3)]
    # plateaus: every pixel of a flat 2x2 top is kept (hmax == heat)
    b = np.zeros((1, 5, 5), dtype=np.float32)
    b[0, 1:3, 1:3] = 0.5
    assert np.count_nonzero(O.box_nms(b, 0.1)) == 4


@pytest.mark.gpu
@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_gpu_find_peaks_exact(path):
    from improved_body_parts_b200 import parse_skeletons as PS
    g = np.load(path)
    hm = _heatmaps(g)
    got = PS.find_peaks(hm, TEST_CFG)
    table, counts = O.peaks_table(got)
    assert np.array_equal(counts, g["part_count"])
    want = g["peaks"]
    assert np.array_equal(table[:, 3], want[:, 3])
    assert np.array_equal(table[:, :2], want[:, :2]), "sub-pixel positions differ (float64, bit-exact expected)"
    assert np.array_equal(table[:, 2].astype(np.float32), want[:, 2].astype(np.float32))


@pytest.mark.gpu
def test_gpu_find_peaks_border_and_plateau():
    from improved_body_parts_b200 import parse_skeletons as PS
    rng = np.random.default_rng(7)
    hm = np.zeros((40, 52, 20), dtype=np.float32)
    hm[..., :18] = rng.uniform(0.0, 0.09, size=(40, 52, 18)).astype(np.float32)
    hm[0, 0, 0], hm[39, 51, 1], hm[1, 20, 2], hm[20, 50, 3] = 0.9, 0.8, 0.7, 0.6   # corner / edge: unrefined
    hm[10:12, 10:12, 4] = 0.5                                                     # 2x2 plateau: four peaks
    hm[25, 25, 5] = 0.1                                                           # exactly thre1: kept (>=)
    got, cg = O.peaks_table(PS.find_peaks(hm, TEST_CFG))
    want, cw = O.peaks_table(O.find_peaks(hm, 0.1, 2))
    assert np.array_equal(cg, cw) and cw[4] == 4 and cw[5] == 1
    assert np.array_equal(got, want)
