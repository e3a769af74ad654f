"""CPU tests: the oracle (numpy + C++) against the golden vectors produced by running the reference itself
(tests/golden/make_golden.py), and against the unmodified reference library when it is present."""
import glob
import hashlib
import os

import numpy as np
import pytest

from oracle import np_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_case(path):
    """-> (net_out or None, heat (h,w,20), paf (h,w,30), golden dict)"""
    from improved_body_parts_b200 import synth
    g = np.load(path)
    name = os.path.basename(path)
    if name == "fixture.npz":
        maps = g["maps"]
        return None, np.ascontiguousarray(maps[30:50].transpose(1, 2, 0)), np.ascontiguousarray(maps[:30].transpose(1, 2, 0)), g
    if name.startswith("crowd"):
        rng = np.random.default_rng(int(g["seed"]))
        rng.uniform(0.25, 0.6, size=0)
        # replay make_golden.crowd_joints' draws so that the noise stream is at the same position
        for _ in range(int(g["people"])):
            rng.uniform(0.25, 0.6)
            rng.uniform(40, 470, size=2)
        clean = g["clean"]
        out = np.stack([synth.add_noise(clean, rng, 0.01), synth.add_noise(synth.flip_twin(clean), rng, 0.01)])
    else:
        out, _ = synth.make_network_output(int(g["seed"]), int(g["people"]))
    assert hashlib.sha256(out.tobytes()).hexdigest() == str(g["net_out_sha256"]), "input generator drifted"
    heat, paf = O.fuse_flip(out)
    return out, heat, paf, g


CASES = sorted(p for p in glob.glob(os.path.join(GOLD, "*.npz")) if os.path.basename(p).startswith(("fixture", "crowd", "synth_")))


def test_golden_files_present():
    assert len(CASES) >= 8


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_numpy_oracle_matches_reference_python(path):
    """heatmap_nms + joint list (parse_skeletons.py:126-176, evaluate.py:99-108) and the x4 limb resize."""
    _, heat, paf, g = load_case(path)
    jl = O.joint_list(O.heatmap_nms(heat, 4))
    assert np.array_equal(jl, g["joint_list"].reshape(-1, 5))
    up = O.limb_upsample(paf, 4)
    assert hashlib.sha256(up.tobytes()).hexdigest() == str(g["paf_up_sha256"])
    ys, xs = np.meshgrid(np.arange(0, up.shape[0], 37), np.arange(0, up.shape[1], 41), indexing="ij")
    assert np.array_equal(up[ys, xs, :], g["paf_up_probe"])


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_cpp_oracle_matches_reference_cpp(path, oracle):
    """process_paf (pafprocess.cpp:26-285): both limb sources of the oracle reproduce the reference's humans."""
    _, heat, paf, g = load_case(path)
    jl = g["joint_list"].reshape(-1, 5)
    if jl.shape[0] == 0:
        pytest.skip("no peaks")
    oracle.process_paf(jl[None], O.limb_upsample(paf, 4), 512)
    for run in (0, 1):
        hs = oracle.humans()
        assert len(hs) == g["human_ids"].shape[0]
        for (ids, score), gid, gsc in zip(hs, g["human_ids"], g["human_scores"]):
            assert np.array_equal(ids, gid)
            assert score == gsc
        if run == 0:
            oracle.process_paf_feat(jl[None], np.ascontiguousarray(paf.transpose(2, 0, 1)), 4, 512)


def test_fixture_values_from_survey():
    """The values SURVEY.md §0 / BASELINE.md §2 recorded for heatmap_test.npz."""
    g = np.load(os.path.join(GOLD, "fixture.npz"))
    jl = g["joint_list"].reshape(-1, 5)
    counts = [int(np.count_nonzero(jl[:, 4] == p)) for p in range(18)]
    assert counts == [1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 1, 1, 0, 1, 1, 0, 1]
    assert (jl[0, 0], jl[0, 1]) == (283, 152) and abs(jl[0, 2] - 1.0088363) < 1e-6
    assert g["human_ids"].tolist() == [[0, 1, 2, 3, 4, 5, 6, 7, 8, 9, -1, 10, 11, -1, 12, 13, -1, 14]]
    assert abs(float(g["human_scores"][0]) - 1.958247) < 1e-6


def test_cubic_sample_c_equals_numpy(oracle):
    rng = np.random.default_rng(5)
    plane = rng.random((9, 13), dtype=np.float32)
    up = O.cubic_upsample(plane, 4)
    for Y in range(up.shape[0]):
        for X in range(up.shape[1]):
            assert oracle.cubic_sample(plane, 4, X, Y) == up[Y, X]


def test_cubic_matches_cv2_native():
    """OpenCV's own cubic kernel (IPP off) bit for bit; the 30-channel call never uses IPP at all."""
    import cv2
    rng = np.random.default_rng(6)
    a = rng.random((24, 32, 30), dtype=np.float32)
    assert np.array_equal(O.cubic_upsample(a, 4), cv2.resize(a, None, fx=4, fy=4, interpolation=cv2.INTER_CUBIC))
    prev = cv2.ipp.useIPP()
    cv2.ipp.setUseIPP(False)
    try:
        for shape in [(5, 5), (4, 5), (5, 3), (3, 3), (40, 56)]:
            b = rng.random(shape, dtype=np.float32)
            assert np.array_equal(O.cubic_upsample(b, 4), cv2.resize(b, None, fx=4, fy=4, interpolation=cv2.INTER_CUBIC))
    finally:
        cv2.ipp.setUseIPP(prev)


def test_cross_nms_matches_scipy():
    from scipy.ndimage import generate_binary_structure, maximum_filter
    rng = np.random.default_rng(7)
    for _ in range(5):
        m = rng.random((17, 23), dtype=np.float32)
        m[rng.random(m.shape) < 0.3] = 0.5  # plateaus
        want = (maximum_filter(m, footprint=generate_binary_structure(2, 1)) == m) * (m > 0.1)
        assert np.array_equal(O.cross_nms(m), np.array(np.nonzero(want)[::-1]).T)


def test_cpu_path_equals_oracle(oracle):
    """The timed CPU baseline route computes the same skeletons as the oracle (scores to 1e-4: cv2 may use IPP)."""
    from improved_body_parts_b200 import synth
    from oracle import cpu_path
    out, _ = synth.make_network_output(77, 5)
    heat, paf = O.fuse_flip(out)
    jl = O.joint_list(O.heatmap_nms(heat, 4))
    oracle.process_paf_feat(jl[None], np.ascontiguousarray(paf.transpose(2, 0, 1)), 4, 512)
    want = oracle.humans()
    got = cpu_path.postprocess_image(out, 512)
    assert len(got) == len(want)
    for (parts, score), (ids, wscore) in zip(got, want):
        assert sorted(parts) == [p for p in range(18) if ids[p] >= 0]
        for p, (x, y, s) in parts.items():
            assert (x, y) == (int(jl[ids[p], 0]), int(jl[ids[p], 1]))
        assert abs(score - wscore) <= 1e-4 * abs(wscore)


RAW = sorted(glob.glob(os.path.join(GOLD, "raw_*.npz")))


def load_raw(path):
    from improved_body_parts_b200 import synth
    g = np.load(path)
    out, _ = synth.make_network_output(int(g["seed"]), int(g["people"]))
    assert hashlib.sha256(out.tobytes()).hexdigest() == str(g["net_out_sha256"]), "input generator drifted"
    return O.fuse_flip(out)[0], int(g["factor"]), g["rows"]


@pytest.mark.parametrize("path", RAW, ids=[os.path.basename(p) for p in RAW])
def test_oracle_heatmap_nms_without_refine(path):
    """heatmap_nms(bool_refine_center=False) (parse_skeletons.py:163-166): goldens from the reference's own function."""
    heat, factor, rows = load_raw(path)
    assert len(RAW) >= 4
    got = O.joint_list(O.heatmap_nms(heat, factor, bool_refine_center=False))
    assert np.array_equal(got, rows)


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_stock_cv2_with_ipp_gives_the_same_coordinates(path):
    """The goldens are generated with cv2.ipp.setUseIPP(False).  Stock cv2 — what the reference really
    runs, IPP's closed-source cubic kernel for 1-channel float patches — must give the SAME peak coordinates and ids;
    only score bits may differ, by far less than the 1e-4 tolerance of north_star.  oracle/cpu_path.heatmap_nms_cv2 is
    the reference's loop issuing the reference's cv2 / scipy calls with cv2's default (IPP on where the build has it)."""
    import cv2
    from oracle import cpu_path
    _, heat, _, g = load_case(path)
    cv2.ipp.setUseIPP(True)
    stock = O.joint_list(cpu_path.heatmap_nms_cv2(heat, 4))
    jl = g["joint_list"].reshape(-1, 5)
    assert stock.shape == jl.shape
    assert np.array_equal(stock[:, [0, 1, 3, 4]], jl[:, [0, 1, 3, 4]])          # x, y, id, part: identical
    assert np.allclose(stock[:, 2], jl[:, 2], rtol=1e-5, atol=0)                  # score: within 1e-5 relative
    differing = int(np.count_nonzero(stock[:, 2] != jl[:, 2]))
    assert differing <= max(int(g["ipp_peak_mismatch"]), 0) + jl.shape[0] // 2   # bookkeeping of make_golden.py stays plausible
