"""Multi-scale / flip fusion at image resolution (utils/parse_skeletons.py:180-283 predict): the numpy oracle against
the golden vectors produced by the reference's own predict() (tests/golden/make_golden_ms.py), and the CUDA kernel
(through the C ABI) against both.  Outputs are float64 and must be bit-exact."""
import glob
import hashlib
import os

import numpy as np
import pytest

from oracle import np_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLD, "ms_*.npz")))


def load_ms(path):
    from improved_body_parts_b200 import synth
    g = np.load(path)
    seed, people, img_h, img_w = int(g["seed"]), int(g["people"]), int(g["img_h"]), int(g["img_w"])
    ss = tuple(float(x) for x in g["scale_search"])
    if ss[0] < 0:  # the reference as shipped: one scale, multiplier 1.0
        ph, pw = O.padded_size(img_h, img_w)
        rng = np.random.default_rng(seed)
        joints = synth.random_people(rng, people, img_w, img_h)
        clean = synth.render_maps(joints, ph, pw, 4)
        outs = [np.stack([synth.add_noise(clean, rng, 0.01), synth.add_noise(synth.flip_twin(clean), rng, 0.01)])]
    else:
        outs, _ = synth.make_multiscale_output(seed, people, img_h, img_w, ss)
    assert hashlib.sha256(b"".join(o.tobytes() for o in outs)).hexdigest() == str(g["net_out_sha256"]), "generator drifted"
    return outs, [float(m) for m in g["multipliers"]], img_h, img_w, g


def check_against_golden(heat, paf, g):
    img_h, img_w = int(g["img_h"]), int(g["img_w"])
    ys, xs = np.meshgrid(np.arange(0, img_h, 29), np.arange(0, img_w, 31), indexing="ij")
    assert heat.dtype == np.float64 and paf.dtype == np.float64
    assert np.array_equal(heat[ys, xs, :], g["heat_probe"])
    assert np.array_equal(paf[ys, xs, :], g["paf_probe"])
    assert hashlib.sha256(np.ascontiguousarray(heat).tobytes()).hexdigest() == str(g["heat_sha256"])
    assert hashlib.sha256(np.ascontiguousarray(paf).tobytes()).hexdigest() == str(g["paf_sha256"])


def test_golden_present():
    assert len(CASES) >= 4


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_oracle_predict_fuse_matches_reference(path):
    outs, mult, img_h, img_w, g = load_ms(path)
    heat, paf = O.predict_fuse(outs, img_h, img_w, mult)
    check_against_golden(heat, paf, g)


def test_cv_resize_cubic_equals_cv2():
    cv2 = pytest.importorskip("cv2")
    cv2.ipp.setUseIPP(False)
    rng = np.random.default_rng(1)
    for (h, w, c, dh, dw) in [(64, 96, 20, 240, 320), (50, 37, 30, 33, 41), (31, 45, 3, 77, 35), (90, 121, 30, 60, 80)]:
        src = rng.uniform(0, 1, (h, w, c)).astype(np.float32)
        assert np.array_equal(cv2.resize(src, (dw, dh), interpolation=cv2.INTER_CUBIC), O.cv_resize_cubic(src, dh, dw))
    cv2.ipp.setUseIPP(True)


@pytest.mark.gpu
@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_gpu_fuse_multiscale_exact(path):
    import torch
    from improved_body_parts_b200.pipeline import PostProcessor
    outs, mult, img_h, img_w, g = load_ms(path)
    pp = PostProcessor(device=0, max_batch=2)
    sizes = [O.scaled_size(img_h, img_w, m) for m in mult]
    ts = [torch.from_numpy(np.stack([o, o])).cuda() for o in outs]  # batch of two identical images
    fused, heat32 = pp.fuse_multiscale(ts, sizes, img_h, img_w)
    fused = fused.cpu().numpy()
    for b in range(2):
        heat = np.ascontiguousarray(fused[b, 30:50].transpose(1, 2, 0))
        paf = np.ascontiguousarray(fused[b, :30].transpose(1, 2, 0))
        check_against_golden(heat, paf, g)
        assert np.array_equal(heat32[b].cpu().numpy(), fused[b, 30:48].astype(np.float32))


@pytest.mark.gpu
def test_gpu_predict_mirror_with_canned_model():
    """The predict() mirror end to end with a stand-in model (canned outputs), scale search on and off, odd sizes
    (row-length tail of the vertical pass: 427*30 and 333*... are not multiples of 4)."""
    import torch
    from improved_body_parts_b200 import parse_skeletons as PS, synth
    from improved_body_parts_b200.config import config_reader
    test_cfg, model_cfg = config_reader()
    # (170, 200): three scales = the division by a non-power of two; multiplier 512/170 = 3.01 = the kernel variant with
    # run-time row strides (wider tiles than the compiled constant)
    for img_h, img_w, ss in [(333, 427, (0.5, 1.0)), (250, 333, None), (170, 200, (0.4, 0.7, 1.0)), (170, 200, (1.0,))]:
        if ss is None:
            shapes = [(1.0, (img_h, img_w), O.padded_size(img_h, img_w), None)]
        else:
            shapes = synth.multiscale_shapes(img_h, img_w, ss)
        rng = np.random.default_rng(5)
        canned = {tuple(s[2]): rng.uniform(0, 1, (2, 50, s[2][0] // 4, s[2][1] // 4)).astype(np.float32) for s in shapes}

        def model(x):
            return [[torch.from_numpy(canned[(int(x.shape[1]), int(x.shape[2]))]).to(x.device)]]

        image = np.zeros((img_h, img_w, 3), dtype=np.uint8)
        heat, paf = PS.predict(image, model, test_cfg, model_cfg, "synthetic", scale_search=ss)
        want_h, want_p = O.predict_fuse([canned[tuple(s[2])] for s in shapes], img_h, img_w, [s[0] for s in shapes])
        assert np.array_equal(heat, want_h) and np.array_equal(paf, want_p)


@pytest.mark.gpu
def test_gpu_original_route_device_resident():
    """fuse_multiscale -> find_peaks -> find_connections -> find_humans without leaving the device, against the oracle
    chain (each link of which is pinned to the reference by its own golden vectors)."""
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    img_h, img_w, ss = 240, 320, (1.0, 2.0)
    shapes = synth.multiscale_shapes(img_h, img_w, ss)
    cases = [synth.make_multiscale_output(70 + b, 2 + 3 * b, img_h, img_w, ss)[0] for b in range(3)]
    pp = PostProcessor(device=0, max_batch=3)
    ts = [torch.from_numpy(np.stack([c[s] for c in cases])).cuda() for s in range(len(ss))]
    got = pp.original_route(ts, [s[1] for s in shapes], img_h, img_w)
    for b, outs in enumerate(cases):
        heat, paf = O.predict_fuse(outs, img_h, img_w, [s[0] for s in shapes])
        peaks = O.find_peaks(heat, 0.1, 2)
        table, counts = O.peaks_table(peaks)
        cl, sp = O.find_connections(peaks, paf, img_h)
        assoc, _ = O.find_humans(cl, sp, peaks)
        assert np.array_equal(got[b][1], counts)
        assert np.array_equal(got[b][0][:, :2], table[:, :2]) and np.array_equal(got[b][0][:, 2], table[:, 2])
        assert np.array_equal(got[b][2], assoc)
        assert assoc.shape[0] >= 1


@pytest.mark.gpu
def test_gpu_bench_workload_configs2_exact():
    """One image of the workload bench.py --workload configs2 times (640x480, flip + 4 scales, seed 2001): the
    four-kernel original route against the oracle, end to end."""
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    img_h, img_w, ss = 480, 640, (0.5, 1.0, 1.5, 2.0)
    shapes = synth.multiscale_shapes(img_h, img_w, ss)
    outs = synth.make_multiscale_output(2001, 4, img_h, img_w, ss)[0]
    pp = PostProcessor(device=0, max_batch=1)
    ts = [torch.from_numpy(o[None]).cuda() for o in outs]
    table_g, counts_g, assoc_g = pp.original_route(ts, [s[1] for s in shapes], img_h, img_w)[0]
    heat, paf = O.predict_fuse(outs, img_h, img_w, [s[0] for s in shapes])
    peaks = O.find_peaks(heat, 0.1, 2)
    table, counts = O.peaks_table(peaks)
    cl, sp = O.find_connections(peaks, paf, img_h)
    assoc, _ = O.find_humans(cl, sp, peaks)
    assert table.shape[0] > 10 and assoc.shape[0] >= 1
    assert np.array_equal(counts_g, counts) and np.array_equal(table_g[:, :3], table[:, :3])
    assert np.array_equal(assoc_g, assoc)


def test_cv2_fusion_route_equals_restatement():
    """oracle/cpu_path.predict_fuse_cv2 (the reference's own cv2.resize calls on 20/30-channel maps, what bench.py's CPU
    baseline of configs[2] times and what the batch parity test uses as its checker) == the numpy restatement that the
    goldens of the reference's predict pin (test_oracle_predict_fuse_matches_reference)."""
    from improved_body_parts_b200 import synth
    from oracle import cpu_path
    for seed, (img_h, img_w, ss) in ((4243, (240, 320, (1.0, 2.0))), (4244, (200, 150, (0.5, 1.5)))):
        shapes = synth.multiscale_shapes(img_h, img_w, ss)
        outs = synth.make_multiscale_output(seed, 3, img_h, img_w, ss)[0]
        a = O.predict_fuse(outs, img_h, img_w, [s[0] for s in shapes])
        b = cpu_path.predict_fuse_cv2(outs, img_h, img_w, [s[0] for s in shapes])
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


@pytest.mark.gpu
def test_gpu_strip_kernel_exact_in_subprocess():
    """The opt-in strip form of the fusion kernel (csrc/k_msstrip.cuh, IBP_MS_KERNEL=strip, read once per process) against
    the same goldens of the reference's predict and the configs2 bench image, in a process of its own."""
    import subprocess
    import sys
    if os.environ.get("IBP_MS_KERNEL") == "strip":
        pytest.skip("already inside the strip-kernel run")
    env = dict(os.environ, IBP_MS_KERNEL="strip")
    res = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-k",
                          "fuse_multiscale_exact or predict_mirror or configs2_exact or device_resident"],
                         env=env, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert " passed" in res.stdout
