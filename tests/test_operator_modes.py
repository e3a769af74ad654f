"""Second operator mode: the operators BASELINE.json's north_star names where the
reference's live path uses others — bilinear interpolation in predict()'s two resizes, and the separable Gaussian of
utils/util.py:103-174 (GaussianSmoothing, unused by the reference) in front of the 3x3 max-suppression.  Floating-point
mode with its own CPU oracle: tolerance 1e-4 relative (north_star) + 1e-6 absolute, written below.

CPU half: the oracle against cv2.resize(INTER_LINEAR) itself and against the goldens made with cv2 / the reference
class's own weights (tests/golden/make_golden_modes.py).  GPU half: ibp_fuse_multiscale_linear / ibp_smooth_gaussian
through the C ABI against the oracle and the goldens."""
import glob
import hashlib
import os

import numpy as np
import pytest

from oracle import np_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
GAUSS = sorted(glob.glob(os.path.join(HERE, "golden", "modes_gauss_*.npz")))
LINEAR = sorted(glob.glob(os.path.join(HERE, "golden", "modes_linear_*.npz")))
RTOL, ATOL = 1e-4, 1e-6  # the tolerance of this mode


def gauss_case(path):
    from improved_body_parts_b200 import synth
    g = np.load(path)
    seed, people, h, w, k = (int(v) for v in g["case"])
    out, _ = synth.make_network_output(seed, people, height=4 * h, width=4 * w)
    heat, _ = O.fuse_flip(out)
    x = np.ascontiguousarray(heat[:, :, :18].transpose(2, 0, 1))
    assert hashlib.sha256(x.tobytes()).hexdigest() == str(g["input_sha"]), "generator drifted"
    return x, k, float(g["sigma"]), g


def linear_case(path):
    from improved_body_parts_b200 import synth
    g = np.load(path)
    seed, people, ih, iw, flip = (int(v) for v in g["case"])
    ss = tuple(float(v) for v in g["scales"])
    outs, _ = synth.make_multiscale_output(seed, people, ih, iw, ss)
    assert hashlib.sha256(b"".join(o.tobytes() for o in outs)).hexdigest() == str(g["input_sha"]), "generator drifted"
    return outs, synth.multiscale_shapes(ih, iw, ss), ih, iw, bool(flip), g


def test_goldens_present():
    assert len(GAUSS) >= 3 and len(LINEAR) >= 3


def test_oracle_linear_resize_equals_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(7)
    for h, w, c, dh, dw in [(32, 40, 30, 128, 160), (128, 160, 20, 480, 640), (100, 130, 30, 77, 93), (17, 9, 20, 68, 36)]:
        src = rng.random((h, w, c), dtype=np.float32)
        want = cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR)
        np.testing.assert_allclose(O.cv_resize_linear(src, dh, dw), want, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("path", LINEAR, ids=os.path.basename)
def test_oracle_linear_fusion_matches_cv2_golden(path):
    outs, shapes, ih, iw, flip, g = linear_case(path)
    heat, paf = O.predict_fuse_linear(outs, ih, iw, [s[0] for s in shapes], flip_avg=flip)
    np.testing.assert_allclose(heat[::7, ::9], g["heat_probe"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(paf[::7, ::9], g["paf_probe"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("path", GAUSS, ids=os.path.basename)
def test_oracle_gaussian_matches_reference_class_golden(path):
    x, k, sigma, g = gauss_case(path)
    np.testing.assert_allclose(O.gaussian_smooth(x, k, sigma)[:, ::3, ::5], g["smoothed_probe"], rtol=1e-5, atol=1e-6)


def test_oracle_gaussian_kernel_is_normalised():
    for k, s in [(3, 0.8), (5, 1.0), (7, 1.5), (15, 3.0)]:
        g = O.gaussian_kernel1d(k, s)
        assert abs(float(g.sum()) - 1.0) < 1e-6 and np.allclose(g, g[::-1])


# ---------------------------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("path", LINEAR, ids=os.path.basename)
def test_gpu_fuse_multiscale_linear(path):
    import torch
    from improved_body_parts_b200.pipeline import PostProcessor
    outs, shapes, ih, iw, flip, g = linear_case(path)
    pp = PostProcessor(device=0, max_batch=2)
    # a batch of two: the image and a copy, so the batch index is exercised
    ts = [torch.from_numpy(np.stack([o, o])).cuda() for o in outs]
    fused, heat32 = pp.fuse_multiscale(ts, [s[1] for s in shapes], ih, iw, flip_avg=flip, interp="linear")
    fused = fused.cpu().numpy()
    assert np.array_equal(fused[0], fused[1])
    heat, paf = O.predict_fuse_linear(outs, ih, iw, [s[0] for s in shapes], flip_avg=flip)
    got_paf, got_heat = fused[0, :30].transpose(1, 2, 0), fused[0, 30:].transpose(1, 2, 0)
    np.testing.assert_allclose(got_heat, heat, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(got_paf, paf, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(got_heat[::7, ::9], g["heat_probe"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(got_paf[::7, ::9], g["paf_probe"], rtol=RTOL, atol=ATOL)
    np.testing.assert_array_equal(heat32.cpu().numpy()[0], fused[0, 30:48].astype(np.float32))


@pytest.mark.gpu
def test_gpu_fuse_multiscale_linear_odd_width():
    """An image whose width is not a multiple of the CTA's 128 columns and a chain split that does not divide the rows."""
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    ih, iw, ss = 203, 277, (0.7, 1.3)
    outs, _ = synth.make_multiscale_output(131, 4, ih, iw, ss)
    shapes = synth.multiscale_shapes(ih, iw, ss)
    pp = PostProcessor(device=0, max_batch=1)
    fused, _ = pp.fuse_multiscale([torch.from_numpy(o[None]).cuda() for o in outs], [s[1] for s in shapes], ih, iw, interp="linear")
    heat, paf = O.predict_fuse_linear(outs, ih, iw, [s[0] for s in shapes])
    fused = fused.cpu().numpy()[0]
    np.testing.assert_allclose(fused[30:].transpose(1, 2, 0), heat, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(fused[:30].transpose(1, 2, 0), paf, rtol=RTOL, atol=ATOL)


@pytest.mark.gpu
@pytest.mark.parametrize("path", GAUSS, ids=os.path.basename)
def test_gpu_smooth_gaussian(path):
    import torch
    from improved_body_parts_b200.pipeline import PostProcessor
    x, k, sigma, g = gauss_case(path)
    pp = PostProcessor(device=0, max_batch=1)
    got = pp.smooth_gaussian(torch.from_numpy(x[None]).cuda(), k, sigma).cpu().numpy()[0]
    np.testing.assert_allclose(got, O.gaussian_smooth(x, k, sigma), rtol=RTOL, atol=1e-5)
    np.testing.assert_allclose(got[:, ::3, ::5], g["smoothed_probe"], rtol=RTOL, atol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("shape,k,sigma", [((3, 96, 70), 5, 1.0), ((2, 50, 131), 3, 0.7), ((4, 77, 96), 11, 2.0), ((1, 200, 260), 9, 1.7)])
def test_gpu_smooth_gaussian_other_shapes(shape, k, sigma):
    """Row lengths that are not 16-byte multiples and kernel sizes outside 3..9 take the scalar variant; partial tiles."""
    import torch
    from improved_body_parts_b200.pipeline import PostProcessor
    x = np.random.default_rng(k).random(shape, dtype=np.float32)
    pp = PostProcessor(device=0, max_batch=1)
    got = pp.smooth_gaussian(torch.from_numpy(x[None]).cuda(), k, sigma).cpu().numpy()[0]
    np.testing.assert_allclose(got, O.gaussian_smooth(x, k, sigma), rtol=RTOL, atol=1e-5)


@pytest.mark.gpu
def test_gpu_second_mode_route_end_to_end():
    """Bilinear fusion -> Gaussian smoothing -> 3x3 max-suppression + centroid -> connections -> people, device resident,
    against the same chain of oracles (peak positions identical, scores to the mode's tolerance)."""
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    ih, iw, ss = 240, 320, (1.0, 2.0)
    outs, _ = synth.make_multiscale_output(141, 4, ih, iw, ss)
    shapes = synth.multiscale_shapes(ih, iw, ss)
    pp = PostProcessor(device=0, max_batch=1)
    ts = [torch.from_numpy(o[None]).cuda() for o in outs]
    B = pp.enqueue_original_route(ts, [s[1] for s in shapes], ih, iw, interp="linear", smooth=(5, 1.0))
    table_g, counts_g, people_g = pp.fetch_original_route(B)[0]
    heat, paf = O.predict_fuse_linear(outs, ih, iw, [s[0] for s in shapes])
    sm = O.gaussian_smooth(heat[:, :, :18].astype(np.float32).transpose(2, 0, 1), 5, 1.0).transpose(1, 2, 0)
    peaks = O.find_peaks(sm, 0.1, 2)
    table, counts = O.peaks_table(peaks)
    assert np.array_equal(counts_g, counts)
    np.testing.assert_allclose(table_g[:, :3], table[:, :3], rtol=1e-3, atol=1e-3)
    cl, sp = O.find_connections(peaks, paf, ih)
    assoc, _ = O.find_humans(cl, sp, peaks)
    assert people_g.shape == assoc.shape
    assert np.array_equal(people_g[:, :18, 0], assoc[:, :18, 0])  # the same peaks in the same people


@pytest.mark.gpu
def test_gpu_linear_fusion_is_linear_at_full_size():
    """Size-independent property at the configs[2] size (640x480, flip + 4 scales): the bilinear fusion is a linear
    operator, fuse(a x + b y) = a fuse(x) + b fuse(y) to the tolerance of this mode; a constant map fuses to that constant
    (every composed tap set sums to 1 / n_scales per scale)."""
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    ss = (0.5, 1.0, 1.5, 2.0)
    shapes = synth.multiscale_shapes(480, 640, ss)
    gen = torch.Generator(device="cuda").manual_seed(3)
    xs = [torch.rand((1, 2, 50) + tuple(s[3]), device="cuda", generator=gen) for s in shapes]
    ys = [torch.rand((1, 2, 50) + tuple(s[3]), device="cuda", generator=gen) for s in shapes]
    pp = PostProcessor(device=0, max_batch=1)
    sizes = [s[1] for s in shapes]
    fx = pp.fuse_multiscale(xs, sizes, 480, 640, flip_avg=True, interp="linear")[0].clone()
    fy = pp.fuse_multiscale(ys, sizes, 480, 640, flip_avg=True, interp="linear")[0].clone()
    fz = pp.fuse_multiscale([0.25 * x + 0.5 * y for x, y in zip(xs, ys)], sizes, 480, 640, flip_avg=True, interp="linear")[0].clone()
    np.testing.assert_allclose(fz.cpu().numpy(), (0.25 * fx + 0.5 * fy).cpu().numpy(), rtol=RTOL, atol=1e-5)
    ones = [torch.full_like(x, 0.375) for x in xs]
    fc = pp.fuse_multiscale(ones, sizes, 480, 640, flip_avg=True, interp="linear")[0]
    np.testing.assert_allclose(fc.cpu().numpy(), 0.375, rtol=RTOL, atol=ATOL)


@pytest.mark.gpu
def test_gpu_gaussian_properties_at_full_size():
    """18 planes of 640x480: a constant plane stays constant (the kernel sums to 1, reflect padding adds no mass), the
    operator is linear and commutes with mirroring the plane (symmetric kernel, reflect border)."""
    import torch
    from improved_body_parts_b200.pipeline import PostProcessor
    pp = PostProcessor(device=0, max_batch=1)
    gen = torch.Generator(device="cuda").manual_seed(4)
    x = torch.rand((1, 18, 480, 640), device="cuda", generator=gen)
    y = torch.rand((1, 18, 480, 640), device="cuda", generator=gen)
    for k, sigma in [(5, 1.0), (9, 2.0)]:
        sx = pp.smooth_gaussian(x, k, sigma).clone()
        sy = pp.smooth_gaussian(y, k, sigma).clone()
        sz = pp.smooth_gaussian(0.5 * x + 2.0 * y, k, sigma).clone()
        np.testing.assert_allclose(sz.cpu().numpy(), (0.5 * sx + 2.0 * sy).cpu().numpy(), rtol=RTOL, atol=1e-5)
        sm = pp.smooth_gaussian(torch.flip(x, dims=(3,)).contiguous(), k, sigma)
        np.testing.assert_allclose(torch.flip(sm, dims=(3,)).cpu().numpy(), sx.cpu().numpy(), rtol=RTOL, atol=1e-6)
        c = pp.smooth_gaussian(torch.full_like(x, 0.625), k, sigma)
        np.testing.assert_allclose(c.cpu().numpy(), 0.625, rtol=1e-5, atol=1e-6)
