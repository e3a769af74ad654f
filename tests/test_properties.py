"""Size-independent properties at BASELINE.json's full sizes (configs[1]: 64 images of 128x128x50 maps + flip twin;
configs[3]: 64 crowd images; configs[2]: 640x480, flip + 4 scales), where the oracle would take minutes: the product
is held against ITSELF across batch layouts (an image's result must not depend on its place in the batch, on its
neighbours, on the batch size or on what ran before), and its output against the invariants the reference's
algorithm guarantees (pafprocess.cpp:277-282: the filter; :290: the score; evaluate.py:99-108: ids in part order).  The small-size tests of
test_gpu_parity.py / test_gpu_golden.py tie the same kernels to the oracle bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CROWD_LIMITS = dict(max_peaks_per_part=256, max_candidates_per_limb=1024, max_bodies=512, max_humans=256)


def _same(a, i, b, j):
    assert a.n_peaks[i] == b.n_peaks[j] and a.n_humans[i] == b.n_humans[j], (i, j)
    assert np.array_equal(a.part_count[i], b.part_count[j]), (i, j)
    assert np.array_equal(a.peaks[i], b.peaks[j]), (i, j)
    assert np.array_equal(a.humans[i], b.humans[j]), (i, j)
    assert a.status[i] == b.status[j], (i, j)


@pytest.fixture(scope="module")
def batch64():
    from improved_body_parts_b200 import synth
    return synth.make_batch(2000, 64, people=(1, 10), height=512, width=512)  # bench.py's configs[1] images


@pytest.fixture(scope="module")
def result64(batch64):
    from improved_body_parts_b200.pipeline import PostProcessor
    pp = PostProcessor(device=0, max_batch=64)
    return pp, pp(batch64, 512)


def test_full_batch_is_deterministic_and_independent_of_history(batch64, result64):
    """The 64-image batch again after other batches ran on the same handle (work counters, tickets and scratch carry
    over from launch to launch): bit-identical."""
    pp, first = result64
    pp(batch64[:7], 512)
    pp(batch64[::-1].copy(), 512)
    again = pp(batch64, 512)
    for i in range(64):
        _same(first, i, again, i)
    assert int(first.n_humans.sum()) > 64


def test_result_does_not_depend_on_place_in_batch(batch64, result64):
    """A permutation of the batch permutes the results: the heaviest-first work queue, the band lists and the grouping
    tables are per image."""
    pp, first = result64
    perm = np.random.default_rng(5).permutation(64)
    res = pp(batch64[perm].copy(), 512)
    for k, i in enumerate(perm):
        _same(first, int(i), res, k)


def test_result_does_not_depend_on_batch_size(batch64, result64):
    """Every image alone (batch 1: a grid of 31 items) and in groups of 5 equals its result inside the batch of 64."""
    from improved_body_parts_b200.pipeline import PostProcessor
    pp, first = result64
    for i in range(0, 64, 3):
        _same(first, i, pp(batch64[i:i + 1], 512), 0)
    small = PostProcessor(device=0, max_batch=5)
    for i0 in range(0, 60, 15):
        res = small(batch64[i0:i0 + 5], 512)
        for k in range(5):
            _same(first, i0 + k, res, k)


def test_same_image_64_times(batch64, result64):
    """One image in all 64 places: 64 identical results (1984 queue items on 296 CTAs working on equal inputs: any
    cross-image interference or ordering dependence would show)."""
    pp, first = result64
    heavy = int(np.argmax(first.n_peaks))
    res = pp(np.repeat(batch64[heavy:heavy + 1], 64, axis=0), 512)
    for k in range(64):
        _same(first, heavy, res, k)


def _invariants(pp, res, b):
    n = int(res.n_peaks[b])
    pk = res.peaks[b][:n]
    # ids are arrival order, part-major (evaluate.py:99-108); the refined positions lie inside the image (:168-176)
    assert np.all(np.diff(pk["part"]) >= 0)
    assert np.array_equal(np.bincount(pk["part"], minlength=18), res.part_count[b])
    assert np.all((pk["x"] >= 0) & (pk["x"] < 512) & (pk["y"] >= 0) & (pk["y"] < 512))
    assert np.all(np.isfinite(pk["score"]))
    # every person passed the filter of pafprocess.cpp:277-282 and carries the score of :290
    par = pp.handle.params
    for h in res.humans[b][:int(res.n_humans[b])]:
        assert h["count"] >= par.min_parts
        assert h["score"] == np.float32(h["total"]) / np.float32(h["count"])
        assert h["score"] >= np.float32(par.min_human_score)
        assert (h["peak_id"] >= 0).sum() >= 2  # a person is at least one connection


def test_invariants_of_the_algorithm_full_batch(result64):
    pp, first = result64
    for b in range(64):
        _invariants(pp, first, b)


def test_crowd_batch_64_layout_independent():
    """configs[3] at its full size (64 images of 40-60 people, the crowd handle of bench.py): batch, reversed batch and
    single images agree; invariants hold."""
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    batch = synth.make_batch(3000, 64, people=(40, 60), height=512, width=512, scale=(0.2, 0.45))
    pp = PostProcessor(device=0, max_batch=64, **CROWD_LIMITS)
    first = pp(batch, 512)
    rev = pp(batch[::-1].copy(), 512)
    for i in range(64):
        _same(first, i, rev, 63 - i)
    for i in (0, 13, 63):
        _same(first, i, pp(batch[i:i + 1], 512), 0)
    for b in range(0, 64, 7):
        _invariants(pp, first, b)
    assert int(first.n_humans.min()) >= 10


def test_float16_full_batch_layout_independent(batch64):
    """The reference's amp default: the 64-image batch in float16, permuted and image by image."""
    import torch
    from improved_body_parts_b200.pipeline import PostProcessor
    pp = PostProcessor(device=0, max_batch=64)
    half = torch.from_numpy(batch64).cuda().half()
    first = pp(half, 512)
    perm = torch.from_numpy(np.random.default_rng(6).permutation(64)).cuda()
    res = pp(half[perm].contiguous(), 512)
    for k, i in enumerate(perm.cpu().numpy()):
        _same(first, int(i), res, k)
    for i in range(0, 64, 9):
        _same(first, i, pp(half[i:i + 1].contiguous(), 512), 0)


def test_multiscale_fusion_full_size_batch_independent():
    """configs[2] (640x480, flip + 4 scales): the fused float64 maps of an image do not depend on the batch it is in, and
    a second scale search made of the same map four times equals the single-scale result (x/4 four times, summed in
    float64, is x exactly: the power-of-two division and the accumulation lose nothing)."""
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    ss = (0.5, 1.0, 1.5, 2.0)
    shapes = synth.multiscale_shapes(480, 640, ss)
    maps = synth.make_multiscale_batch([(4000 + i, 2 + i) for i in range(4)], 480, 640, ss)
    pp = PostProcessor(device=0, max_batch=4)
    ts = [torch.from_numpy(m).cuda() for m in maps]
    fused, _ = pp.fuse_multiscale(ts, [s[1] for s in shapes], 480, 640, flip_avg=True)
    fused = fused.clone()
    for i in (0, 3):
        one, _ = pp.fuse_multiscale([t[i:i + 1].contiguous() for t in ts], [s[1] for s in shapes], 480, 640, flip_avg=True)
        assert torch.equal(one[0], fused[i])
    k = 1  # scale 1.0
    single, _ = pp.fuse_multiscale([ts[k][:1].contiguous()], [shapes[k][1]], 480, 640, flip_avg=True)
    single = single.clone()
    four, _ = pp.fuse_multiscale([ts[k][:1].contiguous()] * 4, [shapes[k][1]] * 4, 480, 640, flip_avg=True)
    assert torch.equal(four, single)
