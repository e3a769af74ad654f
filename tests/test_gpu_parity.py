"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bit-exact for coordinates, ids and assignments; scores compared exactly as well (the tolerance of north_star is
1e-4 relative, the expectation is 0 ulp)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SEEDS = [(11, 1), (12, 4), (13, 10), (14, 25), (15, 50)]


@pytest.fixture(scope="module")
def pp():
    from improved_body_parts_b200.pipeline import PostProcessor
    return PostProcessor(device=0, max_batch=8, max_bodies=1024, max_humans=256, max_candidates_per_limb=4096)


def _oracle_image(oracle, out):
    from oracle import np_oracle as O
    heat, paf = O.fuse_flip(out)
    peaks = O.heatmap_nms(heat, 4)
    jl = O.joint_list(peaks)
    oracle.process_paf_feat(jl[None], np.ascontiguousarray(paf.transpose(2, 0, 1)), 4, 512)
    return heat, paf, peaks, jl, oracle.humans()


def test_fuse_flip_exact(pp):
    import torch
    from improved_body_parts_b200 import synth
    from oracle import np_oracle as O
    out, _ = synth.make_network_output(21, 6)
    fused = pp.fuse_flip(torch.from_numpy(out[None]).cuda())[0].cpu().numpy()
    heat, paf = O.fuse_flip(out)
    assert np.array_equal(fused[:30].transpose(1, 2, 0), paf)
    assert np.array_equal(fused[30:50].transpose(1, 2, 0), heat)
    # fp16 network output (amp O1): the average is taken in half precision like numpy does
    out16 = out.astype(np.float16)
    fused16 = pp.fuse_flip(torch.from_numpy(out16[None]).cuda())[0].cpu().numpy()
    heat16, paf16 = O.fuse_flip(out16)
    assert np.array_equal(fused16[:30].transpose(1, 2, 0), paf16)
    assert np.array_equal(fused16[30:50].transpose(1, 2, 0), heat16)


@pytest.mark.parametrize("seed,people", SEEDS)
def test_peaks_exact(pp, oracle, seed, people):
    import torch
    from improved_body_parts_b200 import synth
    out, _ = synth.make_network_output(seed, people)
    heat, paf, peaks, jl, _ = _oracle_image(oracle, out)
    maps = torch.from_numpy(np.ascontiguousarray(heat[:, :, :18].transpose(2, 0, 1))[None]).cuda()
    pp.find_peaks(maps, 0)
    res = pp.fetch(1)
    assert res.n_peaks[0] == jl.shape[0]
    assert np.array_equal(res.part_count[0], [len(p) for p in peaks])
    got = res.peaks[0]
    assert np.array_equal(got["x"], jl[:, 0].astype(np.int32))
    assert np.array_equal(got["y"], jl[:, 1].astype(np.int32))
    assert np.array_equal(got["score"], jl[:, 2])
    assert np.array_equal(got["part"], jl[:, 4].astype(np.int32))


@pytest.mark.parametrize("seed,people", SEEDS)
def test_pipeline_exact(pp, oracle, seed, people):
    from improved_body_parts_b200 import synth
    out, _ = synth.make_network_output(seed, people)
    _, _, _, jl, humans = _oracle_image(oracle, out)
    res = pp(out[None], 512)
    assert res.n_humans[0] == len(humans)
    for rec, (ids, score) in zip(res.humans[0], humans):
        assert np.array_equal(rec["peak_id"], ids)
        assert rec["score"] == score
    assert res.status[0] & 0xF == 0


@pytest.mark.parametrize("seed,people", SEEDS[:4])
def test_legacy_process_paf_exact(oracle, seed, people):
    """The seven-call protocol with HOST arrays in the reference layout (evaluate.py:106-129)."""
    from improved_body_parts_b200 import pafprocess, synth
    from oracle import np_oracle as O
    out, _ = synth.make_network_output(seed, people)
    heat, paf = O.fuse_flip(out)
    jl = O.joint_list(O.heatmap_nms(heat, 4))[None]
    up = O.limb_upsample(paf, 4)
    oracle.process_paf(jl, up, 512)
    assert pafprocess.process_paf(jl, up, 512) == 0
    assert pafprocess.get_num_humans() == oracle.get_num_humans()
    for h in range(oracle.get_num_humans()):
        for part in range(18):
            assert pafprocess.get_part_peak_id(h, part) == oracle.get_part_peak_id(h, part)
        assert np.float32(pafprocess.get_score(h)) == np.float32(oracle.get_score(h))
    for cid in range(jl.shape[1]):
        assert pafprocess.get_part_x(cid) == oracle.get_part_x(cid)
        assert pafprocess.get_part_y(cid) == oracle.get_part_y(cid)
        assert np.float32(pafprocess.get_part_score(cid)) == np.float32(oracle.get_part_score(cid))


def test_upsample_limbs_exact(pp):
    import torch
    from improved_body_parts_b200 import synth
    from oracle import np_oracle as O
    out, _ = synth.make_network_output(31, 8)
    _, paf = O.fuse_flip(out)
    t = torch.from_numpy(np.ascontiguousarray(paf.transpose(2, 0, 1))[None]).cuda()
    got = pp.upsample_limbs(t, 4)[0].cpu().numpy()
    assert np.array_equal(got, O.limb_upsample(paf, 4))


def test_streaming_kernels_default_limits(oracle):
    """A batch through a handle with DEFAULT limits — the configuration bench.py runs, which takes the persistent
    streaming kernels (peaks, band index, limbs) — against the oracle image by image: peaks, ids, skeletons, scores.
    Twice, with a different batch in between, so that the work counters and tickets carried from launch to launch are
    exercised too."""
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    B = 24
    pp = PostProcessor(device=0, max_batch=B)
    batch = np.stack([synth.make_network_output(400 + b, 1 + (b * 5) % 11)[0] for b in range(B)])
    batch[3] = 0.0                                   # an image without any peak
    batch[7, :, :30] = 0.0                           # peaks but no limb evidence: no connection at all
    want = [_oracle_image(oracle, batch[b]) for b in range(B)]
    other = np.stack([synth.make_network_output(900 + b, 3)[0] for b in range(5)])
    for rnd in range(2):
        res = pp(batch, 512)
        for b in range(B):
            _, _, peaks, jl, humans = want[b]
            assert res.n_peaks[b] == jl.shape[0], (rnd, b)
            assert np.array_equal(res.part_count[b], [len(p) for p in peaks])
            got = res.peaks[b]
            assert np.array_equal(got["x"], jl[:, 0].astype(np.int32))
            assert np.array_equal(got["y"], jl[:, 1].astype(np.int32))
            assert np.array_equal(got["score"], jl[:, 2])
            assert np.array_equal(got["part"], jl[:, 4].astype(np.int32))
            assert res.n_humans[b] == len(humans), (rnd, b)
            for rec, (ids, score) in zip(res.humans[b], humans):
                assert np.array_equal(rec["peak_id"], ids)
                assert rec["score"] == score
            assert res.status[b] & 0xF == 0
        small = pp(other, 512)                       # a different batch size in between
        assert small.n_humans.shape[0] == 5
    assert pp.handle.launches == 2 * (3 + 3)         # peaks + band index + limbs per call: the streaming route ran


def test_streaming_equals_one_shot_kernels_stress():
    """Race hunt: the persistent streaming kernels (default limits) against the one-CTA-per-item kernels (a handle whose
    limits rule the streaming route out) on random batches, everything compared bit for bit, several rounds back to back."""
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    B = 48
    fast = PostProcessor(device=0, max_batch=B)
    slow = PostProcessor(device=0, max_batch=B, max_bodies=1024, max_humans=256, max_candidates_per_limb=4096)
    for rnd in range(5):
        batch = synth.make_batch(7000 + 100 * rnd, B, people=(0, 12) if rnd % 2 else (1, 10))
        for rep in range(2):
            a = fast(batch, 512)
            b = slow(batch, 512)
            assert np.array_equal(a.n_peaks, b.n_peaks) and np.array_equal(a.part_count, b.part_count)
            assert np.array_equal(a.n_humans, b.n_humans), (rnd, rep)
            for i in range(B):
                assert np.array_equal(a.peaks[i], b.peaks[i]), (rnd, rep, i)
                assert np.array_equal(a.humans[i], b.humans[i]), (rnd, rep, i)
            assert not (a.status & 0x4F).any()
    assert fast.handle.launches == 3 * 10 and slow.handle.launches == 3 * 10


@pytest.mark.parametrize("h,w", [(96, 160), (160, 96), (64, 72), (40, 132), (184, 64), (128, 130)])
def test_streaming_kernels_other_map_sizes(oracle, h, w):
    """Default limits (streaming route where the rows are 16-byte multiples, the one-shot kernels for w = 130) on maps
    that are not 128x128: several bands per plane, a last band that is short, planes smaller than the ring."""
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    from oracle import np_oracle as O
    B = 3
    pp = PostProcessor(device=0, max_batch=B)
    outs = [synth.make_network_output(600 + h + b, 2 + 2 * b, height=4 * h, width=4 * w)[0] for b in range(B)]
    res = pp(np.stack(outs), 4 * h)
    for b, out in enumerate(outs):
        heat, paf = O.fuse_flip(out)
        jl = O.joint_list(O.heatmap_nms(heat, 4))
        oracle.process_paf_feat(jl[None], np.ascontiguousarray(paf.transpose(2, 0, 1)), 4, 4 * h)
        want = oracle.humans()
        assert res.n_peaks[b] == jl.shape[0]
        assert np.array_equal(res.peaks[b]["x"], jl[:, 0].astype(np.int32))
        assert np.array_equal(res.peaks[b]["y"], jl[:, 1].astype(np.int32))
        assert np.array_equal(res.peaks[b]["score"], jl[:, 2])
        assert res.n_humans[b] == len(want)
        for rec, (ids, score) in zip(res.humans[b], want):
            assert np.array_equal(rec["peak_id"], ids) and rec["score"] == score


def test_bench_workload_exact(oracle):
    """The batch bench.py times (BASELINE.json configs[1], seeds 2000.., default limits, three batches in flight): every
    image's peaks and skeletons equal the oracle's; with the unmodified reference library present, its getters as well."""
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PipelinedPostProcessor
    from oracle import np_oracle as O
    from oracle.cpu_lib import Reference, reference_available
    n = 24
    batch = synth.make_batch(2000, n, people=(1, 10), height=512, width=512)
    pipe = PipelinedPostProcessor(device=0, depth=3, max_batch=n)
    for _ in range(3):
        pipe.submit(batch, 512)
    results = pipe.drain()
    ref = Reference() if reference_available() else None
    total = 0
    for b in range(n):
        _, paf, _, jl, humans = _oracle_image(oracle, batch[b])
        for res in results:
            assert res.n_peaks[b] == jl.shape[0]
            assert np.array_equal(res.peaks[b]["x"], jl[:, 0].astype(np.int32))
            assert np.array_equal(res.peaks[b]["y"], jl[:, 1].astype(np.int32))
            assert np.array_equal(res.peaks[b]["score"], jl[:, 2])
            assert res.n_humans[b] == len(humans)
            for rec, (ids, score) in zip(res.humans[b], humans):
                assert np.array_equal(rec["peak_id"], ids) and rec["score"] == score
        if ref is not None and b < 8 and jl.shape[0]:
            ref.process_paf(jl[None], O.limb_upsample(paf, 4), 512)
            assert ref.get_num_humans() == len(humans)
            for k, (ids, score) in enumerate(humans):
                assert [ref.get_part_peak_id(k, p) for p in range(18)] == list(ids)
                assert np.float32(ref.get_score(k)) == np.float32(score)
        total += len(humans)
    assert total > n  # the workload has people in it


def test_crowded_limbs_on_default_limits(oracle):
    """Default limits, images with more than 16 people: limbs with more than 256 pairs take the one-thread-per-pair
    path of the streaming limb kernels (in limbs_stream3_kernel, csrc/k_limbs3.cuh: with the candidates in its
    global scratch); mixed in one batch with ordinary images, twice (scratch and counters are re-used), against the oracle."""
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    B = 10
    pp = PostProcessor(device=0, max_batch=B)
    people = [22, 3, 18, 26, 1, 20, 9, 24, 0, 17]
    batch = np.stack([synth.make_network_output(8100 + b, people[b])[0] for b in range(B)])
    want = [_oracle_image(oracle, batch[b]) for b in range(B)]
    crowded = 0
    for rnd in range(2):
        res = pp(batch, 512)
        for b in range(B):
            _, _, peaks, jl, humans = want[b]
            assert res.n_peaks[b] == jl.shape[0], (rnd, b)
            assert np.array_equal(res.peaks[b]["x"], jl[:, 0].astype(np.int32))
            assert np.array_equal(res.peaks[b]["score"], jl[:, 2])
            assert res.n_humans[b] == len(humans), (rnd, b)
            for rec, (ids, score) in zip(res.humans[b], humans):
                assert np.array_equal(rec["peak_id"], ids)
                assert rec["score"] == score
            assert res.status[b] & 0xF == 0
            n = [len(p) for p in peaks]
            crowded += any(n[a] * n[c] > 256 for a, c in zip(_limb_a(), _limb_b()))
    assert crowded >= 6  # the batch does reach the path this test is about
    assert pp.handle.launches == 2 * 3


def _limb_a():
    from improved_body_parts_b200 import config
    return [l[0] for l in config.LIMBS]


def _limb_b():
    from improved_body_parts_b200 import config
    return [l[1] for l in config.LIMBS]


def test_two_ctas_per_sm_limb_kernel_in_subprocess():
    """limbs_stream_kernel (two CTAs per SM) serves host sources; with IBP_LIMBS_NO_STREAM3 (read once per process) it
    takes the device-resident batches as well, which limbs_stream3_kernel (csrc/k_limbs3.cuh) handles by default: the
    streaming tests of this module once more on that kernel, in a process of its own."""
    import os
    import subprocess
    import sys
    if os.environ.get("IBP_LIMBS_NO_STREAM3"):
        pytest.skip("already inside the run without the three-CTA kernel")
    env = dict(os.environ, IBP_LIMBS_NO_STREAM3="1")
    res = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-k",
                          "streaming_kernels or stress or crowded_limbs or bench_workload"],
                         env=env, capture_output=True, text=True, timeout=1500)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert " passed" in res.stdout
