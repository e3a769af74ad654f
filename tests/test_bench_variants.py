"""Parity tests on the exact kernel variants and handle configurations bench.py times (VERDICT r1 "what's weak" 1-2):
the crowd handle (limbs_kernel<float,512> + one-shot peak kernel), the float16 pipeline on whole batches, configs[2] at
batch > 1 through the two-lane path, the 64-image pinned-host configuration — all against the oracle, the crowd
images against the unmodified reference library as well."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CROWD_LIMITS = dict(max_peaks_per_part=256, max_candidates_per_limb=1024, max_bodies=512, max_humans=256)  # bench.py Refactored(crowd)


def _oracle_image(oracle, out, img_h=512):
    from oracle import np_oracle as O
    heat, paf = O.fuse_flip(out)
    peaks = O.heatmap_nms(heat, 4)
    jl = O.joint_list(peaks)
    oracle.process_paf_feat(jl[None], np.ascontiguousarray(paf.transpose(2, 0, 1)), 4, img_h)
    return paf, jl, oracle.humans()


def _check(res, b, jl, humans, tag=""):
    assert res.n_peaks[b] == jl.shape[0], (tag, b)
    got = res.peaks[b]
    assert np.array_equal(got["x"], jl[:, 0].astype(np.int32)), (tag, b)
    assert np.array_equal(got["y"], jl[:, 1].astype(np.int32)), (tag, b)
    assert np.array_equal(got["score"], jl[:, 2]), (tag, b)
    assert np.array_equal(got["part"], jl[:, 4].astype(np.int32)), (tag, b)
    assert res.n_humans[b] == len(humans), (tag, b)
    for rec, (ids, score) in zip(res.humans[b], humans):
        assert np.array_equal(rec["peak_id"], ids), (tag, b)
        assert rec["score"] == score, (tag, b)
    assert res.status[b] & 0x4F == 0, (tag, b)


def test_crowd_bench_handle_exact(oracle):
    """bench.py --workload configs3: handle limits of Refactored(crowd=True), the workload's own images (seed 3000..,
    40-60 people, scale 0.2-0.45), three batches in flight.  First test that executes limbs_kernel<float,512>."""
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PipelinedPostProcessor
    from oracle import np_oracle as O
    from oracle.cpu_lib import Reference, reference_available
    n = 16
    batch = synth.make_batch(3000, n, people=(40, 60), height=512, width=512, scale=(0.2, 0.45))
    pipe = PipelinedPostProcessor(device=0, depth=3, max_batch=n, **CROWD_LIMITS)
    for _ in range(3):
        pipe.submit(batch, 512)
    results = pipe.drain()
    assert pipe.lanes[0].handle.limits.max_peaks_per_part == 256
    ref = Reference() if reference_available() else None
    tie_emulated = 0
    for b in range(n):
        paf, jl, humans = _oracle_image(oracle, batch[b])
        assert len(humans) >= 20
        for k, res in enumerate(results):
            _check(res, b, jl, humans, "lane %d" % k)
            tie_emulated += int(res.status[b] & 0x10 != 0)
        if ref is not None and b < 6:
            ref.process_paf(jl[None], O.limb_upsample(paf, 4), 512)
            assert ref.get_num_humans() == len(humans)
            for k, (ids, score) in enumerate(humans):
                assert [ref.get_part_peak_id(k, p) for p in range(18)] == list(ids)
                assert np.float32(ref.get_score(k)) == np.float32(score)
    assert pipe.launches >= 3 * 2


def test_fp16_batches_exact(oracle):
    """float16 network output (the reference's amp O1 default, parse_skeletons.py:75-80) on whole batches through the
    default handle — the path bench.py's e2e_f16 leg times — plus a batch of an odd map size (rows that are not 16-byte
    multiples in either element type)."""
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PipelinedPostProcessor, PostProcessor
    n = 24
    batch16 = synth.make_batch(2000, n, people=(1, 10), height=512, width=512).astype(np.float16)
    batch16[5] = 0
    pipe = PipelinedPostProcessor(device=0, depth=3, max_batch=n)
    for _ in range(3):
        pipe.submit(batch16, 512)
    results = pipe.drain()
    total = 0
    for b in range(n):
        _, jl, humans = _oracle_image(oracle, batch16[b])
        total += len(humans)
        for k, res in enumerate(results):
            _check(res, b, jl, humans, "f16 lane %d" % k)
    assert total > n
    # odd shape: 46 x 62 maps (184 x 248 input)
    pp = PostProcessor(device=0, max_batch=4)
    odd = np.stack([synth.make_network_output(650 + b, 2 + b, height=184, width=248)[0] for b in range(4)]).astype(np.float16)
    res = pp(odd, 184)
    for b in range(4):
        _, jl, humans = _oracle_image(oracle, odd[b], 184)
        _check(res, b, jl, humans, "f16 odd")
    # float16 and float32 hand-over of the same values agree wherever the half-precision average is exact: values
    # quantised to multiples of 2^-8 in [0, 1] (every sum, and half of it, is representable in float16)
    q = (np.round(synth.make_batch(777, 3, people=(3, 6)) * 256.0) / 256.0).astype(np.float32)
    a = pp(q, 512)
    b16 = pp(q.astype(np.float16), 512)
    for i in range(3):
        assert np.array_equal(a.peaks[i], b16.peaks[i]) and np.array_equal(a.humans[i], b16.humans[i])
    assert a.n_humans.sum() >= 3


def test_bench_configs1_batch64_pinned_host_exact(oracle):
    """The configuration bench.py's e2e leg runs: batch 64, default limits, three lanes, PINNED host input as the very
    first call of every lane (the staging buffer is allocated and zero-filled inside that call) — 2 x 148 persistent CTAs,
    64 x 31 queue items — every image against the oracle."""
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PipelinedPostProcessor
    n = 64
    batch = synth.make_batch(2000, n, people=(1, 10), height=512, width=512)
    host = torch.from_numpy(batch).pin_memory()
    pipe = PipelinedPostProcessor(device=0, depth=3, max_batch=n)
    for _ in range(3):
        pipe.submit(host, 512)          # first call of each lane: host input
    first = pipe.drain()
    dev = host.cuda()
    for _ in range(3):
        pipe.submit(dev, 512)
    second = pipe.drain()
    coco = None
    pipe.submit(host, 512)
    coco = pipe.collect(coco=True)
    from oracle import np_oracle as O
    for b in range(n):
        _, jl, humans = _oracle_image(oracle, batch[b])
        for k, res in enumerate(first + second):
            _check(res, b, jl, humans, "run %d" % k)
        want_rows = O.coco_rows(jl[:, :2].astype(np.int32), humans)
        assert coco[b].shape == want_rows.shape and np.array_equal(coco[b], want_rows)


@pytest.mark.parametrize("dtype", ["float32", "float16"])
@pytest.mark.parametrize("shape", [(512, 512), (384, 640)])
def test_pinned_host_zero_copy_equals_device_and_pageable(dtype, shape):
    """Round 2: from a PINNED host buffer only the keypoint planes are copied and the limb kernel pulls the box of rows x
    columns it needs straight out of the caller's buffer (include/ibp.h §3).  The same batch as a device tensor (whole rows
    staged from HBM) and as a pageable numpy array (staged copy of the 48 planes) must give the same bits, for both
    element types and a non-square map; people near the image border make boxes that start at column 0 / end at w."""
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    n = 12
    batch = synth.make_batch(2600, n, people=(1, 9), height=shape[0], width=shape[1])
    t = torch.from_numpy(batch)
    if dtype == "float16":
        t = t.half()
    pp = PostProcessor(device=0, max_batch=n)
    ref = pp(t.cuda(), shape[0])
    got_pinned = pp(t.pin_memory(), shape[0])
    got_pageable = pp(t.numpy().copy() if dtype == "float32" else t.clone(), shape[0])
    for got in (got_pinned, got_pageable):
        assert np.array_equal(got.n_peaks, ref.n_peaks) and np.array_equal(got.n_humans, ref.n_humans)
        for b in range(n):
            assert np.array_equal(got.peaks[b], ref.peaks[b])
            assert np.array_equal(got.humans[b], ref.humans[b])
    assert ref.n_humans.sum() >= n


def test_configs2_batch_two_lanes_exact():
    """bench.py --workload configs2: 640x480, flip + 4 scales, DISTINCT images at batch 8, two handles on two streams
    (MultiScale.step_dev), every image against the oracle chain.  The fusion oracle here issues the reference's own
    cv2.resize calls (oracle/cpu_path.predict_fuse_cv2, held equal to the numpy restatement by a CPU test)."""
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    from oracle import cpu_path
    from oracle import np_oracle as O
    img_h, img_w, ss = 480, 640, (0.5, 1.0, 1.5, 2.0)
    B = 8
    shapes = synth.multiscale_shapes(img_h, img_w, ss)
    cases = [synth.make_multiscale_output(2000 + i, 1 + (i * 3) % 12, img_h, img_w, ss)[0] for i in range(B)]
    dev = torch.device("cuda", 0)
    sets = [[torch.from_numpy(np.stack([c[s] for c in order])).to(dev) for s in range(len(ss))]
            for order in (cases, cases[::-1])]
    pps = [PostProcessor(device=0, max_batch=B) for _ in range(2)]
    streams = [torch.cuda.Stream(dev) for _ in range(2)]
    cur = torch.cuda.current_stream(dev)
    for rnd in range(2):                      # both lanes busy at the same time, twice
        for k in range(2):
            streams[k].wait_stream(cur)
            with torch.cuda.stream(streams[k]):
                pps[k].enqueue_original_route(sets[k], [s[1] for s in shapes], img_h, img_w)
    got = []
    for k in range(2):
        with torch.cuda.stream(streams[k]):
            got.append(pps[k].fetch_original_route(B))
    people = 0
    for b, outs in enumerate(cases):
        heat, paf = cpu_path.predict_fuse_cv2(outs, img_h, img_w, [s[0] for s in shapes])
        peaks = O.find_peaks(heat, 0.1, 2)
        table, counts = O.peaks_table(peaks)
        cl, sp = O.find_connections(peaks, paf, img_h)
        assoc, _ = O.find_humans(cl, sp, peaks)
        people += assoc.shape[0]
        for k, idx in ((0, b), (1, B - 1 - b)):
            t, c, a = got[k][idx]
            assert np.array_equal(c, counts), (k, b)
            assert np.array_equal(t[:, :3], table[:, :3]), (k, b)
            assert np.array_equal(a, assoc), (k, b)
    assert people >= B


def test_heatmap_nms_mirror_without_refine():
    """heatmap_nms(bool_refine_center=False) against goldens from the reference's own function (raw_8x.npz)."""
    from improved_body_parts_b200 import parse_skeletons as PS
    from test_oracle_golden import RAW, load_raw
    assert len(RAW) >= 4
    for path in RAW:
        heat, factor, rows = load_raw(path)
        got = PS.heatmap_nms(heat, factor, bool_refine_center=False)
        assert len(got) == 18 and all(g.dtype == np.float32 for g in got)
        flat = np.concatenate([np.concatenate([p, np.full((p.shape[0], 1), t, np.float32)], axis=1) for t, p in enumerate(got)])
        assert np.array_equal(flat, rows)


def test_mirrors_grow_capacity_on_crowds(oracle):
    """the single-image mirrors start with 128 peaks per part / 128 bodies; a crowd image must not raise
    IBP_ERR_CAPACITY, the handle is recreated with doubled limits and the call repeated."""
    from improved_body_parts_b200 import parse_skeletons as PS
    from improved_body_parts_b200 import synth
    from oracle import np_oracle as O
    out, _ = synth.make_network_output(9100, 150, scale=(0.08, 0.2))
    heat, _ = O.fuse_flip(out)
    want = O.heatmap_nms(heat, 4)
    assert max(len(p) for p in want) > 128
    PS._level[0] = 0
    got = PS.heatmap_nms(heat, 4)
    assert PS._level[0] >= 1
    for g, w in zip(got, want):
        assert np.array_equal(g, w)


def test_legacy_accepts_cuda_and_pinned_maps(oracle):
    """pafprocess.process_paf with the limb map as a CUDA tensor (sampled in place), as a pinned host tensor (DMA of the
    rows between the extreme peaks) and as a pageable array (threaded staging): same answers."""
    import torch
    from improved_body_parts_b200 import pafprocess, synth
    from oracle import np_oracle as O
    for seed, people in ((421, 3), (422, 14)):
        out, _ = synth.make_network_output(seed, people)
        heat, paf = O.fuse_flip(out)
        jl = O.joint_list(O.heatmap_nms(heat, 4))[None]
        up = O.limb_upsample(paf, 4)
        oracle.process_paf(jl, up, 512)
        want = oracle.humans()
        forms = [up, torch.from_numpy(up).pin_memory().numpy(), torch.from_numpy(up).cuda(), torch.from_numpy(up).cuda().double()]
        for m in forms:
            for peaks in (jl, torch.from_numpy(jl).cuda()):
                assert pafprocess.process_paf(peaks, m, 512) == 0
                got = [([pafprocess.get_part_peak_id(h, p) for p in range(18)], np.float32(pafprocess.get_score(h)))
                       for h in range(pafprocess.get_num_humans())]
                assert len(got) == len(want)
                for (gi, gs), (wi, ws) in zip(got, want):
                    assert gi == list(wi) and gs == ws
