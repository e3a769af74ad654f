"""Several batches in flight (pipeline.PipelinedPostProcessor): the same results as one batch at a time, in
submission order, for device and host inputs; lanes refuse to overwrite an uncollected result."""
import numpy as np
import pytest


def same(a, b):
    assert np.array_equal(a.n_peaks, b.n_peaks) and np.array_equal(a.n_humans, b.n_humans)
    for i in range(a.batch):
        assert np.array_equal(a.peaks[i], b.peaks[i])
        assert np.array_equal(a.humans[i]["peak_id"], b.humans[i]["peak_id"])
        assert np.array_equal(a.humans[i]["score"], b.humans[i]["score"])


@pytest.mark.gpu
def test_pipelined_equals_sequential():
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PipelinedPostProcessor, PostProcessor
    batches = [synth.make_batch(9000 + 10 * k, 4, people=(0, 8)) for k in range(7)]
    one = PostProcessor(device=0, max_batch=4)
    want = [one(b, 512) for b in batches]
    pipe = PipelinedPostProcessor(device=0, depth=3, max_batch=4)
    got = []
    for k, b in enumerate(batches):
        if pipe.full():
            got.append(pipe.collect())
        pipe.submit(torch.from_numpy(b).cuda() if k % 2 else b, 512)  # device and host inputs alternate
    got += pipe.drain()
    assert len(got) == len(want)
    for g, w in zip(got, want):
        same(g, w)
    # COCO rows through the same lanes
    pipe.submit(batches[0], 512)
    rows = pipe.collect(coco=True)
    assert sum(r.shape[0] for r in rows) == int(want[0].n_humans.sum())


@pytest.mark.gpu
def test_pipelined_lane_protection():
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PipelinedPostProcessor
    b = synth.make_batch(9100, 2, people=(1, 3))
    pipe = PipelinedPostProcessor(device=0, depth=2, max_batch=2)
    pipe.submit(b, 512)
    pipe.submit(b, 512)
    assert pipe.full()
    with pytest.raises(RuntimeError):
        pipe.submit(b, 512)
    pipe.submit(b, 512, replace=True)  # drops the oldest result of that lane
    assert len(pipe.drain()) == 2
    with pytest.raises(RuntimeError):
        pipe.collect()


def test_pipelined_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from improved_body_parts_b200 import _lib
    from improved_body_parts_b200.pipeline import PipelinedPostProcessor
    with pytest.raises(_lib.IbpError):
        PipelinedPostProcessor(device=0, depth=2)
