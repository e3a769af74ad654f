"""Host-side cost of one PipelinedPostProcessor.submit (no synchronisation inside the loop) against the GPU time per
batch, and of the bare C call (run on a GPU box: python tools/host_cost.py)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from improved_body_parts_b200 import synth  # noqa: E402
from improved_body_parts_b200.pipeline import PipelinedPostProcessor, PostProcessor  # noqa: E402

dev = torch.device("cuda", 0)
sets = [torch.from_numpy(synth.make_batch(2000 + 1000 * s, 64)).to(dev) for s in range(2)]
for depth in (1, 3):
    pipe = PipelinedPostProcessor(0, depth=depth, max_batch=64)
    for s in range(10):
        pipe.submit(sets[s % 2], 512, replace=True)
    torch.cuda.synchronize()
    n = 300
    t0 = time.perf_counter()
    for s in range(n):
        pipe.submit(sets[s % 2], 512, replace=True)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("depth %d: host %.1f us per submit, wall %.1f us per batch" % (depth, (t1 - t0) / n * 1e6, (t2 - t0) / n * 1e6))
pp = PostProcessor(0, max_batch=64)
st = torch.cuda.current_stream().cuda_stream
torch.cuda.synchronize()
t0 = time.perf_counter()
for s in range(300):
    pp.handle.postprocess_batch(sets[0].data_ptr(), False, 0, 64, 128, 128, 512, st)
t1 = time.perf_counter()
torch.cuda.synchronize()
print("bare ctypes call: host %.1f us per batch" % ((t1 - t0) / 300 * 1e6))
