import os, sys
import numpy as np
sys.path.insert(0, "/root/repo")
os.environ["IBP_DEBUG"] = "8"
os.environ["IBP_TIMELINE"] = "/tmp/ibp_timeline.bin"
import torch
from improved_body_parts_b200 import synth
from improved_body_parts_b200.pipeline import PostProcessor
B = 64
batch = torch.from_numpy(synth.make_batch(2000, B, people=(1, 10))).cuda()
pp = PostProcessor(device=0, max_batch=B)
for _ in range(3):
    res = pp(batch, 512)
t = np.fromfile("/tmp/ibp_timeline.bin", dtype=np.uint64).reshape(B, 31, 4).astype(np.float64)
t0 = t[:, :30, 0].min()
T = (t - t0) / 1000.0
order = np.argsort(-res.n_peaks)
for b in list(order[:4]) + list(order[-3:]):
    print("img %d peaks %d: limb start min %.1f max %.1f, end max %.1f; group %.1f..%.1f" % (b, res.n_peaks[b], T[b, :30, 0].min(), T[b, :30, 0].max(), T[b, :30, 3].max(), T[b, 30, 0], T[b, 30, 1]))
b = order[0]
print("heaviest image per limb: start, staged, scored, end")
for l in range(0, 30, 3):
    print("  limb %d: " % l + " ".join("%.1f" % v for v in T[b, l]))

links = None
print("grouping: staging / chain / filter (us), by image weight")
for b in list(order[:5]) + list(order[30:33]) + list(order[-3:]):
    g = T[b, 30]
    print("  img %2d peaks %3d humans %2d: stage %.2f chain %.2f filter %.2f" % (b, res.n_peaks[b], res.n_humans[b], g[2] - g[0], g[3] - g[2], g[1] - g[3]))
