#!/usr/bin/env python
"""Experiments only: run one bench batch with IBP_DEBUG=8 and summarise the per-CTA %globaltimer stamps of the limb
kernel (start, staged, scored, done per (image, limb); grouping start/end per image)."""
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["IBP_DEBUG"] = str(24 | int(os.environ.get("IBP_DEBUG", "0")))
os.environ["IBP_TIMELINE_PK"] = "/tmp/ibp_timeline_pk.bin"
os.environ["IBP_TIMELINE"] = "/tmp/ibp_timeline.bin"
import torch
from improved_body_parts_b200 import synth
from improved_body_parts_b200.pipeline import PostProcessor

B = 64
CROWD = "--crowd" in sys.argv
if CROWD:
    batch = torch.from_numpy(synth.make_batch(3000, B, people=(40, 60), scale=(0.2, 0.45))).cuda()
    pp = PostProcessor(device=0, max_batch=B, max_peaks_per_part=256, max_candidates_per_limb=1024, max_bodies=512, max_humans=256)
else:
    batch = torch.from_numpy(synth.make_batch(2000, B, people=(1, 10))).cuda()
    pp = PostProcessor(device=0, max_batch=B)
for _ in range(3):
    res = pp(batch, 512)
t = np.fromfile("/tmp/ibp_timeline.bin", dtype=np.uint64).reshape(B, 31, 4).astype(np.float64)
t0 = t[:, :30, 0].min()
T = (t - t0) / 1000.0
limb = T[:, :30, :]
print("limb CTAs: start min/max %.1f %.1f us, end max %.1f" % (limb[:, :, 0].min(), limb[:, :, 0].max(), limb[:, :, 3].max()))
d = limb[:, :, 3] - limb[:, :, 0]
print("limb CTA duration mean %.2f max %.2f; stage %.2f score %.2f tail %.2f (means over CTAs with pairs)" % (
    d.mean(), d.max(), np.nanmean(np.where(t[:, :30, 1] > 0, limb[:, :, 1] - limb[:, :, 0], np.nan)),
    np.nanmean(np.where(t[:, :30, 2] > 0, limb[:, :, 2] - limb[:, :, 1], np.nan)),
    np.nanmean(np.where(t[:, :30, 2] > 0, limb[:, :, 3] - limb[:, :, 2], np.nan))))
g0, g1 = T[:, 30, 0], T[:, 30, 1]
links = np.array([int(h.shape[0]) for h in res.humans])
print("group: start min/max %.1f %.1f, end max %.1f, duration mean %.2f max %.2f" % (g0.min(), g0.max(), g1.max(), (g1 - g0).mean(), (g1 - g0).max()))
order = np.argsort(g1)[-8:]
for b in order:
    print("img %2d limb-done %.1f group %.1f..%.1f (%.1f us) peaks %d humans %d" % (b, limb[b, :, 3].max(), g0[b], g1[b], g1[b] - g0[b], res.n_peaks[b], res.n_humans[b]))

# ---- peak kernel ----
p = np.fromfile("/tmp/ibp_timeline_pk.bin", dtype=np.uint64).astype(np.float64)
p = p.reshape(B, 18, -1, 6)
p0 = p[..., 0].min()
P = (p - p0) / 1000.0
names = ["stage", "nms", "sort+refine", "wait warps", "ticket"]
st = np.sort(P[..., 0].ravel())
print("peak items started in the first microsecond: %d (= CTAs resident at launch)" % int((st < 1.0).sum()))
print("peak CTAs: start max %.1f, end max %.1f us; duration mean %.2f max %.2f" % (P[..., 0].max(), P[..., 5].max(), (P[..., 5] - P[..., 0]).mean(), (P[..., 5] - P[..., 0]).max()))
for i, nme in enumerate(names):
    d = P[..., i + 1] - P[..., i]
    print("  %-12s mean %.2f  p90 %.2f  max %.2f" % (nme, d.mean(), np.percentile(d, 90), d.max()))

# ---- streaming peak kernel: per-CTA view (items c, c + G, c + 2G, ... belong to CTA c) ----
G = int((st < 1.0).sum())
flat = P.reshape(-1, 6)
if G > 0 and flat.shape[0] > G:
    gaps, busy = [], []
    for c in range(G):
        it = flat[c::G]
        busy.append((it[:, 5] - it[:, 0]).sum())
        gaps.extend((it[1:, 0] - it[:-1, 5]).tolist())
    gaps = np.array(gaps)
    print("per CTA: %d items, busy mean %.1f us, gap between items mean %.2f p90 %.2f max %.2f" % (
        flat[0::G].shape[0], np.mean(busy), gaps.mean(), np.percentile(gaps, 90), gaps.max()))
    it = flat[0::G]
    for k in range(min(6, it.shape[0])):
        print("  CTA 0 item %d: " % k + " ".join("%.2f" % v for v in it[k]))
    first = flat[:G, 0]
    last_end = np.array([flat[c::G][-1, 5] for c in range(G)])
    print("first-item start: min %.1f p50 %.1f p90 %.1f max %.1f; last-item end: min %.1f p50 %.1f p90 %.1f max %.1f" % (
        first.min(), np.median(first), np.percentile(first, 90), first.max(), last_end.min(), np.median(last_end),
        np.percentile(last_end, 90), last_end.max()))
    late = np.argsort(last_end)[-3:]
    for c in late:
        it = flat[c::G]
        print("  CTA %d: " % c + " | ".join("%.1f-%.1f" % (v[0], v[5]) for v in it))
    c = int(np.argmax(last_end))
    it = flat[c::G]
    for k in range(min(5, it.shape[0])):
        print("  slow CTA %d item %d: " % (c, k) + " ".join("%.2f" % v for v in it[k]))
    slow = np.where(last_end > 1.5 * np.median(last_end))[0]
    print("slow CTAs (%d): %s" % (len(slow), slow.tolist()))
