"""Synthetic network outputs of the shape the SimplePose head emits (50 channels: 30 limb + 18 keypoint + 2
background, stride 4) for tests and benchmarks.  There is no checkpoint and no dataset in this environment, so
inputs are rendered from randomly placed skeletons the way the reference's label generator draws its training
targets (Gaussian keypoint blobs, Gaussian ridges along limbs — py_cocodata_server/py_data_heatmapper.py:105-244),
plus seeded noise and a "flip twin" so that the flip-fusion step has real work to do.

This is an input generator, not part of the measured path.
"""
import numpy as np

from .config import FLIP_HEAT_ORD, FLIP_PAF_ORD, LIMBS, NUM_LIMBS, NUM_PARTS

f32 = np.float32

# A standing figure, unit = body height; (x, y) per CMU-18 part index (config/config.py:60-62 ordering).
TEMPLATE = np.array([
    [0.00, 0.08],    # nose
    [0.00, 0.20],    # neck
    [-0.11, 0.21],   # Rsho
    [-0.15, 0.38],   # Relb
    [-0.17, 0.53],   # Rwri
    [0.11, 0.21],    # Lsho
    [0.15, 0.38],    # Lelb
    [0.17, 0.53],    # Lwri
    [-0.07, 0.53],   # Rhip
    [-0.08, 0.75],   # Rkne
    [-0.08, 0.97],   # Rank
    [0.07, 0.53],    # Lhip
    [0.08, 0.75],    # Lkne
    [0.08, 0.97],    # Lank
    [-0.025, 0.06],  # Reye
    [0.025, 0.06],   # Leye
    [-0.055, 0.08],  # Rear
    [0.055, 0.08],   # Lear
], dtype=np.float64)


def random_people(rng, n_people, width, height, scale=(0.35, 0.9), drop=0.1):
    """(n,18,3) float32 joints in image pixels; third column 1 = present, 2 = absent (the reference's convention,
    py_data_heatmapper.py:175)."""
    joints = np.zeros((n_people, NUM_PARTS, 3), dtype=f32)
    for p in range(n_people):
        s = rng.uniform(scale[0], scale[1]) * height
        ang = rng.normal(0.0, 0.15)
        ca, sa = np.cos(ang), np.sin(ang)
        t = TEMPLATE * [1.0, 1.0]
        t = t + rng.normal(0.0, 0.012, size=t.shape)  # pose jitter
        if rng.random() < 0.5:
            t[:, 0] = -t[:, 0]  # facing away: left/right swap sides
        xy = np.stack([ca * t[:, 0] - sa * (t[:, 1] - 0.5), sa * t[:, 0] + ca * (t[:, 1] - 0.5)], axis=1) * s
        cx = rng.uniform(0.08 * width, 0.92 * width)
        cy = rng.uniform(0.25 * height, 0.75 * height)
        joints[p, :, 0] = xy[:, 0] + cx
        joints[p, :, 1] = xy[:, 1] + cy
        present = rng.random(NUM_PARTS) >= drop
        inside = (joints[p, :, 0] >= 2) & (joints[p, :, 0] < width - 2) & (joints[p, :, 1] >= 2) & \
                 (joints[p, :, 1] < height - 2)
        joints[p, :, 2] = np.where(present & inside, 1.0, 2.0)
    return joints


def render_maps(joints, height, width, stride=4, sigma=9.0, limb_sigma=7.0, floor=0.015):
    """(50, height/stride, width/stride) float32 maps in [0,1] from joints in image pixels."""
    h, w = height // stride, width // stride
    maps = np.zeros((NUM_LIMBS + NUM_PARTS + 2, h, w), dtype=f32)
    gx = (np.arange(w) * stride + stride / 2 - 0.5).astype(f32)
    gy = (np.arange(h) * stride + stride / 2 - 0.5).astype(f32)
    two_s2 = f32(2.0 * sigma * sigma)
    reach = int(np.ceil(np.sqrt(-2.0 * sigma * sigma * np.log(floor)) / stride))
    # keypoint blobs, combined with max
    for part in range(NUM_PARTS):
        plane = maps[NUM_LIMBS + part]
        for p in range(joints.shape[0]):
            x, y, v = joints[p, part]
            if v >= 2:
                continue
            cx, cy = int(round(x / stride)), int(round(y / stride))
            x0, x1 = max(cx - reach, 0), min(cx + reach + 1, w)
            y0, y1 = max(cy - reach, 0), min(cy + reach + 1, h)
            if x0 >= x1 or y0 >= y1:
                continue
            ex = np.exp(-(gx[x0:x1] - f32(x)) ** 2 / two_s2)
            ey = np.exp(-(gy[y0:y1] - f32(y)) ** 2 / two_s2)
            np.maximum(plane[y0:y1, x0:x1], np.outer(ey, ex), out=plane[y0:y1, x0:x1])
    # limb ridges, averaged where several people overlap
    X, Y = np.meshgrid(gx, gy)
    for limb, (a, b) in enumerate(LIMBS):
        plane = maps[limb]
        cnt = np.zeros((h, w), dtype=f32)
        for p in range(joints.shape[0]):
            if joints[p, a, 2] >= 2 or joints[p, b, 2] >= 2:
                continue
            x1_, y1_ = joints[p, a, :2]
            x2_, y2_ = joints[p, b, :2]
            dx, dy = float(x2_ - x1_), float(y2_ - y1_)
            norm = np.hypot(dx, dy)
            if norm == 0:
                continue
            bx0 = max(int(round((min(x1_, x2_) - stride) / stride)), 0)
            by0 = max(int(round((min(y1_, y2_) - stride) / stride)), 0)
            bx1 = min(int(round((max(x1_, x2_) + stride) / stride)) + 1, w)
            by1 = min(int(round((max(y1_, y2_) + stride) / stride)) + 1, h)
            if bx0 >= bx1 or by0 >= by1:
                continue
            d = np.abs(dx * (y1_ - Y[by0:by1, bx0:bx1]) - (x1_ - X[by0:by1, bx0:bx1]) * dy) / (norm + 1e-6)
            g = np.exp(-d ** 2 / (2.0 * limb_sigma ** 2)).astype(f32)
            g[g <= floor] = 0.01
            plane[by0:by1, bx0:bx1] += g
            cnt[by0:by1, bx0:bx1] += 1
        nz = cnt > 0
        plane[nz] /= cnt[nz]
    maps[NUM_LIMBS + NUM_PARTS + 1] = maps[NUM_LIMBS:NUM_LIMBS + NUM_PARTS].max(axis=0)
    return np.clip(maps, 0.0, 1.0)


def flip_twin(maps):
    """What the network would emit for the mirrored image: mirrored maps with left/right channels exchanged, so
    that `twin[:, :, ::-1][perm] == maps` (the inverse of utils/parse_skeletons.py:92-93; both orders are
    involutions)."""
    perm = np.concatenate([FLIP_PAF_ORD, NUM_LIMBS + FLIP_HEAT_ORD])
    return np.ascontiguousarray(maps[perm][:, :, ::-1])


def add_noise(maps, rng, sigma=0.01):
    return np.clip(maps + rng.normal(0.0, sigma, size=maps.shape).astype(f32), 0.0, 1.0).astype(f32)


def make_network_output(seed, n_people, height=512, width=512, stride=4, noise=0.01, scale=(0.35, 0.9)):
    """One image worth of synthetic head output: (2, 50, h, w) float32 = [upright, flip twin], seeded."""
    rng = np.random.default_rng(seed)
    joints = random_people(rng, n_people, width, height, scale=scale)
    clean = render_maps(joints, height, width, stride)
    out = np.stack([add_noise(clean, rng, noise), add_noise(flip_twin(clean), rng, noise)])
    return out, joints


_shared_out = {}  # arrays in anonymous shared memory, inherited by forked workers (results are written in place, not pickled)


def _shared_array(shape):
    import mmap
    n = int(np.prod(shape)) * 4
    return np.frombuffer(mmap.mmap(-1, max(n, 1)), dtype=f32, count=int(np.prod(shape))).reshape(shape)


def _pool_run(fn, jobs, workers):
    """Run fn(job) for every job over forked worker processes (numpy only; call it before CUDA is initialised)."""
    import multiprocessing as mp
    import os
    workers = min(workers or (os.cpu_count() or 1), len(jobs))
    if workers <= 1:
        for j in jobs:
            fn(j)
        return
    with mp.get_context("fork").Pool(workers) as pool:
        pool.map(fn, jobs, chunksize=1)


def _batch_image(args):
    base_seed, i, people, height, width, stride, noise, scale = args
    rng = np.random.default_rng(base_seed + i)
    n = int(rng.integers(people[0], people[1] + 1))
    _shared_out["batch"][i] = make_network_output(base_seed + i, n, height, width, stride, noise, scale)[0]


def make_batch(base_seed, batch, people=(1, 10), height=512, width=512, stride=4, noise=0.01,
               scale=(0.35, 0.9), workers=1):
    """(B, 2, 50, h, w) float32 — seeds follow SURVEY.md §8(d): base_seed + image index.  workers != 1 renders the
    images in that many forked processes (0 = all cores); the result does not depend on it."""
    shape = (batch, 2, NUM_LIMBS + NUM_PARTS + 2, height // stride, width // stride)
    _shared_out["batch"] = _shared_array(shape) if workers != 1 else np.empty(shape, dtype=f32)
    _pool_run(_batch_image, [(base_seed, i, people, height, width, stride, noise, scale) for i in range(batch)], workers)
    return _shared_out.pop("batch")


def _multiscale_image(args):
    i, seed, n_people, img_h, img_w, scale_search = args
    for k, o in enumerate(make_multiscale_output(seed, n_people, img_h, img_w, scale_search)[0]):
        _shared_out["ms"][k][i] = o


def make_multiscale_batch(seeds_people, img_h=480, img_w=640, scale_search=(0.5, 1.0, 1.5, 2.0), workers=1):
    """[(seed, n_people)] -> list over scales of (B, 2, 50, h_s, w_s) float32 (make_multiscale_output per image)."""
    B = len(seeds_people)
    shapes = [(B, 2, NUM_LIMBS + NUM_PARTS + 2) + tuple(s[3]) for s in multiscale_shapes(img_h, img_w, scale_search)]
    _shared_out["ms"] = [_shared_array(s) if workers != 1 else np.empty(s, dtype=f32) for s in shapes]
    _pool_run(_multiscale_image, [(i, s, n, img_h, img_w, tuple(scale_search)) for i, (s, n) in enumerate(seeds_people)], workers)
    return _shared_out.pop("ms")


def multiscale_shapes(img_h, img_w, scale_search=(0.5, 1.0, 1.5, 2.0), boxsize=512, max_downsample=64, stride=4):
    """Per scale: (multiplier, scaled (h,w), padded (h,w), feature (h,w)) like utils/parse_skeletons.py:186,204-208."""
    out = []
    for x in scale_search:
        m = x * boxsize / img_h
        sh, sw = int(np.rint(img_h * m)), int(np.rint(img_w * m))
        ph, pw = -(-sh // max_downsample) * max_downsample, -(-sw // max_downsample) * max_downsample
        out.append((m, (sh, sw), (ph, pw), (ph // stride, pw // stride)))
    return out


def make_multiscale_output(seed, n_people, img_h=480, img_w=640, scale_search=(0.5, 1.0, 1.5, 2.0), noise=0.01,
                           scale=(0.35, 0.9)):
    """One image through the scale search: list over scales of (2, 50, h_s, w_s) float32 = [upright, flip twin],
    the same people drawn at each scale's padded input size (pad area empty), seeded."""
    rng = np.random.default_rng(seed)
    joints = random_people(rng, n_people, img_w, img_h, scale=scale)
    outs = []
    for m, _, (ph, pw), _ in multiscale_shapes(img_h, img_w, scale_search):
        j = joints.copy()
        j[:, :, :2] *= m
        clean = render_maps(j, ph, pw, 4)
        # the flipped input is the mirror of the PADDED image: padding ends up on the left of the twin
        outs.append(np.stack([add_noise(clean, rng, noise), add_noise(flip_twin(clean), rng, noise)]))
    return outs, joints
