"""TEST INFRASTRUCTURE ONLY — the reference's CPU post-processing route, timed as the baseline.

`postprocess_image` performs, for one image, exactly what evaluate.py:75-129 runs after the model call with
`--run_refactor --run_cpp`: flip average (parse_skeletons.py:80-103), heatmap_nms with scipy's maximum_filter and one
cv2.resize per peak (parse_skeletons.py:106-176), cv2.resize of the limb maps x4 (evaluate.py:77-80), joint-list
packing (evaluate.py:99-108), process_paf and the getter loop (evaluate.py:110-129).

/root/reference does not exist on the GPU box, so the Python half is this restatement issuing the SAME library
calls (scipy.ndimage.maximum_filter, cv2.resize INTER_CUBIC) and the C++ half is the UNMODIFIED reference
pafprocess.cpp prebuilt into oracle/_ref/libpafprocess_ref.so when available (else the C++ oracle port).
bench.py (cpu_baseline leg and --impl reference) is the only caller besides tests.
"""
import os

import numpy as np

from . import np_oracle as O

_state = {}


def _cpp():
    if "cpp" not in _state:
        from .cpu_lib import Oracle, Reference, reference_available
        if reference_available():
            _state["cpp"], _state["kind"] = Reference(), "reference"
        else:
            _state["cpp"], _state["kind"] = Oracle(), "port"
    return _state["cpp"]


def cpp_kind():
    _cpp()
    return _state["kind"]


def heatmap_nms_cv2(heatmaps, up=4):
    """parse_skeletons.py:126-176 with the reference's own library calls."""
    import cv2
    from scipy.ndimage import generate_binary_structure, maximum_filter
    cross = generate_binary_structure(2, 1)
    out, total = [], 0
    for joint in range(O.NUM_KEYPOINTS):
        m = heatmaps[:, :, joint]
        binary = (maximum_filter(m, footprint=cross) == m) * (m > 0.1)
        coords = np.array(np.nonzero(binary)[::-1]).T
        peaks = np.zeros((len(coords), 4), dtype=np.float32)
        for i, peak in enumerate(coords):
            x0, y0 = np.maximum(0, peak - 2)
            x1, y1 = np.minimum(np.array(m.T.shape) - 1, peak + 2)
            patch = m[y0:y1 + 1, x0:x1 + 1]
            big = cv2.resize(patch, None, fx=up, fy=up, interpolation=cv2.INTER_CUBIC)
            loc = np.unravel_index(big.argmax(), big.shape)
            center = (np.array(peak[::-1] - [y0, x0], dtype=np.float32) + 0.5) * up - 0.5
            refined = loc - center
            xy = (np.array(coords[i], dtype=np.float32) + 0.5) * up - 0.5 + refined[::-1]
            peaks[i, :] = (xy[0], xy[1], big[loc], total)
            total += 1
        out.append(peaks)
    return out


def postprocess_image(net_out, img_h=512, up=4):
    """net_out (2,50,h,w) -> [(keypoints {part: (x,y,score)}, human score)]"""
    import cv2
    heat, paf = O.fuse_flip(net_out)
    all_peaks = heatmap_nms_cv2(heat, up)
    paf_up = cv2.resize(paf, None, fx=up, fy=up, interpolation=cv2.INTER_CUBIC)
    jl = O.joint_list(all_peaks)
    humans = []
    if jl.shape[0] > 0:
        cpp = _cpp()
        cpp.process_paf(jl[None], paf_up, img_h)
        for h in range(cpp.get_num_humans()):
            parts = {}
            for part in range(18):
                pid = cpp.get_part_peak_id(h, part)
                if pid < 0:
                    continue
                parts[part] = (cpp.get_part_x(pid), cpp.get_part_y(pid), cpp.get_part_score(pid))
            if parts:
                humans.append((parts, cpp.get_score(h)))
    return humans


def predict_fuse_cv2(net_outs, img_h, img_w, scales, stride=4, flip_avg=True, interp=None):
    """parse_skeletons.py:232-281 (everything predict() does after the model calls) with the reference's own cv2.resize
    calls; net_outs[i] = (2,50,h_i,w_i) network output of scale scales[i].  -> heatmap_avg (H,W,20), paf_avg (H,W,30)
    float64.  Equal to np_oracle.predict_fuse bit for bit (tests/test_multiscale.py holds the two together)."""
    import cv2
    interp = cv2.INTER_CUBIC if interp is None else interp  # the reference's operator; INTER_LINEAR = the second operator mode
    heatmap_avg = np.zeros((img_h, img_w, O.NUM_HEATMAPS))
    paf_avg = np.zeros((img_h, img_w, O.NUM_PAFS))
    n = len(scales)
    for out, scale in zip(net_outs, scales):
        sh, sw = O.scaled_size(img_h, img_w, scale)
        heat, paf = O.fuse_flip(out, flip_avg)                                                    # l.232-248
        heat = cv2.resize(heat, (0, 0), fx=stride, fy=stride, interpolation=interp)               # l.252-256
        paf = cv2.resize(paf, (0, 0), fx=stride, fy=stride, interpolation=interp)                 # l.259-263
        heat, paf = heat[:sh, :sw, :], paf[:sh, :sw, :]                                           # l.272-273
        heat = cv2.resize(heat, (img_w, img_h), interpolation=interp)                             # l.276
        paf = cv2.resize(paf, (img_w, img_h), interpolation=interp)                               # l.277
        heatmap_avg = heatmap_avg + heat / n                                                      # l.280
        paf_avg = paf_avg + paf / n                                                               # l.281
    return heatmap_avg, paf_avg


def original_route_image(net_outs, img_h, img_w, scales):
    """The reference's ORIGINAL route for one image after the model calls (evaluate.py:82-90 without flags): predict's
    fusion (cv2), find_peaks (3x3 NMS + centroid refine; the reference does the NMS with torch on the GPU,
    util.py:177-185 — here numpy), find_connections, find_humans (pure Python/numpy, as in the reference).
    -> number of people."""
    if _shared.get("second_mode"):  # bilinear fusion + Gaussian-smoothed keypoint maps (5 taps, sigma 1), then the same steps
        import cv2
        heat, paf = predict_fuse_cv2(net_outs, img_h, img_w, scales, interp=cv2.INTER_LINEAR)
        heat = O.gaussian_smooth(heat[:, :, :18].astype(np.float32).transpose(2, 0, 1), 5, 1.0).transpose(1, 2, 0)
    else:
        heat, paf = predict_fuse_cv2(net_outs, img_h, img_w, scales)
    peaks = O.find_peaks(heat, 0.1, 2)
    cl, sp = O.find_connections(peaks, paf, img_h)
    assoc, _ = O.find_humans(cl, sp, peaks)
    return assoc.shape[0]


_shared = {}  # the sample is inherited by forked workers instead of being pickled to them


def _worker(args):
    import cv2
    cv2.setNumThreads(1)
    idx, img_h = args
    batch = _shared["batch"]
    return [len(postprocess_image(batch[i], img_h)) for i in idx]


def _worker_orig(args):
    import cv2
    cv2.setNumThreads(1)
    idx, img_h, img_w, scales = args
    cases = _shared["cases"]
    return [original_route_image(cases[i], img_h, img_w, scales) for i in idx]


def time_original_route(cases, img_h, img_w, scales, procs=None, second_mode=False):
    """Seconds to run the original route over `cases` (list over images of per-scale network outputs) once on `procs`
    host processes; returns (seconds, procs)."""
    import multiprocessing as mp
    import time
    procs = max(1, min(procs or os.cpu_count() or 1, len(cases)))
    _shared["cases"] = cases
    _shared["second_mode"] = bool(second_mode)
    if procs == 1:
        t0 = time.perf_counter()
        _worker_orig((list(range(len(cases))), img_h, img_w, scales))
        return time.perf_counter() - t0, 1
    ctx = mp.get_context("fork")
    shards = [(list(range(r, len(cases), procs)), img_h, img_w, scales) for r in range(procs)]
    with ctx.Pool(procs) as pool:
        pool.map(_worker_orig, [([0], img_h, img_w, scales)] * procs, chunksize=1)  # warm the workers
        t0 = time.perf_counter()
        pool.map(_worker_orig, shards, chunksize=1)
        dt = time.perf_counter() - t0
    return dt, procs


def time_batch(batch, img_h=512, procs=None, repeat=1):
    """Seconds to post-process `batch` (n,2,50,h,w) once on `procs` host processes (default: all cores), images
    sharded evenly; returns (seconds, procs)."""
    import multiprocessing as mp
    import time
    procs = procs or os.cpu_count() or 1
    n = batch.shape[0]
    procs = max(1, min(procs, n))
    _shared["batch"] = batch
    if procs == 1:
        t0 = time.perf_counter()
        for _ in range(repeat):
            _worker((list(range(n)), img_h))
        return (time.perf_counter() - t0) / repeat, 1
    ctx = mp.get_context("fork")
    shards = [(list(range(r, n, procs)), img_h) for r in range(procs)]
    with ctx.Pool(procs) as pool:
        pool.map(_worker, [([0], img_h)] * procs, chunksize=1)  # warm the workers (imports, page-in)
        t0 = time.perf_counter()
        for _ in range(repeat):
            pool.map(_worker, shards, chunksize=1)
        dt = (time.perf_counter() - t0) / repeat
    return dt, procs
