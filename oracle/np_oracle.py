"""TEST INFRASTRUCTURE ONLY — numpy restatement of the Python half of the reference's hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
The product package never does (it fails loudly without its CUDA library instead).

Parity status: PINNED against the reference's own code run in the build container (tests/golden/make_golden.py
imports /root/reference/utils/parse_skeletons.py and freezes its outputs; tests/test_oracle_numpy.py replays them),
with ONE documented library switch: cv2's cubic resize is compared with `cv2.ipp.setUseIPP(False)`.  OpenCV's own
C++ cubic kernel is reproduced here bit-for-bit; the closed-source IPP kernel that opencv-python substitutes for
1/3/4-channel images is CPU-dispatch dependent and differs from OpenCV's own code by <= 1.2e-7 abs.
The 30-channel limb-map resize (evaluate.py:77-80) never goes through IPP and matches cv2 with no switch at all.

Every function cites the reference lines it restates (paths relative to /root/reference/).
"""
import numpy as np

f32 = np.float32

NUM_KEYPOINTS = 18  # utils/parse_skeletons.py:17
NUM_HEATMAPS = 20   # :18
NUM_PAFS = 30       # :19

# config/config.py:150-152
FLIP_HEAT_ORD = np.array([0, 1, 5, 6, 7, 2, 3, 4, 11, 12, 13, 8, 9, 10, 15, 14, 17, 16, 18, 19])
FLIP_PAF_ORD = np.array([0, 2, 1, 4, 3, 6, 5, 8, 7, 12, 13, 14, 9, 10, 11, 18, 19, 20, 15, 16, 17,
                         22, 21, 25, 26, 23, 24, 28, 27, 29])
# config/config.py:114-123 == utils/pafprocess/pafprocess.h:20-27
LIMBS = np.array([[1, 0], [1, 14], [1, 15], [1, 16], [1, 17], [0, 14], [0, 15], [14, 16], [15, 17], [1, 2],
                  [2, 3], [3, 4], [1, 5], [5, 6], [6, 7], [1, 8], [8, 9], [9, 10], [1, 11], [11, 12],
                  [12, 13], [0, 2], [0, 5], [2, 8], [8, 12], [5, 11], [11, 9], [16, 2], [17, 5], [8, 11]])
ORDER_COCO = [0, 15, 14, 17, 16, 5, 2, 6, 3, 7, 4, 11, 8, 12, 9, 13, 10]  # evaluate.py:40


# ----------------------------------------------------------------------------------------------------------
# A0  flip fusion — utils/parse_skeletons.py:80-103
# ----------------------------------------------------------------------------------------------------------
def fuse_flip(net_out, flip_avg=True):
    """net_out (2,50,h,w) in the model's dtype -> heatmaps (h,w,20), pafs (h,w,30), float32 HWC.

    The average runs in the input dtype (fp16 under amp O1, parse_skeletons.py:75) and only the result is cast
    (l.103)."""
    out = np.asarray(net_out)
    blob = out[0].transpose(1, 2, 0)
    paf, heat = blob[:, :, :NUM_PAFS], blob[:, :, NUM_PAFS:NUM_PAFS + NUM_HEATMAPS]
    if flip_avg:
        blob_f = out[1].transpose(1, 2, 0)
        paf_f, heat_f = blob_f[:, :, :NUM_PAFS], blob_f[:, :, NUM_PAFS:NUM_PAFS + NUM_HEATMAPS]
        paf = (paf + paf_f[:, ::-1, :][:, :, FLIP_PAF_ORD]) / 2
        heat = (heat + heat_f[:, ::-1, :][:, :, FLIP_HEAT_ORD]) / 2
    return heat.astype(f32), paf.astype(f32)


# ----------------------------------------------------------------------------------------------------------
# OpenCV-native cubic resize, integer up-scaling (cv2.resize(..., fx=up, fy=up, INTER_CUBIC), IPP off)
# ----------------------------------------------------------------------------------------------------------
def _cubic_weights(t):
    """OpenCV interpolateCubic, A = -0.75, float32 arithmetic."""
    t = f32(t)
    A = f32(-0.75)
    one = f32(1)
    w0 = ((A * (t + one) - f32(5) * A) * (t + one) + f32(8) * A) * (t + one) - f32(4) * A
    w1 = ((A + f32(2)) * t - (A + f32(3))) * t * t + one
    w2 = ((A + f32(2)) * (one - t) - (A + f32(3))) * (one - t) * (one - t) + one
    w3 = one - w0 - w1 - w2
    return np.array([w0, w1, w2, w3], dtype=f32)


def _cubic_taps(n_src, n_dst, scale):
    """per destination index: 4 clamped source indices and 4 float32 weights (OpenCV resize.cpp, cubic branch:
    fx = (float)((dx+0.5)*scale - 0.5); sx = floor(fx); fx -= sx; replicate border)."""
    idx = np.empty((n_dst, 4), dtype=np.int64)
    wts = np.empty((n_dst, 4), dtype=f32)
    for d in range(n_dst):
        fx = f32((d + 0.5) * scale - 0.5)
        sx = int(np.floor(fx))
        wts[d] = _cubic_weights(f32(fx - f32(sx)))
        idx[d] = np.clip(np.arange(sx - 1, sx + 3), 0, n_src - 1)
    return idx, wts


def cubic_upsample(src, up):
    """Bit-exact restatement of cv2.resize(src, None, fx=up, fy=up, interpolation=cv2.INTER_CUBIC) for float32
    input on OpenCV's own code path.  Horizontal pass: ((s0*a0 + s1*a1) + s2*a2) + s3*a3 (HResizeCubic).  Vertical
    pass: s0*b0 + (s1*b1 + (s2*b2 + s3*b3)) (VResizeCubicVec_32f; the 4-lane vector loop covers the whole row
    because the destination row length is a multiple of `up` = 4)."""
    s = np.asarray(src, dtype=f32)
    squeeze = s.ndim == 2
    if squeeze:
        s = s[:, :, None]
    h, w, _ = s.shape
    up = int(up)
    xi, xw = _cubic_taps(w, w * up, 1.0 / up)
    yi, yw = _cubic_taps(h, h * up, 1.0 / up)
    p = [s[:, xi[:, k], :] * xw[:, k][None, :, None] for k in range(4)]
    rows = ((p[0] + p[1]) + p[2]) + p[3]
    q = [rows[yi[:, k], :, :] * yw[:, k][:, None, None] for k in range(4)]
    out = q[0] + (q[1] + (q[2] + q[3]))
    if (w * up * s.shape[2]) % 4:
        raise NotImplementedError("row length not a multiple of 4: OpenCV's scalar tail uses another order")
    return out[:, :, 0] if squeeze else out


def limb_upsample(pafs, up=4):
    """A4 — evaluate.py:77-80 / demo_image.py:94-97."""
    return cubic_upsample(pafs, up)


# ----------------------------------------------------------------------------------------------------------
# A2  cross-footprint NMS — utils/parse_skeletons.py:106-119
# ----------------------------------------------------------------------------------------------------------
def cross_nms(img, thr=0.1):
    """maximum_filter(img, footprint=cross) == img, AND img > thr (strict, float32 compare).  scipy's default
    border 'reflect' repeats the edge pixel, so out-of-image neighbours never beat the centre.  Returns (n,2)
    int64 [x, y] in row-major (y, then x) order — the order of np.nonzero."""
    a = np.asarray(img)
    big = np.pad(a, 1, mode="edge")
    c = big[1:-1, 1:-1]
    keep = (c >= big[:-2, 1:-1]) & (c >= big[2:, 1:-1]) & (c >= big[1:-1, :-2]) & (c >= big[1:-1, 2:])
    keep &= a > a.dtype.type(thr)
    ys, xs = np.nonzero(keep)
    return np.stack([xs, ys], axis=1).astype(np.int64)


# ----------------------------------------------------------------------------------------------------------
# A3  per-peak patch refinement — utils/parse_skeletons.py:122-176
# ----------------------------------------------------------------------------------------------------------
def heatmap_nms(heatmaps, upsample_factor=4, bool_refine_center=True):
    """heatmaps (h,w,>=18) float32 -> list of 18 arrays (n,4) float32 [x, y, score, id]."""
    up = int(upsample_factor)
    win = 2
    out = []
    total = 0
    h, w = heatmaps.shape[:2]
    for joint in range(NUM_KEYPOINTS):
        m = np.ascontiguousarray(heatmaps[:, :, joint], dtype=f32)
        coords = cross_nms(m, 0.1)  # literal 0.1 at :139
        peaks = np.zeros((len(coords), 4), dtype=f32)
        for i, (px, py) in enumerate(coords):
            if bool_refine_center:
                x0, y0 = max(0, px - win), max(0, py - win)
                x1, y1 = min(w - 1, px + win), min(h - 1, py + win)
                patch_up = cubic_upsample(m[y0:y1 + 1, x0:x1 + 1], up)  # border = replicate of the PATCH
                flat = int(np.argmax(patch_up))  # first maximum in row-major order
                ry, rx = divmod(flat, patch_up.shape[1])
                # :159-172 reduces to integer arithmetic: (p+0.5)*up-0.5 + (r - ((p-p0+0.5)*up-0.5)) = p0*up + r
                peaks[i] = (f32(x0 * up + rx), f32(y0 * up + ry), patch_up[ry, rx], total)
            else:
                peaks[i] = ((f32(px) + f32(0.5)) * up - f32(0.5), (f32(py) + f32(0.5)) * up - f32(0.5),
                            m[py, px], total)
            total += 1
        out.append(peaks)
    return out


# ----------------------------------------------------------------------------------------------------------
# A5  joint list — evaluate.py:99-108
# ----------------------------------------------------------------------------------------------------------
def joint_list(all_peaks):
    rows = [tuple(p) + (t,) for t, peaks in enumerate(all_peaks) for p in peaks]
    if not rows:
        return np.zeros((0, 5), dtype=f32)
    return np.array(rows).astype(f32)


# ----------------------------------------------------------------------------------------------------------
# A12  3x3 NMS of the original route — utils/util.py:177-185 (keypoint_heatmap_nms)
# ----------------------------------------------------------------------------------------------------------
def box_nms(heat, thre=0.1):
    """heat (C,H,W) float32 -> heat * keep, keep = (3x3 max over the reflect-padded map == heat) * (heat >= thre).
    F.pad(mode='reflect') with pad 1 mirrors WITHOUT repeating the edge, i.e. it adds values that are already inside
    the 3x3 window, so the windowed maximum equals the maximum over the in-image part of the window."""
    a = np.asarray(heat, dtype=f32)
    big = np.pad(a, ((0, 0), (1, 1), (1, 1)), mode="reflect")
    H, W = a.shape[1:]
    hmax = np.full_like(a, -np.inf)
    for dy in range(3):
        for dx in range(3):
            hmax = np.maximum(hmax, big[:, dy:dy + H, dx:dx + W])
    keep = (hmax == a).astype(f32) * (a >= f32(thre)).astype(f32)
    return a * keep


# ----------------------------------------------------------------------------------------------------------
# A13  centroid refinement and the peak lists of the original route — utils/util.py:188-213,
#      utils/parse_skeletons.py:286-321 (find_peaks)
# ----------------------------------------------------------------------------------------------------------
def refine_centroid(scoremap, anchor, radius):
    """(x, y, score): border peaks are returned unchanged with the raw value (util.py:203-204); otherwise the float64
    first moments of the (2r+1)^2 box over its float32 sum — note that np.mgrid puts the ROW offset into x_grid
    (util.py:207), so x moves by the row moment and y by the column moment — and the box mean as the score."""
    x_c, y_c = anchor
    x_min, x_max, y_min, y_max = x_c - radius, x_c + radius + 1, y_c - radius, y_c + radius + 1
    if y_max > scoremap.shape[0] or y_min < 0 or x_max > scoremap.shape[1] or x_min < 0:
        return anchor + (scoremap[y_c, x_c],)
    box = scoremap[y_min:y_max, x_min:x_max]
    x_grid, y_grid = np.mgrid[-radius:radius + 1, -radius:radius + 1]
    offset_x = (box * x_grid).sum() / box.sum()
    offset_y = (box * y_grid).sum() / box.sum()
    return (x_c + offset_x, y_c + offset_y) + (box.mean(),)


def find_peaks(heatmap_avg, thre1=0.1, radius=2):
    """heatmap_avg (H,W,>=18) -> list of 18 lists of (x, y, score, id); x, y float64 (ints at the border)."""
    hm = np.asarray(heatmap_avg).astype(f32)  # parse_skeletons.py:291
    planes = np.ascontiguousarray(hm[:, :, :NUM_KEYPOINTS].transpose(2, 0, 1))
    filt = box_nms(planes, thre1)
    out, counter = [], 0
    for part in range(NUM_KEYPOINTS):
        map_orig = hm[:, :, part]
        ys, xs = np.nonzero(filt[part])  # row-major order (parse_skeletons.py:305)
        refined = [refine_centroid(map_orig, (x, y), radius) for x, y in zip(xs, ys)]
        out.append([refined[i] + (counter + i,) for i in range(len(refined))])
        counter += len(refined)
    return out


def peaks_table(all_peaks):
    """list of 18 lists of (x,y,score,id) -> (N,4) float64 + per-part counts, for comparisons and fixtures."""
    rows = [tuple(float(v) for v in p) for ps in all_peaks for p in ps]
    tab = np.array(rows, dtype=np.float64).reshape(-1, 4)
    return tab, np.array([len(ps) for ps in all_peaks], dtype=np.int64)


# ----------------------------------------------------------------------------------------------------------
# OpenCV-native cubic resize, arbitrary destination size (cv2.resize(src, (dw, dh), INTER_CUBIC), IPP off)
# ----------------------------------------------------------------------------------------------------------
def _cv_cubic_axis(n_src, n_dst, scale):
    """OpenCV resize.cpp, cubic branch, one axis: fx = (float)((d+0.5)*scale - 0.5) with `scale` a double;
    s = cvFloor(fx); fx -= s; four taps s-1..s+2 clamped to the source (replicate border)."""
    idx = np.empty((n_dst, 4), dtype=np.int64)
    wts = np.empty((n_dst, 4), dtype=f32)
    for d in range(n_dst):
        fx = f32((d + 0.5) * scale - 0.5)
        s = int(np.floor(fx))
        wts[d] = _cubic_weights(f32(fx - f32(s)))
        idx[d] = np.clip(np.arange(s - 1, s + 3), 0, n_src - 1)
    return idx, wts


def cv_resize_cubic(src, dst_h, dst_w, scale_y=None, scale_x=None):
    """Bit-exact restatement of OpenCV's own float32 cubic resize for an (H,W,C) image to (dst_h, dst_w).
    scale_* default to what cv2 derives from an explicit dsize: 1. / ((double)dst / src); pass 1./f for the
    fx=f, fy=f call form.  Horizontal pass: ((s0*a0 + s1*a1) + s2*a2) + s3*a3.  Vertical pass: the 4-lane vector
    loop (SSE baseline build) gives s0*b0 + (s1*b1 + (s2*b2 + s3*b3)) for the first 4*floor(W*C/4) elements of a
    row, the scalar tail ((s0*b0 + s1*b1) + s2*b2) + s3*b3 for the rest."""
    s = np.asarray(src, dtype=f32)
    squeeze = s.ndim == 2
    if squeeze:
        s = s[:, :, None]
    h, w, c = s.shape
    if scale_x is None:
        scale_x = 1.0 / (float(dst_w) / w)
    if scale_y is None:
        scale_y = 1.0 / (float(dst_h) / h)
    xi, xw = _cv_cubic_axis(w, dst_w, scale_x)
    yi, yw = _cv_cubic_axis(h, dst_h, scale_y)
    p = [s[:, xi[:, k], :] * xw[:, k][None, :, None] for k in range(4)]
    rows = ((p[0] + p[1]) + p[2]) + p[3]
    q = [rows[yi[:, k], :, :] * yw[:, k][:, None, None] for k in range(4)]
    out = q[0] + (q[1] + (q[2] + q[3]))
    n_vec = (dst_w * c) // 4 * 4
    if n_vec != dst_w * c:
        tail = ((q[0] + q[1]) + q[2]) + q[3]
        flat, flat_t = out.reshape(dst_h, dst_w * c), tail.reshape(dst_h, dst_w * c)
        flat[:, n_vec:] = flat_t[:, n_vec:]
        out = flat.reshape(dst_h, dst_w, c)
    return out[:, :, 0] if squeeze else out


# ----------------------------------------------------------------------------------------------------------
# A1  multi-scale / flip fusion at image resolution — utils/parse_skeletons.py:180-283 (predict)
# ----------------------------------------------------------------------------------------------------------
def scaled_size(img_h, img_w, scale):
    """Size of cv2.resize(image, (0,0), fx=scale, fy=scale) (l.204): saturate_cast<int>(n*scale) = round half to even."""
    return int(np.rint(img_h * scale)), int(np.rint(img_w * scale))


def padded_size(sh, sw, max_downsample=64):
    """util.padRightDownCorner (utils/util.py:44-65): pad bottom/right up to a multiple of max_downsample."""
    return -(-sh // max_downsample) * max_downsample, -(-sw // max_downsample) * max_downsample


def predict_fuse(net_outs, img_h, img_w, scales, stride=4, max_downsample=64, flip_avg=True):
    """Everything predict() does after the model calls.  net_outs[i] = (2,50,padH_i/stride,padW_i/stride) network
    output for scale scales[i] (the multipliers of l.186).  Returns heatmap_avg (H,W,20), paf_avg (H,W,30) float64."""
    heatmap_avg = np.zeros((img_h, img_w, NUM_HEATMAPS))
    paf_avg = np.zeros((img_h, img_w, NUM_PAFS))
    n = len(scales)  # len(multiplier) * len(rotate_angle), rotation_search = [0.0] (utils/config:15)
    for out, scale in zip(net_outs, scales):
        sh, sw = scaled_size(img_h, img_w, scale)
        pad_h, pad_w = padded_size(sh, sw, max_downsample)
        assert out.shape[2] * stride == pad_h and out.shape[3] * stride == pad_w, (out.shape, pad_h, pad_w)
        heat, paf = fuse_flip(out, flip_avg)                                              # l.232-248
        heat = cv_resize_cubic(heat, pad_h, pad_w, 1.0 / stride, 1.0 / stride)            # l.252-256
        paf = cv_resize_cubic(paf, pad_h, pad_w, 1.0 / stride, 1.0 / stride)              # l.259-263
        heat, paf = heat[:sh, :sw, :], paf[:sh, :sw, :]                                   # l.272-273
        heat = cv_resize_cubic(heat, img_h, img_w)                                        # l.276
        paf = cv_resize_cubic(paf, img_h, img_w)                                          # l.277
        heatmap_avg = heatmap_avg + heat / n                                              # l.280 (float32 / int, then + in float64)
        paf_avg = paf_avg + paf / n                                                       # l.281
    return heatmap_avg, paf_avg


# ----------------------------------------------------------------------------------------------------------
# A14  connections, Python semantics — utils/parse_skeletons.py:324-410 (find_connections)
# ----------------------------------------------------------------------------------------------------------
def find_connections(all_peaks, paf_avg, img_height, thre2=0.1, connect_ration=0.8, mid_num_max=20, pairs=LIMBS):
    """all_peaks: 18 lists of (x, y, score, id); paf_avg (H,W,30).  -> (connected_limbs, special_limb):
    connected_limbs[k] = [] when both joint lists are empty (k goes to special_limb), else (n,6) float64 rows
    [src_peak_id, dst_peak_id, score, i, j, length].  Differences from the C++ route: float peak
    positions, np.linspace + half-to-even rounding of the sample points, float64 arithmetic, stable sort."""
    connected, special = [], []
    for k, (pa, pb) in enumerate(pairs):
        src_list, dst_list = all_peaks[pa], all_peaks[pb]
        if len(src_list) == 0 and len(dst_list) == 0:           # l.339-341
            special.append(k)
            connected.append([])
            continue
        plane = paf_avg[:, :, k]
        cands = []
        for i, s in enumerate(src_list):
            s = np.array(s)                                     # mixed tuple -> float64 vector (l.347)
            for j, d in enumerate(dst_list):
                d = np.array(d)
                delta = d[:2] - s[:2]
                length = np.sqrt(np.sum(delta ** 2))            # l.353
                n = min(int(round(length + 1)), mid_num_max)    # l.354: Python round = half to even
                if length == 0.0:                               # l.357
                    continue
                xs = np.round(np.linspace(s[0], d[0], num=n)).astype(np.intp)   # l.361
                ys = np.round(np.linspace(s[1], d[1], num=n)).astype(np.intp)   # l.362
                vals = plane[ys, xs]
                score = vals.mean() + min(0.5 * img_height / length - 1, 0)    # l.367
                if np.count_nonzero(vals > thre2) > n * connect_ration and score > 0:      # l.373-378
                    cands.append((i, j, score, length, 0.5 * score + 0.25 * s[2] + 0.25 * d[2]))   # l.381
        cands.sort(key=lambda c: c[4], reverse=True)            # l.391: stable, ties keep pair order
        limit = min(len(src_list), len(dst_list))
        rows, used_i, used_j = [], set(), set()
        for i, j, score, length, _ in cands:                    # l.397-407
            if i not in used_i and j not in used_j:
                rows.append([src_list[i][3], dst_list[j][3], score, i, j, length])
                used_i.add(i)
                used_j.add(j)
                if len(rows) >= limit:
                    break
        connected.append(np.array(rows, dtype=np.float64).reshape(-1, 6))
    return connected, special


# ----------------------------------------------------------------------------------------------------------
# A15  people, Python semantics — utils/parse_skeletons.py:413-599 (find_humans)
# ----------------------------------------------------------------------------------------------------------
def find_humans(connected_limbs, special_limb, joint_list, len_rate=16, connection_tole=0.7, pairs=LIMBS):
    """-> (person_to_joint_assoc (P,20,2) float64, joint_candidates (N,4) float64).  Rows: [part] = (peak id, limb
    score), [18] = (total score, -1), [19] = (count, last limb length).  remove_recon = 0 (utils/config:27)."""
    cand = np.array([item for sub in joint_list for item in sub])
    people = []  # list of (20,2) float64 arrays, order = creation order minus deletions
    for k, (pa, pb) in enumerate(pairs):
        if k in special_limb:
            continue
        for info in connected_limbs[k]:
            src_id, dst_id, score, length = info[0], info[1], info[2], info[5]
            hit = [p for p in people if p[pa, 0] == src_id or p[pb, 0] == dst_id][:2]   # l.441-449: third match ignored
            if len(hit) == 1:
                p = hit[0]
                cur_id, cur_sc, reach = p[pb, 0], p[pb, 1], p[19, 1]
                if int(cur_id) == -1 and reach * len_rate > length:                       # l.458
                    p[pb] = [dst_id, score]
                    p[19, 0] += 1
                    p[19, 1] = max(length, reach)
                    p[18, 0] += cand[int(dst_id), 2] + score
                elif (int(cur_id) != int(dst_id) and cur_sc <= score and reach * len_rate > length) or \
                        (int(cur_id) == int(dst_id) and cur_sc <= score):                 # l.473-475, l.483-484
                    p[18, 0] -= cand[int(cur_id), 2] + cur_sc
                    p[pb] = [dst_id, score]
                    p[19, 1] = max(length, reach)
                    p[18, 0] += cand[int(dst_id), 2] + score
            elif len(hit) == 2:
                p1, p2 = hit
                m1, m2 = p1[:18, 0] >= 0, p2[:18, 0] >= 0
                if not np.any(m1 & m2):                                                   # l.506
                    tol = min(np.min(p1[:18, 1][m1]), np.min(p2[:18, 1][m2]))
                    if score >= connection_tole * tol and length < p1[19, 1] * len_rate:  # l.511-512
                        reach = p1[19, 1]
                        p1[:18] = np.maximum(p1[:18], p2[:18])
                        p1[19, 0] += p2[19, 0]
                        p1[19, 1] = max(length, reach)
                        p1[18, 0] += p2[18, 0] + score
                        people = [p for p in people if p is not p2]
            else:                                                                         # l.570-585
                row = -1 * np.ones((20, 2))
                row[pa] = [src_id, score]
                row[pb] = [dst_id, score]
                row[19] = [2, length]
                row[18, 0] = sum(cand[[int(src_id), int(dst_id)], 2]) + score
                people.append(row)
    keep = [p for p in people if not (p[19, 0] < 2 or p[18, 0] / p[19, 0] < 0.45)]         # l.591-596
    assoc = np.array(keep, dtype=np.float64).reshape(-1, 20, 2)
    return assoc, cand


# ----------------------------------------------------------------------------------------------------------
# Image preparation — utils/parse_skeletons.py:52-72 (after the resize), utils/util.py:44-65
# ----------------------------------------------------------------------------------------------------------
def prepare_input(image, max_downsample=64, pad_value=128, flip=True):
    """uint8 (h,w,3) -> float32 (2 or 1, hp, wp, 3): the tensor the model is called on (l.72).

    padRightDownCorner pads below and to the right with `img*0 + padValue` (uint8), l.59 divides by 255 in float64
    and rounds once to float32, l.69-70 mirrors the PADDED image (the pad ends up on the left of the twin)."""
    h, w = image.shape[:2]
    ph = 0 if h % max_downsample == 0 else max_downsample - h % max_downsample
    pw = 0 if w % max_downsample == 0 else max_downsample - w % max_downsample
    if h == 1:
        ph = 0  # util.py:59 tiles img[-2:-1], which is empty for a one-row image: nothing is appended
    if w == 1:
        pw = 0  # util.py:61 likewise for a one-column image
    padded = np.full((h + ph, w + pw, image.shape[2]), pad_value, dtype=image.dtype)
    padded[:h, :w] = image
    x = f32(padded / 255)
    if not flip:
        return x[None, ...]
    return np.concatenate((x[None, ...], x[:, ::-1, :][None, ...]), axis=0)


# ----------------------------------------------------------------------------------------------------------
# Image resize of the scale search — utils/parse_skeletons.py:204 / :52: cv2.resize(image, (0,0), fx, fy, INTER_CUBIC) on
# uint8.  Third-party arithmetic (OpenCV 4.13 resize.cpp, 8-bit cubic path), restated; pinned by tests/golden/resize_9x.npz
# (outputs of cv2.resize itself, IPP off) and against the live cv2 in tests/test_resize.py.
# ----------------------------------------------------------------------------------------------------------
def _u8_axis(n_src, n_dst, scale):
    """Per destination index: four clamped source indices and the 11-bit fixed-point taps."""
    d = np.arange(n_dst, dtype=np.float64)
    fx = ((d + 0.5) * scale - 0.5).astype(f32)
    s = np.floor(fx).astype(np.int64)
    t = (fx - s.astype(f32)).astype(f32)
    w = np.stack(_cubic_weights(t), axis=1).astype(f32)                       # (n_dst, 4) float32, A = -0.75
    taps = np.clip(np.rint(w * f32(2048)), -32768, 32767).astype(np.int64)     # saturate_cast<short>(w * 2048)
    idx = np.clip(s[:, None] + np.arange(-1, 3)[None, :], 0, n_src - 1)
    return idx, taps


def cv_resize_cubic_u8(img, fx, fy=None, lanes=8):
    """cv2.resize(img, (0, 0), fx=fx, fy=fy, interpolation=cv2.INTER_CUBIC) for a uint8 (h,w,c) image."""
    fy = fx if fy is None else fy
    h, w, cn = img.shape
    dh, dw = int(np.rint(h * fy)), int(np.rint(w * fx))       # saturate_cast<int>(double): half to even
    xi, xa = _u8_axis(w, dw, 1.0 / fx)                          # dsize empty: scale = 1 / inv_scale
    yi, yb = _u8_axis(h, dh, 1.0 / fy)
    src = img.astype(np.int64)
    H = np.zeros((h, dw, cn), dtype=np.int64)                   # horizontal pass, int
    for k in range(4):
        H += src[:, xi[:, k], :] * xa[:, k][None, :, None]
    H = H.reshape(h, dw * cn)
    rows = [H[yi[:, k]] for k in range(4)]                      # (dh, dw*cn) each
    scale = f32(1.0) / f32(2048 * 2048)
    b = [(yb[:, k].astype(f32) * scale)[:, None] for k in range(4)]
    S = [r.astype(f32) for r in rows]
    acc = (S[0] * b[0]) + ((S[1] * b[1]) + ((S[2] * b[2]) + (S[3] * b[3])))   # float32, separate multiplies and adds
    body = np.clip(np.rint(acc), 0, 255).astype(np.uint8)
    t = sum(rows[k] * yb[:, k][:, None] for k in range(4))
    tail = np.clip((t + (1 << 21)) >> 22, 0, 255).astype(np.uint8)             # FixedPtCast<int, uchar, 22>
    n_vec = (dw * cn // lanes) * lanes
    out = body
    out[:, n_vec:] = tail[:, n_vec:]
    return out.reshape(dh, dw, cn)


# ----------------------------------------------------------------------------------------------------------
# rotation_search branch — utils/parse_skeletons.py:215-218, :265-267 (predict), :63-66, :98-100 (predict_refactor).
# Third-party arithmetic: cv2.getRotationMatrix2D and cv2.warpAffine (INTER_LINEAR, BORDER_CONSTANT 0, same size) of OpenCV
# 4.13 for float32 data — the fixed-point remap path (imgwarp.cpp), restated; checked against the live cv2 for 1/3/4/20/30
# channels (tests/test_rotation.py) and pinned through goldens of the reference's own predict (tests/golden/rot_11x.npz).
# ----------------------------------------------------------------------------------------------------------
def rotation_matrix(cx, cy, angle_deg, scale=1.0):
    import math
    angle = angle_deg * (math.pi / 180)
    alpha, beta = math.cos(angle) * scale, math.sin(angle) * scale
    cx, cy = float(f32(cx)), float(f32(cy))  # Point2f
    return np.array([[alpha, beta, (1 - alpha) * cx - beta * cy], [-beta, alpha, beta * cx + (1 - alpha) * cy]], dtype=np.float64)


def warp_affine(src, matrix):
    """cv2.warpAffine(src (h,w,c) float32, matrix, (0, 0)) with the default flags."""
    src = np.asarray(src, dtype=f32)
    squeeze = src.ndim == 2
    if squeeze:
        src = src[:, :, None]
    h, w = src.shape[:2]
    M = np.asarray(matrix, dtype=np.float64).reshape(-1).copy()
    D = M[0] * M[4] - M[1] * M[3]                       # cv::warpAffine: invert (no WARP_INVERSE_MAP)
    D = 1. / D if D != 0 else 0.
    A11, A22 = M[4] * D, M[0] * D
    M[0] = A11
    M[1] *= -D
    M[3] *= -D
    M[4] = A22
    b1 = -M[0] * M[2] - M[1] * M[5]
    b2 = -M[3] * M[2] - M[4] * M[5]
    M[2], M[5] = b1, b2
    x, y = np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64)
    rint = lambda v: np.rint(v).astype(np.int64)        # saturate_cast<int>(double)
    adelta, bdelta = rint(M[0] * x * 1024), rint(M[3] * x * 1024)
    X0, Y0 = rint((M[1] * y + M[2]) * 1024) + 16, rint((M[4] * y + M[5]) * 1024) + 16
    X, Y = (X0[:, None] + adelta[None, :]) >> 5, (Y0[:, None] + bdelta[None, :]) >> 5
    sx, sy = np.clip(X >> 5, -32768, 32767), np.clip(Y >> 5, -32768, 32767)
    t = (np.arange(32, dtype=f32) * f32(1.0 / 32)).astype(f32)
    c0, c1 = (f32(1) - t).astype(f32), t
    fx, fy = X & 31, Y & 31
    w00, w01 = c0[fy] * c0[fx], c0[fy] * c1[fx]
    w10, w11 = c1[fy] * c0[fx], c1[fy] * c1[fx]

    def px(yy, xx):
        ok = (yy >= 0) & (yy < h) & (xx >= 0) & (xx < w)
        return np.where(ok[..., None], src[np.clip(yy, 0, h - 1), np.clip(xx, 0, w - 1)], f32(0))

    out = px(sy, sx) * w00[..., None]
    out = out + px(sy, sx + 1) * w01[..., None]
    out = out + px(sy + 1, sx) * w10[..., None]
    out = out + px(sy + 1, sx + 1) * w11[..., None]
    return out[:, :, 0] if squeeze else out


def rotated_input(scaled_u8, angle, max_downsample=64, pad_value=128):
    """l.206-224 with a rotation: pad, /255, warpAffine about (rows/2, cols/2) given as (x, y), flipped twin of the rotated
    image.  -> ((2,hp,wp,3) float32, reverse matrix)."""
    x = prepare_input(scaled_u8, max_downsample, pad_value, flip=False)[0]
    if angle != 0:
        m_fwd = rotation_matrix(x.shape[0] / 2, x.shape[1] / 2, angle, 1)
        x = warp_affine(x, m_fwd)
    m_rev = rotation_matrix(x.shape[0] / 2, x.shape[1] / 2, -angle, 1)
    return np.concatenate((x[None, ...], x[:, ::-1, :][None, ...]), axis=0), m_rev


def predict_fuse_rot(items, img_h, img_w, stride=4, flip_avg=True):
    """predict() after the model calls with a rotation search.  items, in the reference's loop order: (net_out (2,50,h,w),
    scaled (sh,sw), angle).  -> heatmap_avg (H,W,20), paf_avg (H,W,30) float64."""
    heatmap_avg = np.zeros((img_h, img_w, NUM_HEATMAPS))
    paf_avg = np.zeros((img_h, img_w, NUM_PAFS))
    n = len(items)
    for out, (sh, sw), angle in items:
        pad_h, pad_w = out.shape[2] * stride, out.shape[3] * stride
        heat, paf = fuse_flip(out, flip_avg)
        heat = cv_resize_cubic(heat, pad_h, pad_w, 1.0 / stride, 1.0 / stride)
        paf = cv_resize_cubic(paf, pad_h, pad_w, 1.0 / stride, 1.0 / stride)
        if angle != 0:                                                                  # l.265-267
            m_rev = rotation_matrix(pad_h / 2, pad_w / 2, -angle, 1)
            heat, paf = warp_affine(heat, m_rev), warp_affine(paf, m_rev)
        heat, paf = heat[:sh, :sw, :], paf[:sh, :sw, :]
        heat = cv_resize_cubic(heat, img_h, img_w)
        paf = cv_resize_cubic(paf, img_h, img_w)
        heatmap_avg = heatmap_avg + heat / n
        paf_avg = paf_avg + paf / n
    return heatmap_avg, paf_avg


# ----------------------------------------------------------------------------------------------------------
# Result rows — evaluate.py:182-209 (append_result, --run_refactor branch)
# ----------------------------------------------------------------------------------------------------------
def coco_rows(peak_xy, humans):
    """peak_xy (N,2) integer peak positions by id; humans = [(ids[18], score)] -> (n, 52) float32: the `keypoints`
    list (17 x (x, y, v), CMU-18 reordered by ORDER_COCO, v = 1 for a found part, zeros otherwise) then `score`."""
    rows = np.zeros((len(humans), 17 * 3 + 1), dtype=np.float32)
    for k, (ids, score) in enumerate(humans):
        cmu = np.zeros((NUM_KEYPOINTS, 3), dtype=np.float64)
        for i in range(NUM_KEYPOINTS):
            if ids[i] >= 0:
                cmu[i] = (peak_xy[ids[i]][0], peak_xy[ids[i]][1], 1)
        rows[k, :51] = cmu[ORDER_COCO, :].reshape(51)
        rows[k, 51] = score
    return rows


# ----------------------------------------------------------------------------------------------------------
# Second operator mode (the operators BASELINE.json names where the reference uses others).
# NOT on the reference's live path: predict() with cv2.INTER_LINEAR in both resizes, and GaussianSmoothing of
# utils/util.py:103-174 (unused by the reference) applied to a mirror-padded map as its docstring says.
# Pinned by: cv_resize_linear == cv2.resize(INTER_LINEAR) (tests/test_operator_modes.py), gaussian_smooth ==
# F.conv2d with the reference class's own weight (tests/golden/modes_12x.npz, made by make_golden_modes.py).
# ----------------------------------------------------------------------------------------------------------
def _cv_linear_axis(n_src, n_dst, scale):
    """cv2 INTER_LINEAR tap set-up of one axis (resize.cpp, generic path): -> (i0, i1, a0, a1) float32 weights."""
    d = np.arange(n_dst, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int64)
    f = (f - s.astype(np.float32)).astype(np.float32)
    lo = s < 0
    f[lo], s[lo] = 0.0, 0
    hi = s >= n_src - 1
    f[hi], s[hi] = 0.0, n_src - 1
    return s, np.minimum(s + 1, n_src - 1), (np.float32(1.0) - f).astype(np.float32), f


def cv_resize_linear(src, dst_h, dst_w, scale_y=None, scale_x=None):
    """cv2.resize(src (h,w,C) float32, (dst_w,dst_h), interpolation=cv2.INTER_LINEAR): horizontal pass
    S[x0]*a0 + S[x1]*a1, then vertical pass S0*b0 + S1*b1, float32 throughout."""
    src = np.ascontiguousarray(src, dtype=np.float32)
    h, w = src.shape[:2]
    sy = scale_y if scale_y is not None else 1.0 / (dst_h / h)
    sx = scale_x if scale_x is not None else 1.0 / (dst_w / w)
    x0, x1, a0, a1 = _cv_linear_axis(w, dst_w, sx)
    y0, y1, b0, b1 = _cv_linear_axis(h, dst_h, sy)
    rows = src[:, x0, :] * a0[None, :, None] + src[:, x1, :] * a1[None, :, None]
    return (rows[y0] * b0[:, None, None] + rows[y1] * b1[:, None, None]).astype(np.float32)


def predict_fuse_linear(net_outs, img_h, img_w, scales, stride=4, max_downsample=64, flip_avg=True):
    """predict_fuse with INTER_LINEAR in both resizes (the second operator mode of ibp_fuse_multiscale_linear)."""
    heatmap_avg = np.zeros((img_h, img_w, NUM_HEATMAPS))
    paf_avg = np.zeros((img_h, img_w, NUM_PAFS))
    n = len(scales)
    for out, scale in zip(net_outs, scales):
        sh, sw = scaled_size(img_h, img_w, scale)
        pad_h, pad_w = padded_size(sh, sw, max_downsample)
        assert out.shape[2] * stride == pad_h and out.shape[3] * stride == pad_w, (out.shape, pad_h, pad_w)
        heat, paf = fuse_flip(out, flip_avg)
        heat = cv_resize_linear(heat, pad_h, pad_w, 1.0 / stride, 1.0 / stride)
        paf = cv_resize_linear(paf, pad_h, pad_w, 1.0 / stride, 1.0 / stride)
        heat, paf = heat[:sh, :sw, :], paf[:sh, :sw, :]
        heat = cv_resize_linear(heat, img_h, img_w)
        paf = cv_resize_linear(paf, img_h, img_w)
        heatmap_avg = heatmap_avg + heat / n
        paf_avg = paf_avg + paf / n
    return heatmap_avg, paf_avg


def gaussian_kernel1d(ksize, sigma):
    """utils/util.py:131-146 for one axis, float32 like torch: the 2-D kernel of GaussianSmoothing is the outer product
    of two of these divided by its sum, i.e. the outer product of the normalised 1-D kernels."""
    i = np.arange(ksize, dtype=np.float32)
    mean = np.float32((ksize - 1) / 2)
    g = (np.float32(1.0) / (np.float32(sigma) * np.float32(np.sqrt(2 * np.pi)))) * np.exp(-(((i - mean) / np.float32(sigma)) ** 2) / 2).astype(np.float32)
    return (g / g.sum(dtype=np.float32)).astype(np.float32)


def gaussian_smooth(maps, ksize=5, sigma=1.0):
    """GaussianSmoothing(C, ksize, sigma)(F.pad(maps, (p,p,p,p), mode='reflect')) for maps (C,H,W) float32 -> (C,H,W)."""
    maps = np.asarray(maps, dtype=np.float32)
    p = ksize // 2
    g = gaussian_kernel1d(ksize, sigma).astype(np.float64)
    x = np.pad(maps, ((0, 0), (p, p), (p, p)), mode="reflect").astype(np.float64)
    H, W = maps.shape[1:]
    tmp = sum(g[k] * x[:, :, k:k + W] for k in range(ksize))
    out = sum(g[k] * tmp[:, k:k + H, :] for k in range(ksize))
    return out.astype(np.float32)
