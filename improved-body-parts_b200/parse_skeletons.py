"""Mirror of the reference's utils/parse_skeletons.py call surface, computed on the GPU through libibp_b200.so.
Same function names, argument order and return containers; no CPU fallback.

    predict_refactor   parse_skeletons.py:28-103
    predict            parse_skeletons.py:180-283 (original route: scale search + flip fusion at image resolution)
    heatmap_nms        parse_skeletons.py:126-176
    find_connections   parse_skeletons.py:324-410 (Python semantics, float64)
    find_humans        parse_skeletons.py:413-599 (Python semantics, float64)
    find_peaks         parse_skeletons.py:286-321 (original route: 3x3 NMS + centroid refinement, utils/util.py:177-213)
    limb_upsample      evaluate.py:77-80 (the cv2.resize between heatmap_nms and process_paf)
"""
import numpy as np
import torch

from . import _lib
from .pipeline import PostProcessor

NUM_KEYPOINTS = 18
NUM_HEATMAPS = NUM_KEYPOINTS + 2
NUM_PAFS = 30

_pp = {}
# capacities of the single-image handles behind the mirrors; doubled on IBP_ERR_CAPACITY (the reference's lists are
# unbounded, a crowd image must not fail here because the defaults were sized for ten people)
_GROWTH = [dict(max_peaks_per_part=128, max_candidates_per_limb=1024, max_bodies=128, max_humans=128),
           dict(max_peaks_per_part=256, max_candidates_per_limb=2048, max_bodies=256, max_humans=256),
           dict(max_peaks_per_part=512, max_candidates_per_limb=4096, max_bodies=512, max_humans=512),
           dict(max_peaks_per_part=1024, max_candidates_per_limb=4096, max_bodies=1024, max_humans=1024)]
_level = [0]


def _processor(device=0, upsample=4, peak_thresh=0.1, refine_window=2):
    """One single-image PostProcessor per (device, peak parameters): the constants live in the handle."""
    key = (device, int(upsample), float(peak_thresh), int(refine_window), _level[0])
    if key not in _pp:
        p = _lib.default_params()
        p.upsample, p.peak_thresh, p.refine_window = key[1], key[2], key[3]
        pp = PostProcessor.__new__(PostProcessor)
        if not torch.cuda.is_available():
            raise _lib.IbpError(-6, "no CUDA device: improved-body-parts_b200 has no CPU fallback")
        pp.device = torch.device("cuda", device)
        pp.handle = _lib.Handle(device, params=p, max_batch=1, **_GROWTH[_level[0]])
        pp.max_batch = 1
        _pp[key] = pp
    return _pp[key]


def _growing(fn):
    """Run fn(); on IBP_ERR_CAPACITY move every mirror to the next capacity level (new handles) and run it again."""
    import functools

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        while True:
            try:
                return fn(*args, **kwargs)
            except _lib.IbpError as e:
                if e.code != -3 or _level[0] + 1 >= len(_GROWTH):
                    raise
                for k in [k for k in _pp if k[4] == _level[0]]:
                    _pp.pop(k).handle.close()
                _level[0] += 1
    return wrapper


def pad_right_down_corner(img, stride, pad_value):
    """utils/util.py:44-65: pad bottom/right to a multiple of `stride` with `pad_value`."""
    h, w = img.shape[:2]
    pad = [0, 0, 0 if h % stride == 0 else stride - h % stride, 0 if w % stride == 0 else stride - w % stride]
    out = np.full((h + pad[2], w + pad[3]) + img.shape[2:], pad_value, dtype=img.dtype)
    out[:h, :w] = img
    return out, pad


def predict_refactor(image, model, test_cfg, model_cfg, input_image_path, flip_avg=True, config=None):
    """-> (heatmaps (h,w,20), pafs (h,w,30)) float32 HWC, like parse_skeletons.py:28-103.  The scale is pinned to 1.0
    (l.36), so the cv2.resize is a copy unless the image is too big; padding, /255 and the flipped twin (l.54-72) and
    the flip average (l.80-96) run on the device."""
    scale = 1.0  # parse_skeletons.py:36
    img_h, img_w, _ = image.shape
    if scale * img_h > 2600 or scale * img_w > 3800:  # l.46-49
        scale = min(2600 / img_h, 3800 / img_w)
        print("Input image: '{}' is too big, shrink it!".format(input_image_path))
    pp = _processor()
    img = _scaled_image(pp, image, scale)  # l.52, on the device (a scale of exactly 1.0 is an identity)
    fused = None
    for angle in test_cfg.get('rotation_search', [0.0]):  # l.45: product([1.], rotate_angle); the LAST item is returned (l.101-102)
        if angle == 0:
            batch = pp.prepare_images(img, model_cfg['max_downsample'], model_cfg['padValue'], flip=True)  # l.54-72
        else:
            batch, m_rev = _rotated_input(pp, img, model_cfg, angle)                                        # l.63-66
        with torch.no_grad():
            out = model(batch)
        if isinstance(out, (list, tuple)):
            out = out[-1][0]  # last stack, scale 0 (l.80)
        fused = pp.fuse_flip(out[None, ...], flip_avg)  # (1,50,h,w)
        if angle != 0:
            # l.98-100: the matrix built for the PADDED INPUT size is applied to the feature-resolution maps as it is
            fused = pp.warp_affine(fused, m_rev, "planar")
    hwc = fused[0].permute(1, 2, 0).contiguous().cpu().numpy()
    return hwc[:, :, NUM_PAFS:NUM_PAFS + NUM_HEATMAPS].copy(), hwc[:, :, :NUM_PAFS].copy()


def _rotated_input(pp, scaled, model_cfg, angle):
    """l.54-72 / l.206-224 with angle != 0: pad, /255, cv2.warpAffine by `angle` about (rows/2, cols/2) — the reference passes
    (shape[0]/2, shape[1]/2) as the (x, y) centre, kept — then the flipped twin of the ROTATED image.  -> (batch (2,hp,wp,3),
    reverse matrix)."""
    x = pp.prepare_images(scaled, model_cfg['max_downsample'], model_cfg['padValue'], flip=False)  # (1,hp,wp,3)
    hp, wp = int(x.shape[1]), int(x.shape[2])
    m_fwd = _lib.rotation_matrix(hp / 2, wp / 2, angle, 1)
    m_rev = _lib.rotation_matrix(hp / 2, wp / 2, -angle, 1)
    x = pp.warp_affine(x, m_fwd, "hwc")
    return torch.cat((x, x.flip(2)), dim=0), m_rev


def _scaled_image(pp, image, scale):
    """cv2.resize(image, (0, 0), fx=scale, fy=scale, interpolation=cv2.INTER_CUBIC) (l.52 / l.204) on the device:
    uint8 (h,w,3) numpy / tensor -> uint8 (1, rint(h*scale), rint(w*scale), 3) CUDA tensor."""
    t = image if isinstance(image, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(image))
    if t.dtype != torch.uint8 or t.dim() != 3 or t.shape[2] != 3:
        raise ValueError("image must be uint8 (h, w, 3) BGR, got %s %s" % (t.dtype, tuple(t.shape)))
    t = t.to(pp.device)
    if scale == 1.0:
        return t[None, ...]
    return pp.resize_images(t, scale)


@_growing
def heatmap_nms(heatmaps, upsample_factor=1., bool_refine_center=True):
    """heatmaps (h,w,>=18) float32 HWC -> list of 18 arrays (n,4) float32 [x, y, score, id] (l.126-176).
    bool_refine_center=False (l.163-166): no patch refine — x, y = compute_resized_coords(pixel) = (p + 0.5) * f - 0.5 in
    float32 (l.122-123), score = the map value at the pixel."""
    pp = _processor(upsample=max(int(upsample_factor), 1))
    maps = torch.from_numpy(np.ascontiguousarray(heatmaps[:, :, :NUM_KEYPOINTS], dtype=np.float32))
    maps = maps.to(pp.device).permute(2, 0, 1).contiguous()[None, ...]
    pp.find_peaks(maps, 0, flavour=_lib.PEAKS_CROSS_REFINE if bool_refine_center else _lib.PEAKS_CROSS_RAW)
    res = pp.fetch(1)
    peaks = res.peaks[0]
    out, pos = [], 0
    for part in range(NUM_KEYPOINTS):
        n = int(res.part_count[0, part])
        rows = np.zeros((n, 4), dtype=np.float32)
        sel = peaks[pos:pos + n]
        if bool_refine_center:
            rows[:, 0], rows[:, 1] = sel["x"], sel["y"]
        else:  # l.122-123 on the integer pixel the device reports
            xy = (np.stack([sel["x"], sel["y"]], axis=1).astype(np.float32) + 0.5) * upsample_factor - 0.5
            rows[:, 0], rows[:, 1] = xy[:, 0], xy[:, 1]
        rows[:, 2] = sel["score"]
        rows[:, 3] = np.arange(pos, pos + n)
        out.append(rows)
        pos += n
    return out


def limb_upsample(pafs, upsample_factor=4):
    """evaluate.py:77-80: cv2.resize(pafs, None, fx=stride, fy=stride, INTER_CUBIC) on (h,w,30) -> (H,W,30)."""
    pp = _processor()
    t = torch.from_numpy(np.ascontiguousarray(pafs, dtype=np.float32)).to(pp.device).permute(2, 0, 1).contiguous()
    return pp.upsample_limbs(t[None, ...], int(upsample_factor))[0].cpu().numpy()


@_growing
def find_peaks(heatmap_avg, test_cfg):
    """heatmap_avg (H,W,>=18) -> list of 18 lists of (x, y, score, id) like parse_skeletons.py:286-321: 3x3 NMS with
    `value >= thre1` (utils/util.py:177-185), 5x5 centroid refinement in float64 (utils/util.py:188-213; border peaks
    keep their integer position and raw value), running ids across joint types."""
    pp = _processor(peak_thresh=test_cfg['thre1'], refine_window=test_cfg['offset_radius'])
    hm = np.asarray(heatmap_avg).astype(np.float32)  # l.291
    maps = torch.from_numpy(np.ascontiguousarray(hm[:, :, :NUM_KEYPOINTS].transpose(2, 0, 1))).to(pp.device)[None, ...]
    pp.find_peaks(maps, 0, flavour=_lib.PEAKS_BOX_CENTROID)
    res = pp.fetch(1)
    n = int(res.n_peaks[0])
    xy = pp.handle.fetch_subpix(0, n)
    peaks = res.peaks[0]
    h, w = hm.shape[:2]
    rad = int(test_cfg['offset_radius'])
    out, pos = [], 0
    for part in range(NUM_KEYPOINTS):
        rows = []
        for k in range(int(res.part_count[0, part])):
            x, y = xy[pos]
            ix, iy = int(x), int(y)
            border = iy + rad + 1 > h or iy - rad < 0 or ix + rad + 1 > w or ix - rad < 0
            # util.py:203-204 returns the integer anchor at the border, float64 positions elsewhere
            rows.append(((ix, iy) if border and x == ix and y == iy else (x, y)) + (peaks[pos]["score"], pos))
            pos += 1
        out.append(rows)
    return out


def predict(image, model, test_cfg, model_cfg, input_image_path, flip_avg=True, config=None, scale_search=None):
    """-> (heatmap_avg (H,W,20), paf_avg (H,W,30)) float64 HWC like parse_skeletons.py:180-283.

    The reference computes the scale-search multipliers (l.186) and then pins them to [1.] (l.188); `scale_search=None`
    does the same.  Pass a list such as test_cfg['scale_search'] to run the multi-scale averaging the reference ships
    but never reaches.  rotation_search is [0.0] in utils/config:15; a non-zero angle takes the rotation branch
    (l.215-218, 265-267; _predict_with_rotation below).  The image goes to the device once (uint8); per scale the cubic resize (l.204), the pad to a multiple of
    max_downsample with padValue, the /255 and the flipped twin (l.206-224) run there; everything after the model calls
    (flip average, cubic x stride, crop, cubic resize to the image, average) is one kernel launch."""
    img_h, img_w, _ = image.shape
    if scale_search is None:
        multiplier = [1.]  # l.188
    else:
        multiplier = [x * model_cfg['boxsize'] / img_h for x in scale_search]  # l.186
    angles = list(test_cfg.get('rotation_search', [0.0]))
    if any(a != 0 for a in angles):
        return _predict_with_rotation(image, model, multiplier, angles, model_cfg, input_image_path, flip_avg)
    pp = _processor()
    img_dev = torch.from_numpy(np.ascontiguousarray(image)).to(pp.device)  # one upload, uint8
    outs, sizes = [], []
    for scale in multiplier:
        if scale * img_h > 2600 or scale * img_w > 3800:  # l.196-199
            scale = min(2600 / img_h, 3800 / img_w)
            print("Input image: '{}' is too big, shrink it!".format(input_image_path))
        scaled = _scaled_image(pp, img_dev, scale)                                                       # l.204
        batch = pp.prepare_images(scaled, model_cfg['max_downsample'], model_cfg['padValue'], flip=True)  # l.206-224
        with torch.no_grad():
            out = model(batch)
        if isinstance(out, (list, tuple)):
            out = out[-1][0]  # last stack, scale 0 (l.232)
        outs.append(out.to(pp.device, torch.float32)[None, ...])
        sizes.append(tuple(int(v) for v in scaled.shape[1:3]))
    fused, _ = pp.fuse_multiscale(outs, sizes, img_h, img_w, flip_avg, want_heat32=False)
    hwc = fused[0].permute(1, 2, 0).contiguous().cpu().numpy()
    return hwc[:, :, NUM_PAFS:NUM_PAFS + NUM_HEATMAPS].copy(), hwc[:, :, :NUM_PAFS].copy()


def _predict_with_rotation(image, model, multiplier, angles, model_cfg, input_image_path, flip_avg):
    """predict() when rotation_search has a non-zero angle (l.196-281, the reference's loop order product(multiplier,
    rotate_angle)).  A warp sits between the two resizes of such an item, so the fused multi-scale kernel cannot take it;
    per item: x4 maps at the padded size (the fused kernel with ONE scale and an identity second resize — exact), backward
    warp (l.265-267), then crop + cubic resize to the image + /n + float64 accumulate (l.272-281) in one kernel."""
    img_h, img_w, _ = image.shape
    pp = _processor()
    img_dev = torch.from_numpy(np.ascontiguousarray(image)).to(pp.device)
    n_items = len(multiplier) * len(angles)
    avg = torch.zeros((1, NUM_PAFS + NUM_HEATMAPS, img_h, img_w), dtype=torch.float64, device=pp.device)
    for scale in multiplier:
        for angle in angles:
            if scale * img_h > 2600 or scale * img_w > 3800:  # l.196-199
                scale = min(2600 / img_h, 3800 / img_w)
                print("Input image: '{}' is too big, shrink it!".format(input_image_path))
            scaled = _scaled_image(pp, img_dev, scale)
            if angle == 0:
                batch = pp.prepare_images(scaled, model_cfg['max_downsample'], model_cfg['padValue'], flip=True)
            else:
                batch, m_rev = _rotated_input(pp, scaled, model_cfg, angle)
            with torch.no_grad():
                out = model(batch)
            if isinstance(out, (list, tuple)):
                out = out[-1][0]
            out = out.to(pp.device, torch.float32)[None, ...]
            ph, pw = int(out.shape[3]) * model_cfg['stride'], int(out.shape[4]) * model_cfg['stride']
            up64, _ = pp.fuse_multiscale([out], [(ph, pw)], ph, pw, flip_avg, want_heat32=False)  # l.232-263
            up = up64.float()  # values are float32 results widened: narrowing is exact
            if angle != 0:
                up = pp.warp_affine(up, m_rev, "planar")
            pp.crop_resize_accumulate(up, scaled.shape[1:3], avg, n_items)
    hwc = avg[0].permute(1, 2, 0).contiguous().cpu().numpy()
    return hwc[:, :, NUM_PAFS:NUM_PAFS + NUM_HEATMAPS].copy(), hwc[:, :, :NUM_PAFS].copy()


def _check_pairs(joint2limb_pairs):
    from .config import LIMBS
    if joint2limb_pairs is not None and [tuple(p) for p in joint2limb_pairs] != [tuple(p) for p in LIMBS]:
        raise ValueError("joint2limb_pairs must be the 30 canonical limbs (config/config.py:114-121)")


@_growing
def find_connections(all_peaks, paf_avg, img_height, test_cfg, joint2limb_pairs=None):
    """-> (connected_limbs, special_limb) like parse_skeletons.py:324-410: connected_limbs[k] is [] when both joint
    lists of limb k are empty (k is listed in special_limb), else an (n,6) float64 array of rows
    [src_peak_id, dst_peak_id, score, i, j, length].  all_peaks: 18 lists of (x, y, score, id); paf_avg (H,W,30)."""
    _check_pairs(joint2limb_pairs)
    pp = _processor()
    rows = np.array([tuple(float(v) for v in p) for ps in all_peaks for p in ps], dtype=np.float64).reshape(-1, 4)
    part = np.array([t for t, ps in enumerate(all_peaks) for _ in ps], dtype=np.int32)
    pp.handle.set_peaks64(0, rows, part)
    paf = torch.from_numpy(np.ascontiguousarray(np.asarray(paf_avg, dtype=np.float64)[:, :, :NUM_PAFS].transpose(2, 0, 1)))
    paf = paf.to(pp.device)
    H, W = paf.shape[1:]
    with torch.cuda.device(pp.device):
        pp.handle.find_connections(_lib.py_params(test_cfg), paf.data_ptr(), NUM_PAFS, 1, H, W, img_height,
                                   torch.cuda.current_stream(pp.device).cuda_stream)
        got, counts = pp.handle.fetch_connections(0, torch.cuda.current_stream(pp.device).cuda_stream)
    from .config import LIMBS
    connected, special = [], []
    for k, (a, b) in enumerate(LIMBS):
        if len(all_peaks[a]) == 0 and len(all_peaks[b]) == 0:
            special.append(k)
            connected.append([])
        else:
            connected.append(got[k, :counts[k]].copy())
    return connected, special


@_growing
def find_humans(connected_limbs, special_limb, joint_list, test_cfg, joint2limb_pairs=None):
    """-> (person_to_joint_assoc (P,20,2) float64, joint_candidates (N,4)) like parse_skeletons.py:413-599."""
    _check_pairs(joint2limb_pairs)
    pp = _processor()
    cand = np.array([item for sub in joint_list for item in sub])  # l.427
    cap = pp.handle.limits.max_peaks_per_part
    rows = np.zeros((NUM_PAFS, cap, 6), dtype=np.float64)
    counts = np.zeros(NUM_PAFS, dtype=np.int32)
    for k in range(NUM_PAFS):
        if k in special_limb or len(connected_limbs[k]) == 0:
            continue
        r = np.asarray(connected_limbs[k], dtype=np.float64).reshape(-1, 6)
        if r.shape[0] > cap:
            raise _lib.IbpError(-3, "limb %d has %d connections (max %d)" % (k, r.shape[0], cap))
        rows[k, :r.shape[0]] = r
        counts[k] = r.shape[0]
    scores = cand[:, 2].astype(np.float64) if cand.size else np.zeros(0)
    pp.handle.set_connections(0, rows, counts, scores)
    with torch.cuda.device(pp.device):
        st = torch.cuda.current_stream(pp.device).cuda_stream
        pp.handle.find_humans(_lib.py_params(test_cfg), 1, st)
        people = pp.handle.fetch_people(0, st)
    return people, cand
