"""ctypes binding of libibp_b200.so (the C ABI declared in include/ibp.h).

There is no Python or CPU implementation behind these calls: if the shared library has not been built, or no CUDA
device is visible, everything here raises — loudly — instead of falling back.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libibp_b200.so")

NUM_PARTS, NUM_LIMBS, NUM_LAYERS = 18, 30, 50
IBP_OK = 0
ERR_NAMES = {-1: "IBP_ERR_SHAPE", -2: "IBP_ERR_PART", -3: "IBP_ERR_CAPACITY", -4: "IBP_ERR_CUDA", -5: "IBP_ERR_ARG",
             -6: "IBP_ERR_NO_DEVICE"}
IBP_F32, IBP_F16 = 0, 1
LIMB_FULLRES_HWC, LIMB_FEAT_PLANAR = 0, 1
PEAKS_CROSS_REFINE, PEAKS_BOX_CENTROID, PEAKS_CROSS_RAW = 0, 1, 2
ST_PEAK_OVERFLOW, ST_CAND_OVERFLOW, ST_BODY_OVERFLOW, ST_HUMAN_OVERFLOW = 0x1, 0x2, 0x4, 0x8
ST_TIE_EMULATED, ST_TIE_UNDEFINED = 0x10, 0x20


class IbpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s (%d): %s" % (ERR_NAMES.get(code, "IBP_ERR"), code, msg))
        self.code = code


class Params(C.Structure):
    _fields_ = [("limb_sample_thresh", C.c_float), ("step_ratio", C.c_float), ("min_parts", C.c_int32),
                ("min_human_score", C.c_float), ("max_steps", C.c_int32), ("len_rate", C.c_int32),
                ("merge_tolerance", C.c_float), ("weight_limb", C.c_float), ("weight_a", C.c_float),
                ("weight_b", C.c_float), ("peak_thresh", C.c_float), ("upsample", C.c_int32),
                ("refine_window", C.c_int32)]


class Limits(C.Structure):
    _fields_ = [("max_batch", C.c_int32), ("max_peaks_per_part", C.c_int32), ("max_candidates_per_limb", C.c_int32),
                ("max_bodies", C.c_int32), ("max_humans", C.c_int32), ("max_feat_h", C.c_int32),
                ("max_feat_w", C.c_int32)]


class PyParams(C.Structure):
    """utils/config:17-27 as Python floats (the Python route computes in float64)."""
    _fields_ = [("thre2", C.c_double), ("connect_ration", C.c_double), ("mid_num", C.c_int32), ("len_rate", C.c_int32),
                ("connection_tole", C.c_double), ("min_score", C.c_double), ("min_parts", C.c_int32)]


class ScaleInput(C.Structure):
    _fields_ = [("net_out", C.c_void_p), ("fh", C.c_int32), ("fw", C.c_int32), ("sh", C.c_int32), ("sw", C.c_int32)]


class BatchView(C.Structure):
    _fields_ = [("batch", C.c_int32), ("peak_stride", C.c_int32), ("human_stride", C.c_int32),
                ("n_peaks", C.POINTER(C.c_int32)), ("part_count", C.POINTER(C.c_int32)), ("peaks", C.c_void_p),
                ("n_humans", C.POINTER(C.c_int32)), ("humans", C.c_void_p), ("status", C.POINTER(C.c_uint32))]


class CocoView(C.Structure):
    _fields_ = [("batch", C.c_int32), ("offsets", C.POINTER(C.c_int32)), ("rows", C.POINTER(C.c_float)),
                ("status", C.POINTER(C.c_uint32))]


COCO_ROW = 52  # 17 x (x, y, v) + score, include/ibp.h IBP_COCO_ROW
PEAK_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("score", "<f4"), ("part", "<i4")])
HUMAN_DTYPE = np.dtype([("peak_id", "<i4", (NUM_PARTS,)), ("score", "<f4"), ("total", "<f4"), ("count", "<i4"),
                        ("reach", "<f4")])
assert PEAK_DTYPE.itemsize == 16 and HUMAN_DTYPE.itemsize == 88

_lib = None


def lib():
    """The loaded library; raises if it is missing (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libibp_b200.so is not built: run `python improved-body-parts_b200/build.py` "
                           "(or __graft_entry__.build()). There is no CPU fallback for this path.")
    L = C.CDLL(LIB_PATH)
    fp, vp, ip = C.POINTER(C.c_float), C.c_void_p, C.c_int
    L.ibp_process_paf.restype = ip
    L.ibp_process_paf.argtypes = [ip, ip, ip, vp, ip, ip, ip, vp, ip]
    L.ibp_get_num_humans.restype = ip
    L.ibp_get_part_peak_id.restype = ip
    L.ibp_get_part_peak_id.argtypes = [ip, ip]
    L.ibp_get_score.restype = C.c_float
    L.ibp_get_score.argtypes = [ip]
    L.ibp_get_part_x.restype = ip
    L.ibp_get_part_x.argtypes = [ip]
    L.ibp_get_part_y.restype = ip
    L.ibp_get_part_y.argtypes = [ip]
    L.ibp_get_part_score.restype = C.c_float
    L.ibp_get_part_score.argtypes = [ip]
    L.ibp_get_status.restype = C.c_uint32
    L.ibp_default_params.argtypes = [C.POINTER(Params)]
    L.ibp_default_limits.argtypes = [C.POINTER(Limits)]
    L.ibp_create.restype = ip
    L.ibp_create.argtypes = [C.POINTER(vp), ip, C.POINTER(Params), C.POINTER(Limits)]
    L.ibp_destroy.argtypes = [vp]
    L.ibp_last_error.restype = C.c_char_p
    L.ibp_device_count.restype = ip
    L.ibp_fuse_flip.restype = ip
    L.ibp_fuse_flip.argtypes = [vp, vp, ip, ip, ip, ip, ip, vp, vp]
    L.ibp_fuse_multiscale.restype = ip
    L.ibp_fuse_multiscale.argtypes = [vp, C.POINTER(ScaleInput), ip, ip, ip, ip, ip, vp, vp, vp]
    L.ibp_fuse_multiscale_linear.restype = ip
    L.ibp_fuse_multiscale_linear.argtypes = [vp, C.POINTER(ScaleInput), ip, ip, ip, ip, ip, vp, vp, vp]
    L.ibp_smooth_gaussian.restype = ip
    L.ibp_smooth_gaussian.argtypes = [vp, vp, vp, ip, ip, ip, ip, C.c_float, vp]
    L.ibp_upsample_limbs.restype = ip
    L.ibp_upsample_limbs.argtypes = [vp, vp, ip, ip, ip, ip, ip, vp, vp]
    L.ibp_find_peaks.restype = ip
    L.ibp_find_peaks.argtypes = [vp, vp, ip, ip, ip, ip, ip, ip, vp]
    L.ibp_set_peaks.restype = ip
    L.ibp_set_peaks.argtypes = [vp, ip, vp, ip, ip]
    L.ibp_assign.restype = ip
    L.ibp_assign.argtypes = [vp, vp, ip, ip, ip, ip, ip, ip, vp]
    L.ibp_postprocess_batch.restype = ip
    L.ibp_postprocess_batch.argtypes = [vp, vp, ip, ip, ip, ip, ip, ip, vp]
    L.ibp_fetch.restype = ip
    L.ibp_fetch.argtypes = [vp, ip, C.POINTER(BatchView), vp]
    L.ibp_fetch_subpix.restype = ip
    L.ibp_fetch_subpix.argtypes = [vp, ip, vp, ip, vp]
    L.ibp_default_py_params.argtypes = [C.POINTER(PyParams)]
    L.ibp_set_peaks64.restype = ip
    L.ibp_set_peaks64.argtypes = [vp, ip, vp, vp, ip]
    L.ibp_find_connections.restype = ip
    L.ibp_find_connections.argtypes = [vp, C.POINTER(PyParams), vp, ip, ip, ip, ip, ip, vp]
    L.ibp_fetch_connections.restype = ip
    L.ibp_fetch_connections.argtypes = [vp, ip, vp, vp, vp]
    L.ibp_set_connections.restype = ip
    L.ibp_set_connections.argtypes = [vp, ip, vp, vp, vp, ip]
    L.ibp_find_humans.restype = ip
    L.ibp_find_humans.argtypes = [vp, C.POINTER(PyParams), ip, vp]
    L.ibp_fetch_people.restype = ip
    L.ibp_fetch_people.argtypes = [vp, ip, vp, ip, vp]
    L.ibp_fetch_coco.restype = ip
    L.ibp_fetch_coco.argtypes = [vp, ip, C.POINTER(CocoView), vp]
    L.ibp_prepare_image.restype = ip
    L.ibp_prepare_image.argtypes = [vp, vp, ip, ip, ip, ip, ip, ip, vp, vp]
    L.ibp_resized_size.restype = None
    L.ibp_resized_size.argtypes = [ip, ip, C.c_double, C.c_double, C.POINTER(ip), C.POINTER(ip)]
    L.ibp_resize_image.restype = ip
    L.ibp_resize_image.argtypes = [vp, vp, ip, ip, ip, C.c_double, C.c_double, vp, ip, ip, vp]
    L.ibp_rotation_matrix.restype = None
    L.ibp_rotation_matrix.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double, C.POINTER(C.c_double)]
    L.ibp_warp_affine.restype = ip
    L.ibp_warp_affine.argtypes = [vp, vp, vp, ip, ip, ip, ip, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.POINTER(C.c_double), vp]
    L.ibp_crop_resize_accumulate.restype = ip
    L.ibp_crop_resize_accumulate.argtypes = [vp, vp, ip, ip, ip, ip, ip, ip, ip, ip, vp, vp]
    L.ibp_find_peaks_noalloc.restype = ip
    L.ibp_find_peaks_noalloc.argtypes = [vp, ip, ip, ip, ip, ip, vp, vp, vp, vp, vp]
    L.ibp_launch_count.restype = C.c_int64
    L.ibp_launch_count.argtypes = [vp]
    L.ibp_set_profiling.restype = ip
    L.ibp_set_profiling.argtypes = [vp, ip]
    L.ibp_get_stage_ms.restype = ip
    L.ibp_get_stage_ms.argtypes = [vp, fp]
    _lib = L
    return L


EXPORTS = ["ibp_process_paf", "ibp_get_num_humans", "ibp_get_part_peak_id", "ibp_get_score", "ibp_get_part_x",
           "ibp_get_part_y", "ibp_get_part_score", "ibp_get_status", "ibp_default_params", "ibp_default_limits",
           "ibp_create", "ibp_destroy", "ibp_last_error", "ibp_device_count", "ibp_fuse_flip", "ibp_fuse_multiscale", "ibp_upsample_limbs",
           "ibp_find_peaks", "ibp_set_peaks", "ibp_assign", "ibp_postprocess_batch", "ibp_fetch", "ibp_fetch_subpix", "ibp_default_py_params", "ibp_set_peaks64", "ibp_find_connections",
           "ibp_fetch_connections", "ibp_set_connections", "ibp_find_humans", "ibp_fetch_people", "ibp_launch_count", "ibp_set_profiling",
           "ibp_get_stage_ms", "ibp_fetch_coco", "ibp_prepare_image", "ibp_find_peaks_noalloc",
           "ibp_resized_size", "ibp_resize_image", "ibp_rotation_matrix", "ibp_warp_affine", "ibp_crop_resize_accumulate",
           "ibp_fuse_multiscale_linear", "ibp_smooth_gaussian"]
NOALLOC_MAX_BATCH, NOALLOC_MAX_BANDS, NOALLOC_CAP = 8, 8, 64  # include/ibp.h §5
STAGES = ["fuse", "peaks", "pack", "limbs", "group", "upsample", "smooth"]


def check(rc):
    if rc != IBP_OK:
        raise IbpError(rc, lib().ibp_last_error().decode("utf-8", "replace"))


def default_params():
    p = Params()
    lib().ibp_default_params(C.byref(p))
    return p


def py_params(test_cfg=None):
    """PyParams from a reference-style test_cfg dict (utils/config_reader.py:53-66); defaults = utils/config."""
    p = PyParams()
    lib().ibp_default_py_params(C.byref(p))
    if test_cfg:
        p.thre2 = float(test_cfg.get('thre2', p.thre2))
        p.connect_ration = float(test_cfg.get('connect_ration', p.connect_ration))
        p.mid_num = int(test_cfg.get('mid_num', p.mid_num))
        p.len_rate = int(test_cfg.get('len_rate', p.len_rate))
        p.connection_tole = float(test_cfg.get('connection_tole', p.connection_tole))
    return p


def resized_size(img_h, img_w, fx, fy):
    """(out_h, out_w) of cv2.resize(img, (0, 0), fx=fx, fy=fy): rint(h * fy), rint(w * fx)."""
    oh, ow = C.c_int(), C.c_int()
    lib().ibp_resized_size(int(img_h), int(img_w), float(fx), float(fy), C.byref(oh), C.byref(ow))
    return oh.value, ow.value


def rotation_matrix(cx, cy, angle_deg, scale=1.0):
    """cv2.getRotationMatrix2D((cx, cy), angle_deg, scale) -> (2,3) float64."""
    m = (C.c_double * 6)()
    lib().ibp_rotation_matrix(float(cx), float(cy), float(angle_deg), float(scale), m)
    return np.array(list(m), dtype=np.float64).reshape(2, 3)


def default_limits():
    l = Limits()
    lib().ibp_default_limits(C.byref(l))
    return l


class BatchResult:
    """Host copy of one ibp_fetch: plain numpy, detached from the handle's pinned buffers."""

    def __init__(self, view):
        B = view.batch
        self.batch = B
        self.n_peaks = np.ctypeslib.as_array(view.n_peaks, (B,)).copy()
        self.part_count = np.ctypeslib.as_array(view.part_count, (B, NUM_PARTS)).copy()
        self.n_humans = np.ctypeslib.as_array(view.n_humans, (B,)).copy()
        self.status = np.ctypeslib.as_array(view.status, (B,)).copy()
        pk = (C.c_char * (B * view.peak_stride * PEAK_DTYPE.itemsize)).from_address(view.peaks)
        hm = (C.c_char * (B * view.human_stride * HUMAN_DTYPE.itemsize)).from_address(view.humans)
        peaks = np.frombuffer(pk, dtype=PEAK_DTYPE).reshape(B, view.peak_stride)
        humans = np.frombuffer(hm, dtype=HUMAN_DTYPE).reshape(B, view.human_stride)
        self.peaks = [peaks[b, :self.n_peaks[b]].copy() for b in range(B)]
        self.humans = [humans[b, :self.n_humans[b]].copy() for b in range(B)]


class Handle:
    """Owner of one ibp_handle (device buffers + pinned result buffers) on one GPU."""

    def __init__(self, device=0, params=None, limits=None, **limit_overrides):
        L = lib()
        if L.ibp_device_count() <= 0:
            raise IbpError(-6, "no CUDA device visible: this library has no CPU path")
        self.params = params or default_params()
        self.limits = limits or default_limits()
        for k, v in limit_overrides.items():
            if not hasattr(self.limits, k):
                raise TypeError("unknown limit %r" % k)
            setattr(self.limits, k, v)
        self._h = C.c_void_p()
        check(L.ibp_create(C.byref(self._h), int(device), C.byref(self.params), C.byref(self.limits)))
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().ibp_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self):
        return int(lib().ibp_launch_count(self._h))

    def set_profiling(self, on=True):
        check(lib().ibp_set_profiling(self._h, int(bool(on))))

    def stage_ms(self):
        """{stage: ms} of the kernels of the last call, from CUDA events on the launching stream."""
        buf = (C.c_float * len(STAGES))()
        check(lib().ibp_get_stage_ms(self._h, buf))
        return {n: float(buf[i]) for i, n in enumerate(STAGES) if buf[i] >= 0}

    # ---- raw stage calls on device pointers (ints) ----
    def fuse_flip(self, net_out_ptr, dtype, batch, fh, fw, fused_ptr, flip_avg=True, stream=0):
        check(lib().ibp_fuse_flip(self._h, net_out_ptr, dtype, batch, fh, fw, int(bool(flip_avg)), fused_ptr, stream))

    def fuse_multiscale(self, scales, batch, img_h, img_w, out_ptr, heat32_ptr=None, flip_avg=True, stream=0, interp="cubic"):
        """scales: [(net_out_ptr, fh, fw, sh, sw)] per scale of the search.  interp: "cubic" = the reference's operator,
        "linear" = the second operator mode (ibp_fuse_multiscale_linear)."""
        arr = (ScaleInput * len(scales))(*[ScaleInput(int(p), int(fh), int(fw), int(sh), int(sw)) for p, fh, fw, sh, sw in scales])
        if interp not in ("cubic", "linear"):
            raise ValueError("interp must be 'cubic' or 'linear', got %r" % (interp,))
        fn = lib().ibp_fuse_multiscale if interp == "cubic" else lib().ibp_fuse_multiscale_linear
        check(fn(self._h, arr, len(scales), batch, img_h, img_w, int(bool(flip_avg)), out_ptr, heat32_ptr, stream))

    def smooth_gaussian(self, maps_ptr, out_ptr, planes, img_h, img_w, ksize, sigma, stream=0):
        check(lib().ibp_smooth_gaussian(self._h, maps_ptr, out_ptr, planes, img_h, img_w, int(ksize), float(sigma), stream))

    def upsample_limbs(self, fused_ptr, chan_stride, batch, fh, fw, up, out_ptr, stream=0):
        check(lib().ibp_upsample_limbs(self._h, fused_ptr, chan_stride, batch, fh, fw, up, out_ptr, stream))

    def find_peaks(self, maps_ptr, chan_stride, first_chan, batch, fh, fw, flavour=PEAKS_CROSS_REFINE, stream=0):
        check(lib().ibp_find_peaks(self._h, maps_ptr, chan_stride, first_chan, batch, fh, fw, flavour, stream))

    def set_peaks(self, image, joint_list):
        a = np.ascontiguousarray(joint_list, dtype=np.float32).reshape(-1, 5)
        check(lib().ibp_set_peaks(self._h, image, a.ctypes.data, a.shape[0], 5))

    def assign(self, limb_ptr, mode, batch, dim0, dim1, dim2, min_img_size, stream=0):
        check(lib().ibp_assign(self._h, limb_ptr, mode, batch, dim0, dim1, dim2, int(min_img_size), stream))

    def postprocess_batch(self, net_out_ptr, on_host, dtype, batch, fh, fw, min_img_size, stream=0):
        check(lib().ibp_postprocess_batch(self._h, net_out_ptr, int(bool(on_host)), dtype, batch, fh, fw,
                                          int(min_img_size), stream))

    def fetch_subpix(self, image, n, stream=0):
        """(n,2) float64 sub-pixel x, y of image `image` in peak-id order (centroid flavour), after fetch()."""
        out = np.zeros((max(int(n), 1), 2), dtype=np.float64)
        rc = lib().ibp_fetch_subpix(self._h, int(image), out.ctypes.data, out.shape[0], stream)
        if rc < 0:
            check(rc)
        return out[:rc]

    # ---- Python-semantics route (find_connections / find_humans) ----
    def set_peaks64(self, image, rows, part):
        r = np.ascontiguousarray(rows, dtype=np.float64).reshape(-1, 4)
        t = np.ascontiguousarray(part, dtype=np.int32).reshape(-1)
        check(lib().ibp_set_peaks64(self._h, int(image), r.ctypes.data, t.ctypes.data, r.shape[0]))

    def find_connections(self, params, paf_ptr, chan_stride, batch, img_h, img_w, img_height, stream=0):
        check(lib().ibp_find_connections(self._h, C.byref(params), paf_ptr, chan_stride, batch, img_h, img_w,
                                         int(img_height), stream))

    def fetch_connections(self, image, stream=0):
        """-> (rows (30, cap, 6) float64, counts (30,) int32)"""
        cap = self.limits.max_peaks_per_part
        rows = np.zeros((NUM_LIMBS, cap, 6), dtype=np.float64)
        counts = np.zeros(NUM_LIMBS, dtype=np.int32)
        check(lib().ibp_fetch_connections(self._h, int(image), rows.ctypes.data, counts.ctypes.data, stream))
        return rows, counts

    def set_connections(self, image, rows, counts, cand_score):
        cap = self.limits.max_peaks_per_part
        r = np.ascontiguousarray(rows, dtype=np.float64)
        assert r.shape == (NUM_LIMBS, cap, 6)
        c = np.ascontiguousarray(counts, dtype=np.int32)
        s = np.ascontiguousarray(cand_score, dtype=np.float64).reshape(-1)
        check(lib().ibp_set_connections(self._h, int(image), r.ctypes.data, c.ctypes.data, s.ctypes.data, s.shape[0]))

    def find_humans(self, params, batch, stream=0):
        check(lib().ibp_find_humans(self._h, C.byref(params), batch, stream))

    def fetch_people(self, image, stream=0):
        """-> person_to_joint_assoc (n, 20, 2) float64"""
        out = np.zeros((self.limits.max_humans, 20, 2), dtype=np.float64)
        n = lib().ibp_fetch_people(self._h, int(image), out.ctypes.data, out.shape[0], stream)
        if n < 0:
            check(n)
        return out[:n].copy()

    def fetch_coco(self, batch, stream=0):
        """-> list over images of (n, 52) float32 arrays: 17 x (x, y, v) in COCO order + score (evaluate.py:182-209)."""
        view = CocoView()
        check(lib().ibp_fetch_coco(self._h, batch, C.byref(view), stream))
        off = np.ctypeslib.as_array(view.offsets, (batch + 1,)).copy()
        total = int(off[batch])
        rows = np.ctypeslib.as_array(view.rows, (total, COCO_ROW)).copy() if total else np.zeros((0, COCO_ROW), np.float32)
        return [rows[off[b]:off[b + 1]] for b in range(batch)]

    def prepare_image(self, img_ptr, batch, img_h, img_w, max_downsample, pad_value, flip, out_ptr, stream=0):
        check(lib().ibp_prepare_image(self._h, img_ptr, batch, img_h, img_w, max_downsample, pad_value, 1 if flip else 0,
                                      out_ptr, stream))

    def resize_image(self, img_ptr, batch, img_h, img_w, fx, fy, out_ptr, out_h, out_w, stream=0):
        check(lib().ibp_resize_image(self._h, img_ptr, batch, img_h, img_w, float(fx), float(fy), out_ptr, out_h, out_w, stream))

    def warp_affine(self, src_ptr, dst_ptr, batch, channels, hh, ww, strides, matrix, stream=0):
        """strides = element strides (batch, channel, row, pixel); matrix = the forward 2x3 matrix as given to cv2.warpAffine."""
        m = (C.c_double * 6)(*[float(v) for v in np.asarray(matrix, dtype=np.float64).reshape(6)])
        check(lib().ibp_warp_affine(self._h, src_ptr, dst_ptr, batch, channels, hh, ww, int(strides[0]), int(strides[1]),
                                    int(strides[2]), int(strides[3]), m, stream))

    def crop_resize_accumulate(self, maps_ptr, batch, ph, pw, sh, sw, img_h, img_w, n_items, out_ptr, stream=0):
        check(lib().ibp_crop_resize_accumulate(self._h, maps_ptr, batch, ph, pw, sh, sw, img_h, img_w, int(n_items), out_ptr, stream))

    def fetch(self, batch, stream=0, raise_on_overflow=True):
        view = BatchView()
        rc = lib().ibp_fetch(self._h, batch, C.byref(view), stream)
        if rc != IBP_OK and (raise_on_overflow or rc != -3):
            check(rc)
        return BatchResult(view)
