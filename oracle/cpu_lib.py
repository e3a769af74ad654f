"""TEST INFRASTRUCTURE ONLY — ctypes bindings for the checker libraries.

`Oracle()`    -> oracle/libibp_oracle.so (our C++ restatement, always buildable)
`Reference()` -> oracle/_ref/libpafprocess_ref.so (the unmodified reference, prebuilt in the build container)
Both expose the reference's seven-call protocol (pafprocess.h:70-76) under the same method names.
"""
import ctypes as C
import os

import numpy as np

from . import build_oracle as _b

_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)


def _as_f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class _SevenCalls:
    """process_paf + the six getters, bound to a prefix ('oracle_' / 'ref_')."""

    def __init__(self, lib, prefix):
        self.lib = lib
        g = lambda n: getattr(lib, prefix + n)
        self._process = g("process_paf")
        self._process.restype = C.c_int
        self._process.argtypes = [C.c_int, C.c_int, C.c_int, _fp, C.c_int, C.c_int, C.c_int, _fp, C.c_int]
        self.get_num_humans = g("get_num_humans")
        self.get_num_humans.restype = C.c_int
        self.get_part_peak_id = g("get_part_peak_id")
        self.get_part_peak_id.restype = C.c_int
        self.get_part_peak_id.argtypes = [C.c_int, C.c_int]
        self.get_score = g("get_score")
        self.get_score.restype = C.c_float
        self.get_score.argtypes = [C.c_int]
        self.get_part_x = g("get_part_x")
        self.get_part_x.restype = C.c_int
        self.get_part_x.argtypes = [C.c_int]
        self.get_part_y = g("get_part_y")
        self.get_part_y.restype = C.c_int
        self.get_part_y.argtypes = [C.c_int]
        self.get_part_score = g("get_part_score")
        self.get_part_score.restype = C.c_float
        self.get_part_score.argtypes = [C.c_int]

    def process_paf(self, peaks, pafmap, min_img_size):
        """Same call shape as the SWIG module: peaks (1,N,5) f32, pafmap (H,W,30) f32 (pafprocess.i:14)."""
        p = _as_f32(peaks)
        m = _as_f32(pafmap)
        assert p.ndim == 3 and m.ndim == 3
        self._keep = (p, m)
        return self._process(p.shape[0], p.shape[1], p.shape[2], p.ctypes.data_as(_fp),
                             m.shape[0], m.shape[1], m.shape[2], m.ctypes.data_as(_fp), int(min_img_size))

    def humans(self):
        """[(ids[18] int, score float32)] through the getter protocol (evaluate.py:111-129)."""
        out = []
        for h in range(self.get_num_humans()):
            ids = np.array([self.get_part_peak_id(h, p) for p in range(18)], dtype=np.int32)
            out.append((ids, np.float32(self.get_score(h))))
        return out


class Oracle(_SevenCalls):
    def __init__(self):
        lib = C.CDLL(_b.build_oracle())
        super().__init__(lib, "oracle_")
        lib.oracle_process_paf_feat.restype = C.c_int
        lib.oracle_process_paf_feat.argtypes = [C.c_int, C.c_int, C.c_int, _fp, C.c_int, C.c_int, _fp, C.c_int,
                                                C.c_int]
        lib.oracle_get_link.argtypes = [C.c_int, C.c_int, _ip, _ip, _fp, _fp]
        lib.oracle_get_human_raw.argtypes = [C.c_int, _ip, _fp, _fp, _ip, _fp]
        lib.oracle_cubic_sample.restype = C.c_float
        lib.oracle_cubic_sample.argtypes = [_fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]

    def process_paf_feat(self, peaks, limb_planes, up, min_img_size):
        """limb_planes (30,h,w) planar feature-resolution maps; values of cv2.resize(x up, CUBIC) are sampled lazily."""
        p = _as_f32(peaks)
        m = _as_f32(limb_planes)
        assert p.ndim == 3 and m.ndim == 3 and m.shape[0] == 30
        return self.lib.oracle_process_paf_feat(p.shape[0], p.shape[1], p.shape[2], p.ctypes.data_as(_fp),
                                                m.shape[1], m.shape[2], m.ctypes.data_as(_fp), int(up),
                                                int(min_img_size))

    def links(self, limb):
        out = []
        for k in range(self.lib.oracle_num_links(limb)):
            a, b, s, l = C.c_int(), C.c_int(), C.c_float(), C.c_float()
            self.lib.oracle_get_link(limb, k, C.byref(a), C.byref(b), C.byref(s), C.byref(l))
            out.append((a.value, b.value, np.float32(s.value), np.float32(l.value)))
        return out

    def num_candidates(self, limb):
        return self.lib.oracle_num_candidates(limb)

    def has_duplicate_keys(self, limb):
        return bool(self.lib.oracle_has_duplicate_keys(limb))

    def human_raw(self, h):
        ids = np.zeros(18, dtype=np.int32)
        sc = np.zeros(18, dtype=np.float32)
        total, count, reach = C.c_float(), C.c_int(), C.c_float()
        self.lib.oracle_get_human_raw(h, ids.ctypes.data_as(_ip), sc.ctypes.data_as(_fp), C.byref(total),
                                      C.byref(count), C.byref(reach))
        return ids, sc, np.float32(total.value), count.value, np.float32(reach.value)

    def cubic_sample(self, plane, up, X, Y):
        p = _as_f32(plane)
        return np.float32(self.lib.oracle_cubic_sample(p.ctypes.data_as(_fp), p.shape[0], p.shape[1], up, X, Y))


class Reference(_SevenCalls):
    def __init__(self):
        path = _b.build_ref()
        if path is None:
            raise FileNotFoundError("oracle/_ref/libpafprocess_ref.so is absent and /root/reference is not here")
        super().__init__(C.CDLL(path), "ref_")


def reference_available():
    return _b.build_ref() is not None
