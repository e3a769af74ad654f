"""Drop-in for the reference's SWIG module `utils.pafprocess.pafprocess` (utils/pafprocess/pafprocess.i:1-16,
pafprocess.h:70-76): same seven names, same argument order, same return types.  evaluate.py:110-127 and
demo_image.py:131-148 run unchanged with

    from improved_body_parts_b200 import pafprocess

Work is done by libibp_b200.so on the GPU (ibp_process_paf and the six getters of include/ibp.h); there is no CPU
path.  Like the reference the results live in process-global state until the next process_paf call, so the module
is not re-entrant.  Array handling follows the numpy.i IN_ARRAY3 typemap (pafprocess.i:14): any array-like that
converts to a 3-D contiguous float32 array is accepted, anything else raises TypeError / ValueError.  On top of that
 both arguments may be CUDA tensors (anything with `__cuda_array_interface__`, e.g. torch): a limb map
that already lives on the GPU is sampled in place — no host copy at all — and a pinned host array is read by DMA
without staging.
"""
import numpy as np

from . import _lib


def _in_array3(obj, name):
    try:
        a = np.ascontiguousarray(obj, dtype=np.float32)
    except (TypeError, ValueError) as e:
        raise TypeError("%s: cannot convert to a float32 array (%s)" % (name, e))
    if a.ndim != 3:
        raise ValueError("%s: array must have 3 dimensions, given array has %d dimensions" % (name, a.ndim))
    return a


def _cuda_array3(obj, name):
    """(keep-alive object, device pointer, shape) of a 3-D float32 C-contiguous CUDA array."""
    if hasattr(obj, "is_cuda"):  # torch: make it float32 + contiguous on the device if it is not
        import torch
        obj = obj.detach().to(torch.float32).contiguous()
    cai = obj.__cuda_array_interface__
    shape = tuple(int(v) for v in cai["shape"])
    if len(shape) != 3:
        raise ValueError("%s: array must have 3 dimensions, given array has %d dimensions" % (name, len(shape)))
    if cai["typestr"] not in ("<f4", "=f4", "|f4") or cai.get("strides") is not None:
        raise TypeError("%s: CUDA arrays must be C-contiguous float32" % name)
    return obj, int(cai["data"][0]), shape


def _is_cuda(obj):
    if hasattr(obj, "is_cuda"):
        return bool(obj.is_cuda)
    return hasattr(obj, "__cuda_array_interface__")


def process_paf(peaks, pafmap, min_img_size):
    """pafprocess.h:70.  peaks (p1,p2,5) rows [x,y,score,id,part]; pafmap (H,W,30) HWC; returns 0."""
    if _is_cuda(peaks):  # the joint list is bucketed on the host (pafprocess.cpp:29-41 is a serial append): a few KB
        peaks = peaks.detach().cpu().numpy() if hasattr(peaks, "detach") else np.asarray(peaks.get())
    p = _in_array3(peaks, "peaks")
    if _is_cuda(pafmap):
        keep, ptr, shape = _cuda_array3(pafmap, "pafmap")
        if hasattr(keep, "device"):  # work queued on the tensor's stream must be visible to the library's stream
            import torch
            torch.cuda.current_stream(keep.device).synchronize()
    else:
        keep = _in_array3(pafmap, "pafmap")
        ptr, shape = keep.ctypes.data, keep.shape
    rc = _lib.lib().ibp_process_paf(p.shape[0], p.shape[1], p.shape[2], p.ctypes.data, shape[0], shape[1],
                                    shape[2], ptr, int(min_img_size))
    _lib.check(rc)
    return rc


def get_num_humans():
    return int(_lib.lib().ibp_get_num_humans())


def get_part_peak_id(skeleton_id, part_id):
    return int(_lib.lib().ibp_get_part_peak_id(int(skeleton_id), int(part_id)))


def get_score(skeleton_id):
    return float(_lib.lib().ibp_get_score(int(skeleton_id)))


def get_part_x(cid):
    return int(_lib.lib().ibp_get_part_x(int(cid)))


def get_part_y(cid):
    return int(_lib.lib().ibp_get_part_y(int(cid)))


def get_part_score(cid):
    return float(_lib.lib().ibp_get_part_score(int(cid)))


def get_status():
    """IBP_ST_* bits of the last call (no reference counterpart)."""
    return int(_lib.lib().ibp_get_status())
