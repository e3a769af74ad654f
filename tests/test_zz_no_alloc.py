"""ibp_find_peaks_noalloc (include/ibp.h §5; no counterpart in the reference):
the allocation-free variant of the heatmap_nms peak kernel, scratch in a file-scope static __device__ array.  Called through
the C ABI and held to the existing entry point (ibp_find_peaks + ibp_fetch, itself bit-exact against the oracle) on two
calls in a row — bit-exact, the tolerance every peak test here uses.  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _noalloc(maps_t, first_chan=0):
    """One call of the new entry point on a (B, C, h, w) float32 CUDA tensor -> host copies of its four outputs."""
    import torch
    from improved_body_parts_b200 import _lib
    L = _lib.lib()
    B, Cs, h, w = maps_t.shape
    dev = maps_t.device
    # outputs are the CALLER's: allocated here, before the call
    peaks = torch.full((B, _lib.NUM_PARTS * _lib.NOALLOC_CAP, 4), -7, dtype=torch.int32, device=dev)
    part_count = torch.full((B, _lib.NUM_PARTS), -7, dtype=torch.int32, device=dev)
    n_peaks = torch.full((B,), -7, dtype=torch.int32, device=dev)
    status = torch.full((B,), 0x7FFFFFFF, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = L.ibp_find_peaks_noalloc(maps_t.data_ptr(), Cs, first_chan, B, h, w, peaks.data_ptr(), part_count.data_ptr(),
                                      n_peaks.data_ptr(), status.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(rc)
    torch.cuda.synchronize(dev)
    rows = peaks.cpu().numpy().view(_lib.PEAK_DTYPE).reshape(B, -1)
    return rows, part_count.cpu().numpy(), n_peaks.cpu().numpy(), status.cpu().numpy()


def _existing(pp, maps_t, first_chan=0):
    pp.find_peaks(maps_t, first_chan)
    return pp.fetch(maps_t.shape[0])


def _heat_batch(seeds_people, height=512, width=512):
    import torch
    from improved_body_parts_b200 import synth
    from oracle import np_oracle as O
    planes = []
    for seed, people in seeds_people:
        out, _ = synth.make_network_output(seed, people, height=height, width=width)
        heat, _ = O.fuse_flip(out)
        planes.append(np.ascontiguousarray(heat[:, :, :18].transpose(2, 0, 1)))
    return torch.from_numpy(np.stack(planes)).cuda()


def _compare(got, want, B):
    rows, part_count, n_peaks, status = got
    assert np.array_equal(n_peaks, want.n_peaks[:B])
    assert np.array_equal(part_count, want.part_count[:B])
    for b in range(B):
        n = int(n_peaks[b])
        assert np.array_equal(rows[b, :n], want.peaks[b]), b        # x, y, score, part: identical bits
        assert status[b] == want.status[b] == 0


def test_noalloc_matches_existing_entry_point_two_calls_in_a_row():
    from improved_body_parts_b200 import _lib
    from improved_body_parts_b200.pipeline import PostProcessor
    L = _lib.lib()
    # set-up-only call (what a caller does before capturing a graph): nothing to compute, nothing allocated
    assert L.ibp_find_peaks_noalloc(None, 18, 0, 0, 128, 128, None, None, None, None, None) == 0
    pp = PostProcessor(device=0, max_batch=8, max_peaks_per_part=_lib.NOALLOC_CAP)
    first = _heat_batch([(9301, 9), (9302, 1), (9303, 6), (9304, 10), (9305, 4)])
    second = _heat_batch([(9311, 2), (9312, 7), (9313, 0)])          # fewer images, other peaks: stale scratch must not leak
    want1, want2 = _existing(pp, first), _existing(pp, second)
    assert want1.n_peaks.sum() > 100 and want2.n_peaks[2] == 0
    _compare(_noalloc(first), want1, 5)
    _compare(_noalloc(second), want2, 3)
    _compare(_noalloc(first), want1, 5)                               # and back again
    # another map size (2 bands of 24 rows, a width that is not a power of two), keypoints at a channel offset
    import torch
    odd = _heat_batch([(9321, 3), (9322, 5)], height=192, width=416)  # 48 x 104 maps
    padded = torch.cat([torch.zeros_like(odd[:, :3]), odd, torch.zeros_like(odd[:, :2])], dim=1).contiguous()
    _compare(_noalloc(padded, first_chan=3), _existing(pp, padded, 3), 2)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    from improved_body_parts_b200 import _lib
    L = _lib.lib()
    t = torch.zeros((1, 18, 8, 8), dtype=torch.float32, device="cuda")
    o = torch.zeros(4096, dtype=torch.int32, device="cuda")
    args = lambda batch, fh, fw: (t.data_ptr(), 18, 0, batch, fh, fw, o.data_ptr(), o.data_ptr(), o.data_ptr(), o.data_ptr(), None)
    # validation happens before anything is launched: the (too small) buffers are never touched
    assert L.ibp_find_peaks_noalloc(*args(_lib.NOALLOC_MAX_BATCH + 1, 8, 8)) == -5
    assert b"IBP_NOALLOC_MAX_BATCH" in L.ibp_last_error()
    assert L.ibp_find_peaks_noalloc(*args(1, 1024, 128)) == -1        # 32 bands of 32 rows > IBP_NOALLOC_MAX_BANDS
    assert b"IBP_NOALLOC_MAX_BANDS" in L.ibp_last_error()
    assert L.ibp_find_peaks_noalloc(*args(1, 8, 4096)) == -1
    assert L.ibp_find_peaks_noalloc(t.data_ptr(), 18, 0, 1, 8, 8, None, None, None, None, None) == -5   # null outputs
    torch.cuda.synchronize()
