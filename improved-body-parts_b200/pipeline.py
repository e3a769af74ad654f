"""Batched, device-resident post-processing: everything evaluate.py:75-129 does after the model call, for B images
per launch set, plus the multi-GPU sharding helpers (images shard, no collective on the data path).

PyTorch is used for device memory, streams and (optionally) torch.distributed rendezvous only.
"""
import numpy as np
import torch

from . import _lib
from .common import humans_from_result
from .config import ORDER_COCO

_DTYPES = {torch.float32: _lib.IBP_F32, torch.float16: _lib.IBP_F16}


def _stream_ptr(device):
    return torch.cuda.current_stream(device).cuda_stream


class PostProcessor:
    """One instance per GPU (and per stream / thread).  `net_out` is what `NetworkEval` returns for the last stack at
    scale 0 on the batch [img, flip(img)] (models/posenet.py:184-202, parse_skeletons.py:76-80): (B, 2, 50, h, w)."""

    def __init__(self, device=0, max_batch=64, max_feat=(128, 128), **limit_overrides):
        if not torch.cuda.is_available():
            raise _lib.IbpError(-6, "no CUDA device: improved-body-parts_b200 has no CPU fallback")
        self.device = torch.device("cuda", device) if not isinstance(device, torch.device) else device
        self.handle = _lib.Handle(self.device.index or 0, max_batch=max_batch, max_feat_h=max_feat[0],
                                  max_feat_w=max_feat[1], **limit_overrides)
        self.max_batch = max_batch

    # ---- full pipeline -------------------------------------------------------------------------------------
    def enqueue(self, net_out, min_img_size):
        """Launch the pipeline on the current stream.  net_out: CUDA tensor, pinned/pageable CPU tensor or numpy."""
        if isinstance(net_out, np.ndarray):
            net_out = torch.from_numpy(net_out)
        if net_out.dim() != 5 or net_out.shape[1] != 2 or net_out.shape[2] != _lib.NUM_LAYERS:
            raise ValueError("net_out must be (B, 2, 50, h, w), got %s" % (tuple(net_out.shape),))
        if net_out.dtype not in _DTYPES:
            raise TypeError("net_out dtype must be float32 or float16")
        net_out = net_out.contiguous()
        B, _, _, h, w = net_out.shape
        on_host = not net_out.is_cuda
        with torch.cuda.device(self.device):
            self.handle.postprocess_batch(net_out.data_ptr(), on_host, _DTYPES[net_out.dtype], B, h, w, min_img_size,
                                          _stream_ptr(self.device))
        self._keep = net_out  # keep the source alive until fetch
        return B

    def fetch(self, batch, raise_on_overflow=True):
        with torch.cuda.device(self.device):
            res = self.handle.fetch(batch, _stream_ptr(self.device), raise_on_overflow)
        self._keep = None
        return res

    def __call__(self, net_out, min_img_size):
        return self.fetch(self.enqueue(net_out, min_img_size))

    def fetch_coco(self, batch):
        """Instead of fetch(): the humans of the batch as COCO rows built on the device (evaluate.py:182-209), one
        compacted copy.  -> list over images of (n, 52) float32: 17 x (x, y, v) in ORDER_COCO + score."""
        with torch.cuda.device(self.device):
            rows = self.handle.fetch_coco(batch, _stream_ptr(self.device))
        self._keep = None
        return rows

    def prepare_images(self, images, max_downsample=64, pad_value=128, flip=True):
        """utils/parse_skeletons.py:52-72 after the resize, on the device: uint8 (B,h,w,3) or (h,w,3) HWC -> float32
        (B*(2 if flip else 1), hp, wp, 3) CUDA tensor [img0, flip(img0), img1, ...], padded right/down with `pad_value`
        to multiples of `max_downsample` (utils/util.py:44-65) and divided by 255 like np.float32(img / 255)."""
        if isinstance(images, np.ndarray):
            images = torch.from_numpy(np.ascontiguousarray(images))
        if images.dim() == 3:
            images = images[None, ...]
        if images.dtype != torch.uint8 or images.dim() != 4 or images.shape[3] != 3:
            raise ValueError("images must be uint8 (B, h, w, 3), got %s %s" % (images.dtype, tuple(images.shape)))
        images = images.to(self.device).contiguous()
        B, h, w, _ = images.shape
        hp = h if h == 1 else -(-h // max_downsample) * max_downsample  # util.py:59-62: a dimension of 1 stays unpadded
        wp = w if w == 1 else -(-w // max_downsample) * max_downsample
        out = torch.empty((B * (2 if flip else 1), hp, wp, 3), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self.handle.prepare_image(images.data_ptr(), B, h, w, max_downsample, pad_value, flip, out.data_ptr(),
                                      _stream_ptr(self.device))
        return out

    def resize_images(self, images, fx, fy=None):
        """cv2.resize(img, (0, 0), fx=fx, fy=fy, interpolation=cv2.INTER_CUBIC) for uint8 (B,h,w,3) / (h,w,3) HWC images on
        the device (utils/parse_skeletons.py:204, :52) -> uint8 (B, rint(h*fy), rint(w*fx), 3) CUDA tensor."""
        fy = fx if fy is None else fy
        if isinstance(images, np.ndarray):
            images = torch.from_numpy(np.ascontiguousarray(images))
        if images.dim() == 3:
            images = images[None, ...]
        if images.dtype != torch.uint8 or images.dim() != 4 or images.shape[3] != 3:
            raise ValueError("images must be uint8 (B, h, w, 3), got %s %s" % (images.dtype, tuple(images.shape)))
        images = images.to(self.device).contiguous()
        B, h, w, _ = images.shape
        oh, ow = _lib.resized_size(h, w, fx, fy)
        out = torch.empty((B, oh, ow, 3), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            self.handle.resize_image(images.data_ptr(), B, h, w, fx, fy, out.data_ptr(), oh, ow, _stream_ptr(self.device))
        return out

    def warp_affine(self, t, matrix, layout="planar"):
        """cv2.warpAffine(x, matrix, (0, 0)) (INTER_LINEAR, constant 0 border, same size) on a float32 CUDA tensor:
        layout "planar" = (B,C,h,w) maps, "hwc" = (B,h,w,C) images (utils/parse_skeletons.py:215-218, :265-267, :98-100)."""
        t = t.to(self.device, torch.float32).contiguous()
        out = torch.empty_like(t)
        if layout == "planar":
            B, Cn, h, w = t.shape
            strides = (Cn * h * w, h * w, w, 1)
        else:
            B, h, w, Cn = t.shape
            strides = (h * w * Cn, 1, w * Cn, Cn)
        with torch.cuda.device(self.device):
            self.handle.warp_affine(t.data_ptr(), out.data_ptr(), B, Cn, h, w, strides, matrix, _stream_ptr(self.device))
        return out

    def crop_resize_accumulate(self, maps, scaled_size, out, n_items):
        """maps (B,50,ph,pw) float32 x4 maps at the padded size -> cropped to scaled_size, cubic-resized to out's (H,W),
        / n_items in float32, added to out (B,50,H,W) float64 (utils/parse_skeletons.py:272-281)."""
        maps = maps.to(self.device, torch.float32).contiguous()
        B, _, ph, pw = maps.shape
        with torch.cuda.device(self.device):
            self.handle.crop_resize_accumulate(maps.data_ptr(), B, ph, pw, int(scaled_size[0]), int(scaled_size[1]), out.shape[2],
                                               out.shape[3], n_items, out.data_ptr(), _stream_ptr(self.device))
        return out

    # ---- stages on tensors (used by the parse_skeletons mirror and by tests) ---------------------------------
    def fuse_flip(self, net_out, flip_avg=True):
        """(B,2,50,h,w) CUDA tensor -> (B,50,h,w) float32 CUDA tensor (K1)."""
        net_out = net_out.to(self.device).contiguous()
        B, _, _, h, w = net_out.shape
        fused = torch.empty((B, _lib.NUM_LAYERS, h, w), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self.handle.fuse_flip(net_out.data_ptr(), _DTYPES[net_out.dtype], B, h, w, fused.data_ptr(), flip_avg,
                                  _stream_ptr(self.device))
        return fused

    def fuse_multiscale(self, net_outs, scaled_sizes, img_h, img_w, flip_avg=True, want_heat32=True, interp="cubic"):
        """predict()'s fusion (utils/parse_skeletons.py:232-281) for a batch: net_outs = one (B,2,50,h_s,w_s) float32
        CUDA tensor per scale, scaled_sizes = [(sh,sw)] the scaled image sizes before padding.
        -> (fused (B,50,img_h,img_w) float64 planar [0,30) limb / [30,50) keypoint, heat32 (B,18,img_h,img_w) or None).
        interp="linear": the second operator mode (bilinear in both resizes; not the reference's operator)."""
        ts = [t.to(self.device, torch.float32).contiguous() for t in net_outs]
        B = ts[0].shape[0]
        out = torch.empty((B, _lib.NUM_LAYERS, img_h, img_w), dtype=torch.float64, device=self.device)
        heat32 = torch.empty((B, _lib.NUM_PARTS, img_h, img_w), dtype=torch.float32, device=self.device) if want_heat32 else None
        scales = [(t.data_ptr(), t.shape[3], t.shape[4], sh, sw) for t, (sh, sw) in zip(ts, scaled_sizes)]
        with torch.cuda.device(self.device):
            self.handle.fuse_multiscale(scales, B, img_h, img_w, out.data_ptr(), heat32.data_ptr() if want_heat32 else None,
                                        flip_avg, _stream_ptr(self.device), interp)
        self._keep_ms = ts
        return out, heat32

    def smooth_gaussian(self, maps, ksize=5, sigma=1.0):
        """Second operator mode: GaussianSmoothing(C, ksize, sigma) of utils/util.py:103-174 on F.pad(maps, reflect) —
        (B,C,H,W) float32 planar CUDA tensor -> smoothed tensor of the same shape."""
        maps = maps.to(self.device, torch.float32).contiguous()
        B, Cn, H, W = maps.shape
        out = torch.empty_like(maps)
        with torch.cuda.device(self.device):
            self.handle.smooth_gaussian(maps.data_ptr(), out.data_ptr(), B * Cn, H, W, ksize, sigma, _stream_ptr(self.device))
        return out

    def original_route(self, net_outs, scaled_sizes, img_h, img_w, test_cfg=None, flip_avg=True):
        """The reference's ORIGINAL route for a batch, device resident end to end (evaluate.py:82-90 without flags):
        predict (scale search + flip fusion at image resolution) -> find_peaks (3x3 NMS, centroid refinement) ->
        find_connections -> find_humans (Python semantics, float64).  The handle's peak_thresh / refine_window must be
        the route's thre1 / offset_radius (both equal the defaults).  Returns a list per image of
        (all_peaks table (N,4) float64 [x,y,score,id], part_count (18,), person_to_joint_assoc (P,20,2) float64)."""
        B = self.enqueue_original_route(net_outs, scaled_sizes, img_h, img_w, test_cfg, flip_avg)
        return self.fetch_original_route(B)

    def enqueue_original_route(self, net_outs, scaled_sizes, img_h, img_w, test_cfg=None, flip_avg=True, interp="cubic", smooth=None):
        """Launch the four kernels of the original route on the current stream; results stay on the device.
        interp="linear" and smooth=(ksize, sigma) select the second operator mode (bilinear fusion, Gaussian-smoothed
        keypoint maps before the 3x3 max-suppression); the defaults are the reference's operators."""
        fused, heat32 = self.fuse_multiscale(net_outs, scaled_sizes, img_h, img_w, flip_avg, want_heat32=True, interp=interp)
        B = fused.shape[0]
        if smooth is not None:
            heat32 = self.smooth_gaussian(heat32, smooth[0], smooth[1])
        params = _lib.py_params(test_cfg)
        with torch.cuda.device(self.device):
            st = _stream_ptr(self.device)
            self.handle.find_peaks(heat32.data_ptr(), _lib.NUM_PARTS, 0, B, img_h, img_w, _lib.PEAKS_BOX_CENTROID, st)
            self.handle.find_connections(params, fused.data_ptr(), _lib.NUM_LAYERS, B, img_h, img_w, img_h, st)
            self.handle.find_humans(params, B, st)
        self._keep_or = (fused, heat32)
        return B

    def fetch_original_route(self, B):
        with torch.cuda.device(self.device):
            st = _stream_ptr(self.device)
            res = self.handle.fetch(B, st)
            out = []
            for b in range(B):
                n = int(res.n_peaks[b])
                xy = self.handle.fetch_subpix(b, n, st)
                table = np.zeros((n, 4), dtype=np.float64)
                table[:, :2] = xy
                table[:, 2] = res.peaks[b]["score"]
                table[:, 3] = np.arange(n)
                out.append((table, res.part_count[b].copy(), self.handle.fetch_people(b, st)))
        return out

    def upsample_limbs(self, fused, up=4):
        """(B,C>=30,h,w) planar -> (B, h*up, w*up, 30) HWC float32 (A4, evaluate.py:77-80)."""
        fused = fused.to(self.device, torch.float32).contiguous()
        B, Cs, h, w = fused.shape
        out = torch.empty((B, h * up, w * up, _lib.NUM_LIMBS), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self.handle.upsample_limbs(fused.data_ptr(), Cs, B, h, w, up, out.data_ptr(), _stream_ptr(self.device))
        return out

    def find_peaks(self, maps, first_chan, flavour=_lib.PEAKS_CROSS_REFINE):
        """maps (B,C,h,w) planar float32 CUDA tensor; peaks stay on the device for assign()/fetch()."""
        maps = maps.to(self.device, torch.float32).contiguous()
        B, Cs, h, w = maps.shape
        with torch.cuda.device(self.device):
            self.handle.find_peaks(maps.data_ptr(), Cs, first_chan, B, h, w, flavour, _stream_ptr(self.device))
        self._keep = maps
        return B

    def assign_feat(self, limb_planes, min_img_size):
        """limb_planes (B,C>=30,h,w) planar feature-resolution maps, bicubic evaluated at the samples."""
        t = limb_planes.to(self.device, torch.float32).contiguous()
        B, Cs, h, w = t.shape
        with torch.cuda.device(self.device):
            self.handle.assign(t.data_ptr(), _lib.LIMB_FEAT_PLANAR, B, h, w, Cs, min_img_size,
                               _stream_ptr(self.device))
        self._keep2 = t

    def assign_fullres(self, pafmap_hwc, min_img_size):
        """pafmap_hwc (B,H,W,C>=30) float32 CUDA tensor in the reference layout."""
        t = pafmap_hwc.to(self.device, torch.float32).contiguous()
        B, H, W, Cc = t.shape
        with torch.cuda.device(self.device):
            self.handle.assign(t.data_ptr(), _lib.LIMB_FULLRES_HWC, B, H, W, Cc, min_img_size,
                               _stream_ptr(self.device))
        self._keep2 = t


# ---- result conversions ------------------------------------------------------------------------------------------
class PipelinedPostProcessor:
    """Several batches in flight on one GPU.  The limb kernel is a latency chain with a long, thin tail (the grouping
    of the last images) and the peak kernel a bandwidth kernel: with `depth` independent PostProcessors — each with its
    own handle (scratch, result buffers) and its own CUDA stream — the tail of one batch runs under the head of the
    next, and with host inputs the upload of a batch runs under the kernels and the read-back of the previous ones.
    Batches are independent, results come back in submission order.

        pipe = PipelinedPostProcessor(device=0, depth=3, max_batch=64)
        for net_out in stream_of_batches:
            pipe.submit(net_out, img_h)
            if pipe.full():
                handle_result(pipe.collect())
        for res in pipe.drain():
            handle_result(res)
    """

    def __init__(self, device=0, depth=3, max_batch=64, max_feat=(128, 128), **limit_overrides):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.lanes = [PostProcessor(device, max_batch, max_feat, **limit_overrides) for _ in range(depth)]
        self.device = self.lanes[0].device
        self.streams = [torch.cuda.Stream(self.device) for _ in range(depth)]
        self.depth = depth
        self._pending = []  # [(lane, batch size)] oldest first
        self._next = 0

    def full(self):
        return len(self._pending) == self.depth

    def submit(self, net_out, min_img_size, replace=False):
        """Launch a batch on the next lane (its stream first waits for the caller's current stream, where `net_out` was
        produced).  A lane whose previous result was not collected refuses, unless `replace` drops that result."""
        k = self._next
        busy = [i for i, (lane, _) in enumerate(self._pending) if lane == k]
        if busy:
            if not replace:
                raise RuntimeError("every lane holds an uncollected batch: collect() before submitting more")
            del self._pending[busy[0]]
        if isinstance(net_out, np.ndarray):
            net_out = torch.from_numpy(net_out)
        st = self.streams[k]
        st.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(st):
            B = self.lanes[k].enqueue(net_out, min_img_size)
        if net_out.is_cuda:
            net_out.record_stream(st)  # the caller may drop its reference before the lane has read the tensor
        self._pending.append((k, B))
        self._next = (k + 1) % self.depth
        return B

    def collect(self, coco=False):
        """Result of the oldest batch in flight (waits for that lane's stream only): a BatchResult, or with coco=True the
        COCO rows built on the device (PostProcessor.fetch_coco)."""
        if not self._pending:
            raise RuntimeError("nothing in flight")
        k, B = self._pending.pop(0)
        with torch.cuda.stream(self.streams[k]):
            return self.lanes[k].fetch_coco(B) if coco else self.lanes[k].fetch(B)

    def drain(self, coco=False):
        return [self.collect(coco) for _ in range(len(self._pending))]

    def synchronize(self):
        """The caller's current stream waits for everything in flight (results stay collectable)."""
        for st in self.streams:
            torch.cuda.current_stream(self.device).wait_stream(st)

    @property
    def launches(self):
        return sum(lane.handle.launches for lane in self.lanes)


def to_humans(result):
    """BatchResult -> list (per image) of [Human] exactly as evaluate.py:111-129 builds them."""
    return [humans_from_result(result.peaks[b], result.humans[b]) for b in range(result.batch)]


def to_coco_rows(result, b):
    """evaluate.py:182-209 for image b: [(keypoints[17*3] CMU-18 -> COCO-17 via ORDER_COCO, score)]."""
    rows = []
    peaks = result.peaks[b]
    for rec in result.humans[b]:
        kp = np.zeros((18, 3), dtype=np.float32)
        for part in range(18):
            pid = int(rec["peak_id"][part])
            if pid >= 0:
                kp[part] = (peaks[pid]["x"], peaks[pid]["y"], 1)
        rows.append((kp[ORDER_COCO].reshape(-1), float(rec["score"])))
    return rows


# ---- multi-GPU: images shard, no data-path collective ---------------------------------------------
def shard_indices(n_images, rank, world):
    """rank r of W takes images[r::W]."""
    return list(range(rank, n_images, world))


def merge_shards(shards, n_images):
    """shards: list over ranks of [(image_index, payload)] -> list of payloads ordered by image index, independent of
    the world size."""
    out = [None] * n_images
    for part in shards:
        for idx, payload in part:
            out[idx] = payload
    if any(o is None for o in out):
        raise ValueError("missing image results after merge")
    return out


def gather_results(local, n_images):
    """Host-side gather of per-image results over torch.distributed (any backend; the payload is a few KB per
    image, so this is an object gather, not a tensor collective).  Every rank returns the full ordered list."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return merge_shards([local], n_images)
    buckets = [None] * dist.get_world_size()
    dist.all_gather_object(buckets, local)
    return merge_shards(buckets, n_images)
