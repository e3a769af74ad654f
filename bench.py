#!/usr/bin/env python
"""bench.py — post-processing images/s of the keypoint-assignment path (BASELINE.json metric) on N B200s.

A step = one pass of the hot path (flip fusion -> peaks -> limb scoring -> grouping) over one batch of synthetic
network outputs.  Workload at every N: BASELINE.json configs[1] ("synthetic batch 64, 512x512 input, stride-4
128x128 maps, 50 channels"), per rank (weak scaling; images shard, no data-path collective — SURVEY.md §8e).

    python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (under torchrun when N > 1)
    python bench.py --impl reference [--gpus N] ...                the reference's CPU route on the host cores

One JSON line on stdout (rank 0).  See DESIGN.md §6 for what each key means.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

BATCH = 64
FEAT = 128
IMG = 512
WORKLOAD = "configs[1]: synthetic batch 64, 512x512 input, stride-4 128x128x50 maps, flip twin, 1-10 people/image"
METRIC = "post-proc images/sec"
UNIT = "images/s"
SEED = 2000  # 1000 * config_id + image index
# algorithmic bytes per image at 128x128 fp32 (SURVEY.md §8d / DESIGN.md §5)
ALGO_BYTES = {
    "peaks": 2 * 18 * FEAT * FEAT * 4,         # keypoint planes of the upright and the flipped output, read once
    "limbs": 2 * 30 * FEAT * FEAT * 4,         # limb planes of the upright and the flipped output, read once
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="images in the cpu_baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="configs1", choices=["configs1", "configs2", "configs3"],
                    help="configs1 (default, the headline): batch 64 of 128x128x50 maps, refactored + C++ route; "
                         "configs2: flip + 4-scale fusion to 640x480, original route; configs3: crowd 40-60 people")
    ap.add_argument("--depth", type=int, default=3, help="configs1/3: batches in flight (PipelinedPostProcessor lanes); 1 = one "
                                                         "batch at a time on one stream")
    ap.add_argument("--batch", type=int, default=0, help="images per step and GPU (0 = the workload's default)")
    return ap.parse_args()


def make_inputs(rank, n_sets=2):
    """n_sets different batches per rank so that consecutive steps never re-read the same 419 MB."""
    from improved_body_parts_b200 import synth
    return [synth.make_batch(SEED + 100000 * rank + 1000 * s, BATCH, people=(1, 10), height=IMG, width=IMG)
            for s in range(n_sets)]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the timed regions run."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_baseline(n_images, wl=None):
    """The reference's CPU route (oracle/cpu_path.py) on a bounded sample of the same workload, all host cores."""
    from improved_body_parts_b200 import synth
    from oracle import cpu_path
    crowd = wl is not None and getattr(wl, "crowd", False)
    if crowd:
        n_images = min(n_images, 2 * (os.cpu_count() or 1))  # 40-60 people per image: seconds of host work each
    sample = wl.cpu_sample(n_images) if wl is not None else synth.make_batch(SEED, n_images, people=(1, 10), height=IMG, width=IMG)
    passes = 1 if crowd else 6  # about 10-20 s of CPU work over all cores
    secs, procs = cpu_path.time_batch(sample, IMG, repeat=passes)
    return {"value": n_images / secs, "unit": UNIT, "cores": procs,
            "kind": "reference" if cpu_path.cpp_kind() == "reference" else "port",
            "sample": "%d images of the same workload (seed %d..), %d passes after a warm-up, %d processes x 1 thread; C++ stage = %s, "
                      "Python stage = restated scipy.maximum_filter + cv2.resize calls (oracle/cpu_path.py)"
                      % (n_images, SEED, passes, procs, "unmodified reference pafprocess.cpp (oracle/_ref)"
                         if cpu_path.cpp_kind() == "reference" else "oracle port")}


def run_reference(args, rank, world):
    if rank != 0:
        return None
    n = args.cpu_sample or max(8, 4 * (os.cpu_count() or 1))
    from improved_body_parts_b200 import synth
    from oracle import cpu_path
    sample = synth.make_batch(SEED, n, people=(1, 10), height=IMG, width=IMG)
    for _ in range(max(args.warmup, 1) if args.warmup else 0):
        cpu_path.time_batch(sample[:max(2, n // 8)], IMG)
    t_all, procs = 0.0, 1
    for _ in range(args.steps):
        secs, procs = cpu_path.time_batch(sample, IMG)
        t_all += secs
    value = n * args.steps / t_all
    kind = "reference" if cpu_path.cpp_kind() == "reference" else "port"
    sample_txt = ("each step = %d images of the workload on %d host processes (cv2 threads = 1 per process); C++ stage = %s"
                  % (n, procs, "unmodified reference pafprocess.cpp (oracle/_ref)" if kind == "reference" else "oracle port"))
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * t_all / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "images_per_step": n},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": kind, "sample": sample_txt},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


class Refactored:
    """configs[1] / configs[3]: B images of 128x128x50 maps with flip twin through the refactored + C++ route
    (flip fusion -> heatmap_nms peaks -> limb scoring on the fly-upsampled maps -> pafprocess grouping): two kernels."""

    def __init__(self, args, rank, dev, crowd):
        import torch
        from improved_body_parts_b200 import synth
        from improved_body_parts_b200.pipeline import PipelinedPostProcessor
        self.batch = args.batch or BATCH
        self.crowd = crowd
        people, scale = ((40, 60), (0.2, 0.45)) if crowd else ((1, 10), (0.35, 0.9))
        self.people = people
        base = (3 if crowd else 2) * 1000  # 1000 * config_id + image index
        sets = [synth.make_batch(base + 100000 * rank + 1000 * s, self.batch, people=people, height=IMG, width=IMG, scale=scale)
                for s in range(2)]
        self.host_sets = [torch.from_numpy(b).pin_memory() for b in sets]
        self.dev_sets = [h.to(dev) for h in self.host_sets]
        lim = dict(max_peaks_per_part=256, max_candidates_per_limb=1024, max_bodies=512, max_humans=256) if crowd else {}
        # `depth` batches in flight, each lane with its own handle and stream (pipeline.PipelinedPostProcessor)
        self.depth = max(1, getattr(args, "depth", 3))
        self.pipe = PipelinedPostProcessor(device=dev.index, depth=self.depth, max_batch=self.batch, **lim)
        self.pp = self.pipe.lanes[0]
        self.pps = self.pipe.lanes
        self.dev_obj = dev
        self.e2e_mode = False
        self.workload = ("configs[3]: crowd stress, synthetic batch %d, 128x128x50 maps, flip twin, %d-%d people/image"
                         if crowd else "configs[1]: synthetic batch %d, 512x512 input, stride-4 128x128x50 maps, flip twin, "
                                       "%d-%d people/image") % (self.batch, people[0], people[1])
        self.l2 = "2 rotating input sets of %d MB per rank, each larger than the 126 MB L2" % (self.host_sets[0].numel() * 4 // 1000000)
        self.algo_bytes = dict(ALGO_BYTES)
        self.route = "refactored + C++ (predict_refactor, heatmap_nms, process_paf)"

    def step_dev(self, s):
        self.pipe.submit(self.dev_sets[s % 2], IMG, replace=True)  # kernels only: results of earlier batches are dropped

    def join(self):
        """End of a timed region: everything in flight has finished (and, end to end, has been read back)."""
        if self.e2e_mode:
            self.last = (self.pipe.drain() or [self.last])[-1]
        self.pipe.synchronize()

    def _step_host(self, sets, s):
        self.e2e_mode = True
        if self.pipe.full():
            self.last = self.pipe.collect()               # packed skeletons of the oldest batch, device -> host
        self.pipe.submit(sets[s % 2], IMG)                # pinned host -> device inside the timed region

    def step_e2e(self, s):
        self._step_host(self.host_sets, s)

    def step_e2e_f16(self, s):
        """The same call on float16 network output — what the reference's default amp O1 model hands over
        (parse_skeletons.py:75-80); the flip average is then taken in half precision like numpy does."""
        if not hasattr(self, "host16"):
            self.host16 = [h.half().pin_memory() for h in self.host_sets]
        self._step_host(self.host16, s)

    def finish(self):
        res = self.pipe.drain()
        if res:
            self.last = res[-1]
        self.e2e_mode = False

    def h2d(self):
        # ibp_postprocess_batch uploads the 48 limb + keypoint planes of each half; the 2 background planes stay behind
        return self.host_sets[0].numel() * self.host_sets[0].element_size() // 50 * 48

    def d2h(self):
        L = self.pp.handle.limits
        return self.batch * (4 + 18 * 4 + L.max_peaks_per_part * 18 * 16 + 4 + L.max_humans * 88 + 4)

    def humans(self):
        return int(self.last.n_humans.sum())

    def cpu_sample(self, n):
        from improved_body_parts_b200 import synth
        people, scale = ((40, 60), (0.2, 0.45)) if self.crowd else ((1, 10), (0.35, 0.9))
        return synth.make_batch((3 if self.crowd else 2) * 1000, n, people=people, height=IMG, width=IMG, scale=scale)


class MultiScale:
    """configs[2]: flip + 4-scale fusion up-sampled to the 640x480 image, original route (predict -> find_peaks ->
    find_connections -> find_humans, Python semantics in float64): four kernels."""
    H, W, SS = 480, 640, (0.5, 1.0, 1.5, 2.0)

    def __init__(self, args, rank, dev, crowd=False):
        import torch
        from improved_body_parts_b200 import synth
        from improved_body_parts_b200.pipeline import PostProcessor
        self.batch = args.batch or 32
        self.shapes = synth.multiscale_shapes(self.H, self.W, self.SS)
        distinct = min(8, self.batch)
        imgs = [synth.make_multiscale_output(2000 + 100 * rank + i, 1 + (i * 3) % 12, self.H, self.W, self.SS)[0] for i in range(distinct)]
        reps = -(-self.batch // distinct)
        self.host = [torch.from_numpy(np.concatenate([np.stack([im[s] for im in imgs])] * reps)[:self.batch]).pin_memory()
                     for s in range(len(self.SS))]
        self.dev = [[h.to(dev) for h in self.host] for _ in range(2)]
        self.sizes = [s[1] for s in self.shapes]
        self.pp = PostProcessor(device=dev.index, max_batch=self.batch)
        # device-resident run: two batches in flight (own handle and stream each), the three thin kernels at the end of
        # a batch then run under the multi-scale kernel of the next one
        self.depth = 2 if getattr(args, "depth", 3) > 1 else 1
        self.pps = [self.pp] + [PostProcessor(device=dev.index, max_batch=self.batch) for _ in range(self.depth - 1)]
        self.streams = [torch.cuda.Stream(dev) for _ in range(self.depth)]
        self.torch = torch
        self.in_bytes = sum(h.numel() * 4 for h in self.host)
        self.workload = ("configs[2]: flip + 4-scale fusion (0.5/1/1.5/2 x 512/480: maps 64x96..256x352) up-sampled to 640x480, "
                         "batch %d (%d distinct images tiled), 1-12 people/image" % (self.batch, distinct))
        self.l2 = "2 rotating input sets of %d MB per rank plus %d MB of float64 output per step, larger than the 126 MB L2" % (
            self.in_bytes // 1000000, self.batch * 50 * self.H * self.W * 8 // 1000000)
        px = self.H * self.W
        self.algo_bytes = {"fuse": self.in_bytes // self.batch + 50 * px * 8 + 18 * px * 4, "peaks": 18 * px * 4}
        self.route = "original (predict, find_peaks, find_connections, find_humans)"
        self.e2e_note = ("each step uploads one full input set from pinned host memory; the upload of step k+1 runs on a copy "
                         "stream under the kernels of step k and step k does not end before it has")
        self.dev_obj = dev

    def step_dev(self, s):
        if self.depth == 1:
            self.pp.enqueue_original_route(self.dev[s % 2], self.sizes, self.H, self.W)
            return
        k = s % self.depth
        self.streams[k].wait_stream(self.torch.cuda.current_stream(self.dev_obj))
        with self.torch.cuda.stream(self.streams[k]):
            self.pps[k].enqueue_original_route(self.dev[s % 2], self.sizes, self.H, self.W)

    def join(self):
        for st in self.streams:
            self.torch.cuda.current_stream(self.dev_obj).wait_stream(st)

    def _upload(self, k):
        """Pinned host -> device copy of one step's inputs (all scales) on the copy stream."""
        import torch
        with torch.cuda.stream(self.copy_stream):
            for d, h in zip(self.dev[k], self.host):
                d.copy_(h, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        return k, ev

    def step_e2e(self, s):
        """One step = one upload of a full input set + the four kernels + the read-back of the result.  The upload of
        the NEXT step's inputs runs on a copy stream under this step's kernels (two device input sets); the step does
        not end before that upload has, so K timed steps contain K whole uploads."""
        import torch
        cur = torch.cuda.current_stream(self.dev_obj)
        if not hasattr(self, "copy_stream"):
            self.copy_stream = torch.cuda.Stream(self.dev_obj)
            self.copy_stream.wait_stream(cur)
            self.inflight = self._upload(0)
        k, ev = self.inflight
        cur.wait_event(ev)
        self.pp.enqueue_original_route(self.dev[k], self.sizes, self.H, self.W)
        self.inflight = self._upload(1 - k)
        cur.wait_event(self.inflight[1])
        self.last = self.pp.fetch_original_route(self.batch)

    def finish(self):
        self.last = self.pp.fetch_original_route(self.batch)

    def h2d(self):
        return self.in_bytes

    def d2h(self):
        return int(sum(t.nbytes + c.nbytes + p.nbytes for t, c, p in self.last))

    def humans(self):
        return int(sum(p.shape[0] for _, _, p in self.last))


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    wl = {"configs1": lambda: Refactored(args, rank, dev, False), "configs3": lambda: Refactored(args, rank, dev, True),
          "configs2": lambda: MultiScale(args, rank, dev)}[args.workload]()
    pp, batch = wl.pp, wl.batch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps):
        """K steps bracketed by barrier + synchronize; device time by CUDA events; max over ranks."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(steps):
            fn(s)
        if hasattr(wl, "join"):
            wl.join()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        barrier()
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- device-resident throughput (`value`) ----
    for s in range(args.warmup):
        wl.step_dev(s)
    wl.finish()
    handles = [p.handle for p in getattr(wl, "pps", [pp])]
    launches0 = sum(h.launches for h in handles)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms = timed(wl.step_dev, args.steps)
    launches = sum(h.launches for h in handles) - launches0
    wl.finish()

    # ---- end to end through the public API with HOST buffers (`e2e`) ----
    for s in range(min(args.warmup, 3)):
        wl.step_e2e(s)
    wl.finish()  # nothing in flight when the timed region starts; K uploads, K read-backs inside it
    ms_e2e = timed(wl.step_e2e, args.steps)
    wl.finish()
    ms_e2e16 = None
    if hasattr(wl, "step_e2e_f16"):  # informational: float16 hand-over (the reference's amp default), same public call
        for s in range(3):
            wl.step_e2e_f16(s)
        wl.finish()
        ms_e2e16 = timed(wl.step_e2e_f16, args.steps)
        wl.finish()

    # ---- per-kernel device times for the roofline (events on the launching stream inside the library) ----
    # per-kernel times: one batch at a time on the first lane
    step_one = (lambda s: pp.enqueue(wl.dev_sets[s % 2], IMG)) if hasattr(wl, "pipe") else wl.step_dev
    depth_used = getattr(wl, "depth", 1)
    if not hasattr(wl, "pipe") and hasattr(wl, "depth"):
        wl.depth = 1
    pp.handle.set_profiling(True)
    acc = {}
    for s in range(args.steps):
        step_one(s)
        for k, v in pp.handle.stage_ms().items():
            acc.setdefault(k, []).append(v)
    pp.handle.set_profiling(False)
    wl.finish()
    stage_ms = {k: float(np.mean(v)) for k, v in acc.items()}
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return None

    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_file):
        peak, peak_src = float(json.load(open(peaks_file))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"
    algo = wl.algo_bytes
    traffic = None  # dram bytes read + written per launch of the dominant kernel, from the committed ncu --set full capture
    tfile = os.path.join(ROOT, "profiles", "r1_traffic.json")
    if os.path.exists(tfile):
        t = json.load(open(tfile))
        # keyed by kernel (peaks / index / limbs: configs1 at batch 64; fuse: configs2 at batch 32); per launch, so only
        # meaningful at the batch size it was captured on
        if (args.workload, batch) in (("configs1", 64), ("configs2", 32)):
            traffic = t["kernels"]
    dominant = max((k for k in stage_ms if k in algo), key=lambda k: stage_ms[k])
    achieved = algo[dominant] * batch / (stage_ms[dominant] * 1e-3) / 1e9
    traffic_bytes = (traffic[dominant]["dram_bytes_read"] + traffic[dominant]["dram_bytes_write"]) if traffic and dominant in traffic else None
    out = {"metric": METRIC, "value": world * batch * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f64" if args.workload == "configs2" else "f32", "data": "synthetic",
           "config": {"workload": wl.workload, "route": wl.route, "batch_per_gpu": batch, "global_batch": world * batch,
                      "parallelism": "images sharded, %d rank(s), no collective" % world, "l2_policy": wl.l2,
                      **({"batches_in_flight": depth_used} if hasattr(wl, "depth") else {})},
           "e2e": {"value": world * batch * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
                   "h2d_bytes_per_step": world * wl.h2d(), "d2h_bytes_per_step": world * wl.d2h(),
                   **({"note": wl.e2e_note} if hasattr(wl, "e2e_note") else {})},
           "e2e_f16": ({"value": world * batch * args.steps / (ms_e2e16 * 1e-3), "unit": UNIT,
                        "h2d_bytes_per_step": world * wl.h2d() // 2,
                        "note": "same call, float16 network output (reference default amp O1); not the headline dtype"}
                       if ms_e2e16 else None),
           "gpu_launches": int(launches),
           "clocks": clocks,
           "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "frac_of_nominal_8000": achieved / 8000.0, "traffic": traffic_bytes,
                        # the limb kernel stages only the rows its samples can touch: what it really moves per second
                        "achieved_on_traffic": (traffic_bytes / (stage_ms[dominant] * 1e-3) / 1e9) if traffic_bytes else None, "peak_source": peak_src,
                        "algorithmic_bytes_per_launch": algo[dominant] * batch},
           "stage_ms": stage_ms,
           "stage_gbs": {k: algo[k] * batch / (stage_ms[k] * 1e-3) / 1e9 for k in stage_ms if k in algo},
           "whole_step_gbs": sum(algo.values()) * batch / (ms / args.steps * 1e-3) / 1e9,
           "humans_last_batch": wl.humans()}
    if not args.no_cpu_baseline and hasattr(wl, "cpu_sample") and world == 1:  # rank 0 at N = 1 only
        out["cpu_baseline"] = cpu_baseline(args.cpu_sample or max(8, 8 * (os.cpu_count() or 1)), wl)
    return out


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        out = run_reference(args, rank, world)
    else:
        import torch
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        out = run_ours(args, rank, world, local_rank)
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
            dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
