#!/usr/bin/env python
"""bench.py — post-processing images/s of the keypoint-assignment path (BASELINE.json metric) on N B200s.

A step = one pass of the hot path over one batch of synthetic network outputs.  Headline workload at every N:
BASELINE.json configs[1] ("synthetic batch 64, 512x512 input, stride-4 128x128 maps, 50 channels"), per rank (weak
scaling; images shard, no data-path collective — SURVEY.md §8e).  At N = 1 the same JSON line also carries, under
`workloads`, configs[2] (640x480, flip + 4 scales, original route), configs[3] (crowds) and configs[4] (model-fed sweep,
when the reference's network definition travelled in baseline/_ref), each with its own value / e2e / roofline /
cpu_baseline, plus the `legacy` block (the seven-call drop-in against the unmodified reference library) and the
remaining CPU baselines of BASELINE.md §3 under `cpu_baselines`.

    python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (under torchrun when N > 1)
    python bench.py --impl reference [--gpus N] ...                the reference's CPU route on the host cores
    python bench.py --workload configs2|configs3|configs4          one workload as the headline line

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
METRIC = "post-proc images/sec"
UNIT = "images/s"
MIN_REGION_MS = 60.0  # every reported time is the median of enough K-step regions to cover at least this much
CPU_BUDGET_S = 12.0   # host work per cpu_baseline leg


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="images in the cpu_baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="configs1", choices=["configs1", "configs2", "configs3", "configs4"],
                    help="the headline line's workload.  configs1 (default): batch 64 of 128x128x50 maps, refactored + C++ "
                         "route; configs2: flip + 4-scale fusion to 640x480, original route; configs3: crowd 40-60 people; "
                         "configs4: sweep fed in-stream by the random-init IMHN forward")
    ap.add_argument("--no-extras", action="store_true", help="N = 1 default run: skip the `workloads`, `legacy` and "
                                                             "`cpu_baselines` blocks (headline only)")
    ap.add_argument("--depth", type=int, default=3, help="configs1/3: batches in flight (PipelinedPostProcessor lanes); 1 = one "
                                                         "batch at a time on one stream")
    ap.add_argument("--batch", type=int, default=0, help="images per step and GPU (0 = the workload's default)")
    ap.add_argument("--sweep-images", type=int, default=0, help="configs4: images of the whole sweep (0 = 10000 / scaled down "
                                                                "to fit the default run)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------ workload descriptions
def workload_text(name, batch):
    return {
        "configs1": "configs[1]: synthetic batch %d, 512x512 input, stride-4 128x128x50 maps, flip twin, 1-10 people/image" % batch,
        "configs2": "configs[2]: flip + 4-scale fusion (0.5/1/1.5/2 x 512/480: maps 64x96..256x352) up-sampled to 640x480, "
                    "batch %d distinct images, 1-12 people/image" % batch,
        "configs2_second_mode": "configs[2] inputs (flip + 4-scale fusion to 640x480, batch %d) through the SECOND operator mode — the "
                                "north_star's operators, not the reference's: bilinear in both resizes, 5-tap Gaussian (sigma 1) in "
                                "front of the 3x3 max-suppression" % batch,
        "configs3": "configs[3]: crowd stress, synthetic batch %d, 128x128x50 maps, flip twin, 40-60 people/image" % batch,
        "configs4": "configs[4]: image sweep fed in-stream by the random-init IMHN forward (fp16 autocast), %d images x 2 flips "
                    "per forward, 512x512 input" % batch,
    }[name]


ROUTES = {"configs1": "refactored + C++ (predict_refactor, heatmap_nms, process_paf)",
          "configs3": "refactored + C++ (predict_refactor, heatmap_nms, process_paf)",
          "configs2": "original (predict, find_peaks, find_connections, find_humans)",
          "configs2_second_mode": "original route with ibp_fuse_multiscale_linear + ibp_smooth_gaussian (second operator mode, own oracle, 1e-4)",
          "configs4": "IMHN forward -> refactored + C++ (predict_refactor, heatmap_nms, process_paf), fetch_coco"}
DEFAULT_BATCH = {"configs1": BATCH, "configs2": 64, "configs2_second_mode": 64, "configs3": BATCH, "configs4": 16}


def config_of(name, world, batch, depth):
    """The `config` object — the SAME for our arm and the reference arm (it names the workload, not the implementation)."""
    return {"workload": workload_text(name, batch), "route": ROUTES[name], "batch_per_gpu": batch, "global_batch": world * batch,
            "parallelism": "images sharded, %d rank(s), no collective" % world,
            "l2_policy": "2 rotating input sets per rank, each larger than the 126 MB L2",
            "batches_in_flight": depth}


def seeds_of(name, rank, s):
    """1000 * config_id + image index (+ per-rank / per-set offsets)."""
    return {"configs1": 2000, "configs3": 3000, "configs2": 2000, "configs4": 5000}[name] + 100000 * rank + 1000 * s


def gen_refactored(name, rank, batch, n_sets=2):
    from improved_body_parts_b200 import synth
    people, scale = ((40, 60), (0.2, 0.45)) if name == "configs3" else ((1, 10), (0.35, 0.9))
    return [synth.make_batch(seeds_of(name, rank, s), batch, people=people, height=IMG, width=IMG, scale=scale, workers=0)
            for s in range(n_sets)]


MS_H, MS_W, MS_SS = 480, 640, (0.5, 1.0, 1.5, 2.0)


def ms_seeds(rank, n):
    return [(2000 + 100 * rank + i, 1 + (i * 3) % 12) for i in range(n)]


def gen_multiscale(rank, batch):
    from improved_body_parts_b200 import synth
    return synth.make_multiscale_batch(ms_seeds(rank, batch), MS_H, MS_W, MS_SS, workers=0)


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
        # "under load": samples above the idle clock; the bench alternates short GPU bursts with host work
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU baselines
def cpu_kind():
    from oracle import cpu_path
    return "reference" if cpu_path.cpp_kind() == "reference" else "port"


def cpp_stage_text():
    return ("unmodified reference pafprocess.cpp (oracle/_ref)" if cpu_kind() == "reference" else "oracle port of pafprocess.cpp")


def cpu_baseline_refactored(sample, passes=None):
    """The reference's refactored + C++ route (oracle/cpu_path.py) on a bounded sample of the workload, all host cores."""
    from oracle import cpu_path
    n = sample.shape[0]
    secs1, procs = cpu_path.time_batch(sample, IMG, repeat=1)
    passes = passes or max(1, min(8, int(CPU_BUDGET_S / max(secs1, 1e-3)) - 1))
    secs, procs = cpu_path.time_batch(sample, IMG, repeat=passes)
    return {"value": n / secs, "unit": UNIT, "cores": procs, "kind": cpu_kind(),
            "kind_detail": "reference C++ + restated Python",
            "sample": "%d images of the same workload, %d passes after a warm-up pass, %d processes x 1 thread; C++ stage = %s, "
                      "Python stage = restatement issuing the reference's own scipy.maximum_filter + cv2.resize calls "
                      "(oracle/cpu_path.py)" % (n, passes, procs, cpp_stage_text())}


def cpu_baseline_original(cases_by_scale, n_images, second_mode=False):
    """The reference's ORIGINAL route on the host (configs[2]): predict's fusion with the reference's cv2.resize calls,
    find_peaks, find_connections, find_humans in numpy/Python (oracle/cpu_path.original_route_image)."""
    from improved_body_parts_b200 import synth
    from oracle import cpu_path
    shapes = synth.multiscale_shapes(MS_H, MS_W, MS_SS)
    cases = [[cases_by_scale[s][i] for s in range(len(MS_SS))] for i in range(n_images)]
    secs, procs = cpu_path.time_original_route(cases, MS_H, MS_W, [s[0] for s in shapes], second_mode=second_mode)
    if second_mode:
        return {"value": n_images / secs, "unit": UNIT, "cores": procs, "kind": "port",
                "kind_detail": "the same route with cv2.INTER_LINEAR in predict's resizes and a numpy 5-tap Gaussian (sigma 1) in front of find_peaks",
                "sample": "%d images of the same workload, one pass after a warm-up image per process, %d processes x 1 thread" % (n_images, procs)}
    return {"value": n_images / secs, "unit": UNIT, "cores": procs, "kind": "port",
            "kind_detail": "restated Python issuing the reference's cv2.resize calls (the route has no C++ stage)",
            "sample": "%d images of the same workload, one pass after a warm-up image per process, %d processes x 1 thread: "
                      "predict fusion (cv2.resize INTER_CUBIC x4, crop, resize to 640x480, float64 average) + find_peaks + "
                      "find_connections + find_humans (oracle/cpu_path.original_route_image)" % (n_images, procs)}


_legs_sample = {}


def cpu_baseline_extra_legs(sample):
    """BASELINE.md §3 legs that are not the all-cores number: the single-process C++ route the way evaluate.py runs it
    (cv2 at 1 thread and at cv2's default), and the pure-Python find_connections + find_humans on a subsample.
    Runs in a forked child: cv2's worker threads must never exist in THIS process, which forks pools elsewhere."""
    import multiprocessing as mp
    _legs_sample["s"] = sample
    with mp.get_context("fork").Pool(1) as pool:
        return pool.apply(_extra_legs_child)


def _extra_legs_child():
    import cv2
    sample = _legs_sample["s"]
    from oracle import cpu_path
    from oracle import np_oracle as O
    out = {}
    n = min(sample.shape[0], 12)
    default_threads = cv2.getNumThreads()
    for label, thr in (("cv2_1_thread", 1), ("cv2_default_threads", default_threads)):
        cv2.setNumThreads(thr)
        cpu_path.postprocess_image(sample[0], IMG)
        t0 = time.perf_counter()
        for i in range(n):
            cpu_path.postprocess_image(sample[i], IMG)
        dt = time.perf_counter() - t0
        out["single_process_cpp_route_" + label] = {"value": n / dt, "unit": UNIT, "cores": 1, "cv2_threads": thr, "kind": cpu_kind(),
                                                    "sample": "%d images of configs[1], one process (evaluate.py --run_refactor --run_cpp "
                                                              "after the model call, evaluate.py:75-129)" % n}
    cv2.setNumThreads(default_threads)
    # pure-Python route (evaluate.py:88-90 with --run_refactor and without --run_cpp): find_connections + find_humans on the
    # refactored route's peaks and x4 limb maps
    m = min(sample.shape[0], 6)
    prepared = []
    for i in range(m):
        heat, paf = O.fuse_flip(sample[i])
        peaks = cpu_path.heatmap_nms_cv2(heat, 4)
        paf_up = cv2.resize(paf, None, fx=4, fy=4, interpolation=cv2.INTER_CUBIC)
        prepared.append(([[tuple(r) for r in p] for p in peaks], paf_up))
    t0 = time.perf_counter()
    people = 0
    for peaks, paf_up in prepared:
        cl, sp = O.find_connections(peaks, paf_up, IMG)
        people += O.find_humans(cl, sp, peaks)[0].shape[0]
    dt = time.perf_counter() - t0
    out["python_find_connections_find_humans"] = {"value": m / dt, "unit": UNIT, "cores": 1, "kind": "port",
                                                  "sample": "%d images of configs[1] (%d people found), find_connections + find_humans only "
                                                            "(parse_skeletons.py:324-599 restated in oracle/np_oracle.py)" % (m, people)}
    return out


def legacy_block():
    """The seven-call drop-in (pafprocess.process_paf + getters, HOST arrays in the reference layout) against the
    unmodified reference library on the same arrays: 1 person, 10 people, 50 people."""
    import cv2
    import torch
    from improved_body_parts_b200 import pafprocess, synth
    cv2.setNumThreads(1)  # (input preparation only) no cv2 worker threads in a process that forks pools
    from oracle import np_oracle as O
    from oracle.cpu_lib import Oracle, Reference, reference_available
    ref = Reference() if reference_available() else Oracle()
    cases = []
    for people, seed in ((1, 11), (10, 13), (50, 15)):
        out, _ = synth.make_network_output(seed, people)
        heat, paf = O.fuse_flip(out)
        jl = O.joint_list(O.heatmap_nms(heat, 4))[None]
        up = cv2.resize(paf, None, fx=4, fy=4, interpolation=cv2.INTER_CUBIC)

        def getters(mod):
            tot = 0
            for h in range(mod.get_num_humans()):
                for part in range(18):
                    pid = mod.get_part_peak_id(h, part)
                    if pid >= 0:
                        tot += mod.get_part_x(pid) + mod.get_part_y(pid)
                        mod.get_part_score(pid)
                mod.get_score(h)
            return tot

        def clock(fn, reps):
            fn()
            ts = []
            for _ in range(reps):
                t0 = time.perf_counter()
                fn()
                ts.append(time.perf_counter() - t0)
            return 1e3 * float(np.median(ts))

        reps = 15 if people < 50 else 7
        row = {"people": people, "peaks": int(jl.shape[1]), "map_mb": up.nbytes / 1e6}
        row["reference_ms"] = clock(lambda: (ref.process_paf(jl, up, IMG), getters(ref)), reps)
        row["reference_process_paf_only_ms"] = clock(lambda: ref.process_paf(jl, up, IMG), reps)
        row["ours_pageable_ms"] = clock(lambda: (pafprocess.process_paf(jl, up, IMG), getters(pafprocess)), reps)
        row["ours_process_paf_only_ms"] = clock(lambda: pafprocess.process_paf(jl, up, IMG), reps)
        pinned = torch.from_numpy(up).pin_memory().numpy()
        row["ours_pinned_ms"] = clock(lambda: (pafprocess.process_paf(jl, pinned, IMG), getters(pafprocess)), reps)
        dev = torch.from_numpy(up).cuda()
        row["ours_cuda_tensor_ms"] = clock(lambda: (pafprocess.process_paf(jl, dev, IMG), getters(pafprocess)), reps)
        row["humans"] = int(pafprocess.get_num_humans())
        row["speedup_pageable"] = row["reference_ms"] / row["ours_pageable_ms"]
        cases.append(row)
    return {"what": "process_paf + the getter loop of evaluate.py:110-129 on the same host arrays; median ms per image",
            "reference": cpp_stage_text() if isinstance(ref, Reference) else "oracle port", "cases": cases}


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args, rank, world):
    """The reference's own CPU implementation of the path on the box's host cores; rank 0 alone."""
    if rank != 0:
        return None
    from oracle import cpu_path
    name = args.workload
    batch = args.batch or DEFAULT_BATCH[name]
    ncores = os.cpu_count() or 1
    if name == "configs2":
        n = args.cpu_sample or max(4, ncores)
        cases = gen_multiscale(0, n)
        t_all = 0.0
        for _ in range(args.steps):
            cb = cpu_baseline_original(cases, n)
            t_all += n / cb["value"]
        procs, sample_txt, kind = cb["cores"], cb["sample"], "port"
    else:
        n = args.cpu_sample or (max(8, 4 * ncores) if name != "configs3" else max(4, ncores))
        from improved_body_parts_b200 import synth
        people, scale = ((40, 60), (0.2, 0.45)) if name == "configs3" else ((1, 10), (0.35, 0.9))
        sample = synth.make_batch(seeds_of(name if name != "configs4" else "configs1", 0, 0), n, people=people, height=IMG, width=IMG,
                                  scale=scale, workers=0)
        for _ in range(1 if args.warmup else 0):
            cpu_path.time_batch(sample[:max(2, n // 8)], IMG)
        t_all, procs = 0.0, 1
        for _ in range(args.steps):
            secs, procs = cpu_path.time_batch(sample, IMG)
            t_all += secs
        kind = cpu_kind()
        sample_txt = ("each step = %d images of the workload on %d host processes (cv2 threads = 1 per process); C++ stage = %s, "
                      "Python stage = restated scipy/cv2 calls" % (n, procs, cpp_stage_text()))
    value = n * args.steps / t_all
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * t_all / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64" if name == "configs2" else "f32",
            "data": "synthetic", "config": config_of(name, world, batch, 3 if name != "configs2" else 2),
            "images_per_step": n,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": kind, "sample": sample_txt},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


# ------------------------------------------------------------------------------------------------ our workloads
class Refactored:
    """configs[1] / configs[3]: B images of 128x128x50 maps with flip twin through the refactored + C++ route
    (flip fusion -> heatmap_nms peaks -> limb scoring on the fly-upsampled maps -> pafprocess grouping)."""

    def __init__(self, name, args, dev, sets, rank=0):
        import torch
        from improved_body_parts_b200.pipeline import PipelinedPostProcessor
        self.name, self.crowd = name, name == "configs3"
        self.batch = sets[0].shape[0]
        self.sets_np = sets
        self.host_sets = [torch.from_numpy(b).pin_memory() for b in sets]
        self.dev_sets = [h.to(dev) for h in self.host_sets]
        lim = dict(max_peaks_per_part=256, max_candidates_per_limb=1024, max_bodies=512, max_humans=256) if self.crowd else {}
        self.depth = max(1, args.depth)
        self.pipe = PipelinedPostProcessor(device=dev.index, depth=self.depth, max_batch=self.batch, **lim)
        self.pp = self.pipe.lanes[0]
        self.pps = self.pipe.lanes
        self.dev_obj = dev
        self.e2e_mode = False
        self.dtype = "f32"
        fb = 2 * FEAT * FEAT * 4
        self.e2e_note = ("the caller hands over the whole (B,2,50,h,w) pinned host tensor; h2d_bytes_per_step is what crosses the bus: the 36 "
                         "keypoint planes per image by cudaMemcpy2DAsync plus the limb boxes the limb kernel pulls straight from the pinned "
                         "buffer (rows x columns between the extreme peaks of a limb's joint types, counted from the fetched peaks); "
                         "h2d_bytes_offered_per_step = the 96 limb + keypoint planes per image a full copy would move")
        self.algo_bytes = {"peaks": 18 * fb, "limbs": 30 * fb}  # planes of the upright and the flipped output, read once

    def step_dev(self, s):
        self.pipe.submit(self.dev_sets[s % 2], IMG, replace=True)  # kernels only: results of earlier batches are dropped

    def step_one(self, s):
        self.pp.enqueue(self.dev_sets[s % 2], IMG)

    def step_dev_f16(self, s):
        """Device-resident float16 network output (the reference's amp O1 default): same kernels, half the plane bytes."""
        if not hasattr(self, "dev16"):
            self.dev16 = [d.half() for d in self.dev_sets]
        self.pipe.submit(self.dev16[s % 2], IMG, replace=True)

    def join(self):
        """End of a timed region: everything in flight has finished (and, end to end, has been read back)."""
        if self.e2e_mode:
            self.last = (self.pipe.drain(coco=True) or [self.last])[-1]
        self.pipe.synchronize()

    def _step_host(self, sets, s):
        self.e2e_mode = True
        if self.pipe.full():
            self.last = self.pipe.collect(coco=True)      # COCO rows of the oldest batch, built and compacted on the device
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
        res = self.pipe.drain(coco=self.e2e_mode)
        if res:
            self.last = res[-1]
        self.e2e_mode = False

    def h2d(self, esz=4):
        """Bytes that cross the bus per step.  ibp_postprocess_batch copies the 18 keypoint planes of each half; the limb
        kernel pulls the limb rows it needs straight out of the pinned host buffer (zero copy) — for limb l of an image the
        box of feature rows x columns between the extreme peaks of its two joint types plus the bicubic support (the
        formula of index_bands_kernel, csrc/k_peaks.cuh), of the upright and of the twin plane.  Counted here from the fetched peaks,
        averaged over the two input sets; the 2 background planes never move.  With IBP_NO_ZEROCOPY the 48 planes are copied."""
        full = self.host_sets[0].numel() * esz // 50
        if os.environ.get("IBP_NO_ZEROCOPY") or self.crowd:  # the crowd handle runs the one-CTA-per-item limb kernel: staged copy
            return full * 48
        if not hasattr(self, "_limb_elems"):
            from improved_body_parts_b200.config import LIMBS
            rows = []
            for d in self.dev_sets:
                self.pp.enqueue(d, IMG)
                res = self.pp.fetch(self.batch)
                n = 0
                for b in range(self.batch):
                    pk = res.peaks[b]
                    lo, hi, xl, xh, cnt = {}, {}, {}, {}, {}
                    for part in range(18):
                        sel = pk["part"] == part
                        ys, xs = pk["y"][sel], pk["x"][sel]
                        cnt[part] = len(ys)
                        if len(ys):
                            lo[part], hi[part], xl[part], xh[part] = int(ys.min()), int(ys.max()), int(xs.min()), int(xs.max())
                    for pa, pb in LIMBS:
                        pa, pb = int(pa), int(pb)
                        if cnt[pa] and cnt[pb]:
                            r0 = max(0, ((min(lo[pa], lo[pb]) - 2) >> 2) - 2)
                            r1 = min(FEAT, ((max(hi[pa], hi[pb]) - 2) >> 2) + 4)
                            c0 = max(0, ((min(xl[pa], xl[pb]) - 2) >> 2) - 2) & ~7
                            c1 = min(FEAT, (((max(xh[pa], xh[pb]) - 2) >> 2) + 4 + 7) & ~7)
                            n += (r1 - r0) * (c1 - c0)
                rows.append(n)
            self._limb_elems = float(np.mean(rows))
        return int(full * 18 + 2 * self._limb_elems * esz)

    def d2h(self):
        # ibp_fetch_coco: offsets + status + one 52-float row per human of the batch
        rows = sum(r.shape[0] for r in self.last) if isinstance(self.last, list) else 0
        return 4 * (self.batch + 1) + 4 * self.batch + rows * 52 * 4

    def humans(self):
        if isinstance(self.last, list):
            return int(sum(r.shape[0] for r in self.last))
        return int(self.last.n_humans.sum())

    def cpu_baseline(self, args):
        n = args.cpu_sample or (max(8, 8 * (os.cpu_count() or 1)) if not self.crowd else 2 * (os.cpu_count() or 1))
        n = min(n, 2 * self.batch)
        sample = np.concatenate(self.sets_np)[:n]
        return cpu_baseline_refactored(sample, passes=1 if self.crowd else None)

    def release(self):
        del self.pipe, self.pp, self.pps, self.dev_sets, self.host_sets
        for k in ("dev16", "host16"):
            if hasattr(self, k):
                delattr(self, k)


class MultiScale:
    """configs[2]: flip + 4-scale fusion up-sampled to the 640x480 image, original route (predict -> find_peaks ->
    find_connections -> find_humans, Python semantics in float64): four kernels."""
    H, W, SS = MS_H, MS_W, MS_SS

    def __init__(self, name, args, dev, by_scale, rank=0, route_kw=None):
        import torch
        from improved_body_parts_b200 import synth
        from improved_body_parts_b200.pipeline import PostProcessor
        self.name = name
        # {} = the reference's operators; {"interp": "linear", "smooth": (5, 1.0)} = the second operator mode
        self.route_kw = route_kw or {}
        self.by_scale = by_scale
        self.batch = by_scale[0].shape[0]
        self.shapes = synth.multiscale_shapes(self.H, self.W, self.SS)
        self.host = [torch.from_numpy(a).pin_memory() for a in by_scale]
        # second rotating set: the same images in another order (every set is tens of times the L2)
        self.dev = [[h.to(dev) for h in self.host], [h.to(dev).roll(self.batch // 2 + 1, 0).contiguous() for h in self.host]]
        self.sizes = [s[1] for s in self.shapes]
        self.pp = PostProcessor(device=dev.index, max_batch=self.batch)
        # device-resident run: two batches in flight (own handle and stream each), the three thin kernels at the end of
        # a batch then run under the multi-scale kernel of the next one
        self.depth = 2 if args.depth > 1 else 1
        self.pps = [self.pp] + [PostProcessor(device=dev.index, max_batch=self.batch) for _ in range(self.depth - 1)]
        self.streams = [torch.cuda.Stream(dev) for _ in range(self.depth)]
        self.torch = torch
        self.in_bytes = sum(h.numel() * 4 for h in self.host)
        px = self.H * self.W
        self.algo_bytes = {"fuse": self.in_bytes // self.batch + 50 * px * 8 + 18 * px * 4, "peaks": 18 * px * 4, "smooth": 2 * 18 * px * 4}
        self.e2e_note = ("each step uploads one full input set from pinned host memory; the upload of step k+1 runs on a copy "
                         "stream under the kernels of step k and step k does not end before it has")
        self.dev_obj = dev
        self.dtype = "f64"
        self.one_at_a_time = False

    def step_dev(self, s):
        if self.depth == 1 or self.one_at_a_time:
            self.pp.enqueue_original_route(self.dev[s % 2], self.sizes, self.H, self.W, **self.route_kw)
            return
        k = s % self.depth
        self.streams[k].wait_stream(self.torch.cuda.current_stream(self.dev_obj))
        with self.torch.cuda.stream(self.streams[k]):
            self.pps[k].enqueue_original_route(self.dev[s % 2], self.sizes, self.H, self.W, **self.route_kw)

    def step_one(self, s):
        self.pp.enqueue_original_route(self.dev[s % 2], self.sizes, self.H, self.W, **self.route_kw)

    def join(self):
        for st in self.streams:
            self.torch.cuda.current_stream(self.dev_obj).wait_stream(st)

    def _upload(self, k):
        """Pinned host -> device copy of one step's inputs (all scales) on the copy stream."""
        torch = self.torch
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
        torch = self.torch
        cur = torch.cuda.current_stream(self.dev_obj)
        if not hasattr(self, "copy_stream"):
            self.copy_stream = torch.cuda.Stream(self.dev_obj)
            self.copy_stream.wait_stream(cur)
            self.inflight = self._upload(0)
        k, ev = self.inflight
        cur.wait_event(ev)
        self.pp.enqueue_original_route(self.dev[k], self.sizes, self.H, self.W, **self.route_kw)
        self.inflight = self._upload(1 - k)
        cur.wait_event(self.inflight[1])
        self.last = self.pp.fetch_original_route(self.batch)

    def finish(self):
        self.join()
        self.last = self.pp.fetch_original_route(self.batch)

    def h2d(self):
        return self.in_bytes

    def d2h(self):
        return int(sum(t.nbytes + c.nbytes + p.nbytes for t, c, p in self.last))

    def humans(self):
        return int(sum(p.shape[0] for _, _, p in self.last))

    def cpu_baseline(self, args):
        n = min(args.cpu_sample or max(4, os.cpu_count() or 1), self.batch)
        return cpu_baseline_original(self.by_scale, n, second_mode=bool(self.route_kw))

    def release(self):
        del self.pps, self.pp, self.dev, self.host


def measure(wl, args, dev, world, rank, traffic_key=None):
    """value / e2e / roofline / stage times of one workload.  Every number is the median over enough K-step regions to
    cover MIN_REGION_MS; each region is bracketed by barrier + synchronize, timed with CUDA events, max over ranks."""
    import torch
    import torch.distributed as dist

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(steps):
            fn(s)
        wl.join()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        barrier()
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def timed_median(fn, steps, after=None, max_reps=40):
        first = timed(fn, steps)
        if after:
            after()
        reps = int(min(max_reps, max(1, np.ceil(MIN_REGION_MS / max(first, 1e-3)))))
        if world > 1:  # every rank must run the same number of regions
            t = torch.tensor([reps], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            reps = int(t.item())
        all_ms = [first]
        for _ in range(reps - 1):
            all_ms.append(timed(fn, steps))
            if after:
                after()
        return float(np.median(all_ms)), all_ms

    batch = wl.batch
    handles = [p.handle for p in wl.pps]
    # ---- device-resident throughput (`value`) ----
    for s in range(args.warmup):
        wl.step_dev(s)
    wl.finish()
    launches0 = sum(h.launches for h in handles)
    ms, regions = timed_median(wl.step_dev, args.steps, after=wl.finish)
    launches = (sum(h.launches for h in handles) - launches0) // len(regions)
    ms_dev16 = None
    if hasattr(wl, "step_dev_f16"):
        for s in range(3):
            wl.step_dev_f16(s)
        wl.finish()
        ms_dev16, _ = timed_median(wl.step_dev_f16, args.steps, after=wl.finish)

    # ---- end to end through the public API with HOST buffers (`e2e`) ----
    for s in range(min(args.warmup, 3)):
        wl.step_e2e(s)
    wl.finish()  # nothing in flight when the timed region starts; K uploads, K read-backs inside it
    ms_e2e, regions_e2e = timed_median(wl.step_e2e, args.steps, after=wl.finish, max_reps=3)
    ms_e2e16 = None
    if hasattr(wl, "step_e2e_f16"):  # float16 hand-over (the reference's amp O1 default), same public call
        for s in range(3):
            wl.step_e2e_f16(s)
        wl.finish()
        ms_e2e16, _ = timed_median(wl.step_e2e_f16, args.steps, after=wl.finish, max_reps=3)

    # ---- per-kernel device times for the roofline: one batch at a time on the first lane, events on the launching
    #      stream inside the library ----
    pp = wl.pp
    pp.handle.set_profiling(True)
    acc = {}
    for s in range(max(args.steps, 10)):
        wl.step_one(s)
        for k, v in pp.handle.stage_ms().items():
            acc.setdefault(k, []).append(v)
    pp.handle.set_profiling(False)
    wl.finish()
    stage_ms = {k: float(np.median(v)) for k, v in acc.items()}

    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_file):
        peak, peak_src = float(json.load(open(peaks_file))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth, burst)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback 6.65 TB/s"
    algo = wl.algo_bytes
    # dram bytes read + written per launch of each kernel, from the committed `ncu --set full` captures: per launch, so only
    # meaningful at the batch size it was captured on (profiles/r2_traffic.json: {"<workload>@<batch>": {kernel: {...}}})
    traffic = None
    key = "%s@%d" % (wl.name.replace("_second_mode", ""), batch)
    alias = {"fuse": "fuse_linear"} if getattr(wl, "route_kw", None) else {}  # the stage timer is shared, the kernel is not
    tfile = os.path.join(ROOT, "profiles", "r2_traffic.json")
    if os.path.exists(tfile):
        traffic = json.load(open(tfile)).get(key)
    if traffic is None and key in ("configs1@64", "configs2@32"):
        tfile = os.path.join(ROOT, "profiles", "r1_traffic.json")
        if os.path.exists(tfile):
            traffic = json.load(open(tfile))["kernels"]
    dominant = max((k for k in stage_ms if k in algo), key=lambda k: stage_ms[k])
    achieved = algo[dominant] * batch / (stage_ms[dominant] * 1e-3) / 1e9
    tdom = alias.get(dominant, dominant)
    traffic_bytes = (traffic[tdom]["dram_bytes_read"] + traffic[tdom]["dram_bytes_write"]) if traffic and tdom in traffic else None
    out = {"value": world * batch * args.steps / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / args.steps,
           "repeats": len(regions), "timed_regions_ms": [round(r, 4) for r in regions],
           "dtype": wl.dtype,
           "config": config_of(wl.name, world, batch, wl.depth),
           "e2e": {"value": world * batch * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
                   "h2d_bytes_per_step": world * wl.h2d(), "d2h_bytes_per_step": world * wl.d2h(),
                   "ms_per_step": ms_e2e / args.steps, "repeats": len(regions_e2e),
                   "h2d_gbs_per_rank": wl.h2d() / (ms_e2e / args.steps * 1e-3) / 1e9,
                   **({"h2d_bytes_offered_per_step": world * wl.host_sets[0].numel() * 4 // 50 * 48} if hasattr(wl, "host_sets") else {}),
                   **({"note": wl.e2e_note} if hasattr(wl, "e2e_note") else {})},
           "gpu_launches": int(launches),
           "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "frac_of_nominal_8000": achieved / 8000.0, "traffic": traffic_bytes,
                        # the limb kernel stages only the rows its samples can touch: what it really moves per second
                        "achieved_on_traffic": (traffic_bytes / (stage_ms[dominant] * 1e-3) / 1e9) if traffic_bytes else None,
                        "peak_source": peak_src, "algorithmic_bytes_per_launch": algo[dominant] * batch},
           "stage_ms": stage_ms,
           "stage_gbs": {k: algo[k] * batch / (stage_ms[k] * 1e-3) / 1e9 for k in stage_ms if k in algo},
           "whole_step_gbs": sum(algo.values()) * batch / (ms / args.steps * 1e-3) / 1e9,
           "humans_last_batch": wl.humans()}
    if ms_dev16:
        out["value_f16"] = {"value": world * batch * args.steps / (ms_dev16 * 1e-3), "unit": UNIT, "ms_per_step": ms_dev16 / args.steps,
                            "note": "device-resident float16 network output (the reference's amp O1 default), same streaming kernels"}
    if ms_e2e16:
        out["e2e_f16"] = {"value": world * batch * args.steps / (ms_e2e16 * 1e-3), "unit": UNIT,
                          "h2d_bytes_per_step": world * wl.h2d(2),
                          "note": "same call, float16 network output (the reference's amp O1 default, parse_skeletons.py:75-80)"}
    return out


def h2d_calibration(dev, world, host_tensor):
    """Pure pinned host -> device copy rate of this box with all `world` ranks copying at once (no kernels): the ceiling
    of every end-to-end number, GB/s per rank (max-time over ranks)."""
    import torch
    import torch.distributed as dist
    dst = torch.empty_like(host_tensor, device=dev)
    for _ in range(2):
        dst.copy_(host_tensor, non_blocking=True)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 8
    e0.record()
    for _ in range(reps):
        dst.copy_(host_tensor, non_blocking=True)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return reps * host_tensor.numel() * host_tensor.element_size() / (float(ms.item()) * 1e-3) / 1e9


def bind_rank(local_rank):
    """One process per GPU: keep the rank's threads (and therefore the first-touch placement of its pinned buffers) on
    the cores nvidia-smi reports as local to its GPU.  Best effort; returns what it did for the JSON line."""
    info = {}
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        for line in out.splitlines():
            cols = line.split()
            if cols and cols[0] == "GPU%d" % local_rank:
                # ... CPU Affinity, NUMA Affinity, GPU NUMA ID are the trailing columns
                tail = [c for c in cols if "-" in c and c.replace("-", "").replace(",", "").isdigit()]
                if tail:
                    info["cpu_affinity"] = tail[0]
                    cores = set()
                    for part in tail[0].split(","):
                        a, b = part.split("-") if "-" in part else (part, part)
                        cores.update(range(int(a), int(b) + 1))
                    avail = os.sched_getaffinity(0)
                    use = cores & avail
                    if use and use != avail:
                        os.sched_setaffinity(0, use)
                        info["bound_to"] = len(use)
                nums = [c for c in cols[1:] if c.isdigit()]
                if nums:
                    info["numa_node"] = int(nums[0])
    except Exception as e:  # noqa: BLE001
        info["error"] = str(e)[:80]
    return info


def run_ours(args, rank, world, local_rank, inputs):
    import torch

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    name = args.workload
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    def build(nm):
        if nm == "configs2":
            return MultiScale(nm, args, dev, inputs[nm], rank)
        if nm == "configs2_second_mode":  # same inputs, the north_star's operators (bilinear fusion, Gaussian-smoothed peaks)
            return MultiScale(nm, args, dev, inputs["configs2"], rank, route_kw={"interp": "linear", "smooth": (5, 1.0)})
        if nm == "configs4":
            import copy
            import sweep
            a4 = copy.copy(args)
            a4.batch = args.batch if name == "configs4" else 0
            # as an extra of the default run: a bounded sweep (the 10 000-image sweep is `--workload configs4`)
            a4.sweep_images = args.sweep_images or (10000 if name == "configs4" else 768)
            return sweep.SweepWorkload(nm, a4, dev, rank, world)
        return Refactored(nm, args, dev, inputs[nm], rank)

    wl = build(name)
    if name == "configs4":
        head = wl.run(args)
    else:
        head = measure(wl, args, dev, world, rank)
    clocks = sampler.stop() if rank == 0 else None  # sampled DURING the headline's timed regions only (the GPU idles under the CPU legs)
    out = {"metric": METRIC, "value": head.pop("value"), "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": head.pop("ms_per_step"), "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": head.pop("dtype"), "data": "synthetic"}
    out.update(head)
    if name in ("configs1", "configs3") and world >= 1:
        out["e2e"]["h2d_pinned_copy_gbs_per_rank_all_ranks_busy"] = h2d_calibration(dev, world, wl.host_sets[0])
    if world > 1:
        out["rank_binding"] = inputs.get("binding")
    if rank == 0 and world == 1 and not args.no_cpu_baseline and hasattr(wl, "cpu_baseline"):
        out["cpu_baseline"] = wl.cpu_baseline(args)
    first_sample = wl.sets_np[0] if hasattr(wl, "sets_np") else None
    wl.release()
    del wl
    torch.cuda.empty_cache()

    if world == 1 and not args.no_extras and name == "configs1":
        extras = {}
        import sweep
        have_model = sweep.model_available()
        for nm in ("configs2", "configs2_second_mode", "configs3", "configs4"):
            if (nm.split("_")[0] not in inputs and nm != "configs4") or (nm == "configs4" and not have_model):
                continue
            t0 = time.perf_counter()
            try:
                w2 = build(nm)
                if nm == "configs4":
                    r = w2.run(args)
                else:
                    r = measure(w2, args, dev, world, rank)
                    if not args.no_cpu_baseline:
                        r["cpu_baseline"] = w2.cpu_baseline(args)
                w2.release()
                del w2
            except Exception as e:  # noqa: BLE001 — an extra must never take the headline line down
                r = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
            torch.cuda.empty_cache()
            r["bench_seconds"] = round(time.perf_counter() - t0, 1)
            extras[nm] = r
        out["workloads"] = extras
        if not args.no_cpu_baseline:
            try:
                out["legacy"] = legacy_block()
            except Exception as e:  # noqa: BLE001
                out["legacy"] = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
            try:
                out["cpu_baselines"] = cpu_baseline_extra_legs(first_sample)
            except Exception as e:  # noqa: BLE001
                out["cpu_baselines"] = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}
        out["host"] = {"cpu_count": os.cpu_count(), "cpu_model": cpu_model()}
    out["clocks"] = clocks
    return out if rank == 0 else None


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return None


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
        # synthetic inputs first: they are rendered by forked worker processes, which must exist before CUDA does
        inputs = {}
        if world > 1:
            inputs["binding"] = bind_rank(local_rank)
        name = args.workload
        batch = args.batch or DEFAULT_BATCH[name]
        todo = [name] + (["configs2", "configs3"] if (world == 1 and name == "configs1" and not args.no_extras) else [])
        for nm in todo:
            b = batch if nm == name else DEFAULT_BATCH[nm]
            if nm == "configs2":
                inputs[nm] = gen_multiscale(rank, b)
            elif nm in ("configs1", "configs3"):
                inputs[nm] = gen_refactored(nm, rank, b)
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        out = run_ours(args, rank, world, local_rank, inputs)
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
            dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
