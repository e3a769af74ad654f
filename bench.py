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
    "fuse": (2 * 50 + 50) * FEAT * FEAT * 4,   # read both flips, write the fused planes
    "peaks": 18 * FEAT * FEAT * 4,             # keypoint planes read once
    "limbs": 30 * FEAT * FEAT * 4,             # limb planes staged once
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="images in the cpu_baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def make_inputs(rank, n_sets=2):
    """n_sets different batches per rank so that consecutive steps never re-read the same 419 MB."""
    from improved_body_parts_b200 import synth
    return [synth.make_batch(SEED + 100000 * rank + 1000 * s, BATCH, people=(1, 10), height=IMG, width=IMG)
            for s in range(n_sets)]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
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


def cpu_baseline(n_images):
    """The reference's CPU route (oracle/cpu_path.py) on a bounded sample of the same workload, all host cores."""
    from improved_body_parts_b200 import synth
    from oracle import cpu_path
    sample = synth.make_batch(SEED, n_images, people=(1, 10), height=IMG, width=IMG)
    secs, procs = cpu_path.time_batch(sample, IMG)
    return {"value": n_images / secs, "unit": UNIT, "cores": procs,
            "kind": "reference" if cpu_path.cpp_kind() == "reference" else "port",
            "sample": "%d images of the same workload (seed %d..), one pass, %d processes x 1 thread; C++ stage = %s, "
                      "Python stage = restated scipy.maximum_filter + cv2.resize calls (oracle/cpu_path.py)"
                      % (n_images, SEED, procs, "unmodified reference pafprocess.cpp (oracle/_ref)"
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


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from improved_body_parts_b200.pipeline import PostProcessor

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    host_sets = [torch.from_numpy(b).pin_memory() for b in make_inputs(rank)]
    dev_sets = [h.to(dev) for h in host_sets]
    pp = PostProcessor(device=local_rank, max_batch=BATCH)

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
        e1.record()
        torch.cuda.synchronize(dev)
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        barrier()
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- device-resident throughput (`value`) ----
    def step_dev(s):
        pp.enqueue(dev_sets[s % len(dev_sets)], IMG)

    for s in range(args.warmup):
        step_dev(s)
    pp.fetch(BATCH)
    launches0 = pp.handle.launches
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms = timed(step_dev, args.steps)
    launches = pp.handle.launches - launches0
    last = pp.fetch(BATCH)

    # ---- end to end through the public API with HOST buffers (`e2e`) ----
    def step_e2e(s):
        pp.enqueue(host_sets[s % len(host_sets)], IMG)   # pinned host -> device inside the timed region
        pp.fetch(BATCH)                                  # packed skeletons device -> host
    for s in range(min(args.warmup, 3)):
        step_e2e(s)
    ms_e2e = timed(step_e2e, args.steps)
    clocks = sampler.stop() if rank == 0 else None

    # ---- per-kernel device times for the roofline (events on the launching stream inside the library) ----
    pp.handle.set_profiling(True)
    acc = {}
    for s in range(args.steps):
        pp.enqueue(dev_sets[s % len(dev_sets)], IMG)
        for k, v in pp.handle.stage_ms().items():
            acc.setdefault(k, []).append(v)
    pp.handle.set_profiling(False)
    stage_ms = {k: float(np.mean(v)) for k, v in acc.items()}
    if rank != 0:
        return None

    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_file):
        peak, peak_src = float(json.load(open(peaks_file))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"
    dominant = max((k for k in stage_ms if k in ALGO_BYTES), key=lambda k: stage_ms[k])
    achieved = ALGO_BYTES[dominant] * BATCH / (stage_ms[dominant] * 1e-3) / 1e9
    h2d = host_sets[0].numel() * host_sets[0].element_size()
    view_bytes = BATCH * (4 + 18 * 4 + pp.handle.limits.max_peaks_per_part * 18 * 16 + 4 +
                          pp.handle.limits.max_humans * 88 + 4)
    out = {"metric": METRIC, "value": world * BATCH * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": WORKLOAD, "batch_per_gpu": BATCH, "global_batch": world * BATCH,
                      "parallelism": "images sharded, %d rank(s), no collective" % world,
                      "l2_policy": "2 rotating input sets of 419 MB per rank, each larger than the 126 MB L2"},
           "e2e": {"value": world * BATCH * args.steps / (ms_e2e * 1e-3), "unit": UNIT,
                   "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": view_bytes},
           "gpu_launches": int(launches),
           "clocks": clocks,
           "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                        "algorithmic_bytes_per_launch": ALGO_BYTES[dominant] * BATCH},
           "stage_ms": stage_ms,
           "stage_gbs": {k: ALGO_BYTES[k] * BATCH / (stage_ms[k] * 1e-3) / 1e9 for k in stage_ms if k in ALGO_BYTES},
           "humans_last_batch": int(last.n_humans.sum())}
    if not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args.cpu_sample or max(8, 4 * (os.cpu_count() or 1)))
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
