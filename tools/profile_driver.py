"""A handful of launches of every hot kernel, for ncu (one --set full capture covers them all):
    configs[1] batch 64: 3 batches float32, 3 batches float16 (peaks_stream, index_bands, limbs_stream, each element type)
    configs[3] batch 64 crowd handle: 2 batches (peaks_kernel / limbs_kernel<float,512>)
    configs[2] batch 64: 2 batches (fuse_multiscale, peaks_stream<1>, py_connections, py_humans), then 2 in the second operator mode
                        (fuse_linear, gauss_smooth)
Usage: python tools/profile_driver.py [configs1] [configs3] [configs2]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    which = set(sys.argv[1:]) or {"configs1", "configs3", "configs2"}
    c1 = bench.gen_refactored("configs1", 0, 64, n_sets=1)[0] if "configs1" in which else None
    c3 = bench.gen_refactored("configs3", 0, 64, n_sets=1)[0] if "configs3" in which else None
    c2 = bench.gen_multiscale(0, 64) if "configs2" in which else None
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    if c1 is not None:
        pp = PostProcessor(device=0, max_batch=64)
        d32 = torch.from_numpy(c1).cuda()
        d16 = d32.half()
        for d in (d32, d32, d32, d16, d16, d16):
            pp.enqueue(d, 512)
            torch.cuda.synchronize()
        print("configs1:", int(pp.fetch(64).n_humans.sum()), "humans", flush=True)
    if c3 is not None:
        pp = PostProcessor(device=0, max_batch=64, max_peaks_per_part=256, max_candidates_per_limb=1024, max_bodies=512, max_humans=256)
        d = torch.from_numpy(c3).cuda()
        for _ in range(2):
            pp.enqueue(d, 512)
            torch.cuda.synchronize()
        print("configs3:", int(pp.fetch(64).n_humans.sum()), "humans", flush=True)
    if c2 is not None:
        shapes = synth.multiscale_shapes(bench.MS_H, bench.MS_W, bench.MS_SS)
        pp = PostProcessor(device=0, max_batch=64)
        ts = [torch.from_numpy(a).cuda() for a in c2]
        for _ in range(2):
            pp.enqueue_original_route(ts, [s[1] for s in shapes], bench.MS_H, bench.MS_W)
            torch.cuda.synchronize()
        print("configs2:", sum(p.shape[0] for _, _, p in pp.fetch_original_route(64)), "people", flush=True)
        for _ in range(2):  # second operator mode: fuse_linear_kernel, gauss_smooth_kernel (+ the same three kernels behind them)
            pp.enqueue_original_route(ts, [s[1] for s in shapes], bench.MS_H, bench.MS_W, interp="linear", smooth=(5, 1.0))
            torch.cuda.synchronize()
        print("configs2 second mode:", sum(p.shape[0] for _, _, p in pp.fetch_original_route(64)), "people", flush=True)


if __name__ == "__main__":
    main()
