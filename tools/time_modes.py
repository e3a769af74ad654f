"""Kernel times of the second operator mode on the configs[2] shapes (batch 32): python tools/time_modes.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    c2 = bench.gen_multiscale(0, B)
    shapes = synth.multiscale_shapes(bench.MS_H, bench.MS_W, bench.MS_SS)
    pp = PostProcessor(device=0, max_batch=B)
    pp.handle.set_profiling(True)
    ts = [torch.from_numpy(a).cuda() for a in c2]
    in_bytes = sum(t.numel() * 4 for t in ts)
    out_bytes = B * (50 * 8 + 18 * 4) * bench.MS_H * bench.MS_W
    for interp in ("linear", "cubic"):
        for _ in range(3):
            fused, heat32 = pp.fuse_multiscale(ts, [s[1] for s in shapes], bench.MS_H, bench.MS_W, interp=interp)
            torch.cuda.synchronize()
        ms = pp.handle.stage_ms()["fuse"]
        print("%s fuse: %.3f ms  %.0f GB/s algorithmic (%.2f GB)" % (interp, ms, (in_bytes + out_bytes) / ms / 1e6, (in_bytes + out_bytes) / 1e9))
    for _ in range(3):
        sm = pp.smooth_gaussian(heat32, 5, 1.0)
        torch.cuda.synchronize()
    ms = pp.handle.stage_ms()["smooth"]
    print("gauss 5x5 on (%d,18,%d,%d): %.3f ms  %.0f GB/s" % (B, bench.MS_H, bench.MS_W, ms, 2 * heat32.numel() * 4 / ms / 1e6))


if __name__ == "__main__":
    main()
