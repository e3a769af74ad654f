"""Small invocations of every hand-synchronised kernel for compute-sanitizer (tools/sanitize.sh): one configs[1] batch
through the persistent streaming trio (twice: carried work counters / tickets), one crowd image through the handle
bench.py sizes for configs[3] (limbs_kernel<float,512>), one float16 batch, one configs[2] image through the four-kernel
original route.  Results are checked against the oracle so that a tool-induced slowdown cannot hide a wrong answer."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from improved_body_parts_b200 import synth
    from improved_body_parts_b200.pipeline import PostProcessor
    from oracle import np_oracle as O
    from oracle.cpu_lib import Oracle
    orc = Oracle()
    which = set(sys.argv[1:]) or {"configs1", "crowd", "f16", "configs2"}

    def check(res, batch, img_h=512):
        for b in range(batch.shape[0]):
            heat, paf = O.fuse_flip(batch[b])
            jl = O.joint_list(O.heatmap_nms(heat, 4))
            orc.process_paf_feat(jl[None], np.ascontiguousarray(paf.transpose(2, 0, 1)), 4, img_h)
            want = orc.humans()
            assert res.n_peaks[b] == jl.shape[0] and np.array_equal(res.peaks[b]["score"], jl[:, 2])
            assert [tuple(r["peak_id"]) for r in res.humans[b]] == [tuple(i) for i, _ in want]

    if "configs1" in which:
        batch = synth.make_batch(2000, 6, people=(1, 10))
        pp = PostProcessor(device=0, max_batch=6)
        for _ in range(2):
            check(pp(batch, 512), batch)
        print("configs1 ok: %d launches" % pp.handle.launches, flush=True)
    if "f16" in which:
        batch = synth.make_batch(2000, 4, people=(1, 10)).astype(np.float16)
        pp = PostProcessor(device=0, max_batch=4)
        check(pp(batch, 512), batch)
        print("f16 ok: %d launches" % pp.handle.launches, flush=True)
    if "crowd" in which:
        batch = synth.make_batch(3000, 1, people=(40, 60), scale=(0.2, 0.45))
        pp = PostProcessor(device=0, max_batch=1, max_peaks_per_part=256, max_candidates_per_limb=1024, max_bodies=512, max_humans=256)
        check(pp(batch, 512), batch)
        print("crowd ok: %d launches" % pp.handle.launches, flush=True)
    if "configs2" in which:
        img_h, img_w, ss = 240, 320, (0.5, 1.0, 2.0)
        shapes = synth.multiscale_shapes(img_h, img_w, ss)
        outs = synth.make_multiscale_output(2001, 4, img_h, img_w, ss)[0]
        pp = PostProcessor(device=0, max_batch=1)
        ts = [torch.from_numpy(o[None]).cuda() for o in outs]
        table, counts, assoc = pp.original_route(ts, [s[1] for s in shapes], img_h, img_w)[0]
        from oracle import cpu_path
        heat, paf = cpu_path.predict_fuse_cv2(outs, img_h, img_w, [s[0] for s in shapes])
        peaks = O.find_peaks(heat, 0.1, 2)
        cl, sp = O.find_connections(peaks, paf, img_h)
        want, _ = O.find_humans(cl, sp, peaks)
        assert np.array_equal(assoc, want)
        print("configs2 ok: %d launches, %d people" % (pp.handle.launches, assoc.shape[0]), flush=True)


if __name__ == "__main__":
    main()
