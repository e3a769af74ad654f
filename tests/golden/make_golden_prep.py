"""Golden vectors for the two ends of the path, produced by RUNNING THE REFERENCE in the build container:

  * prep_6x.npz — the tensor utils/parse_skeletons.py:28-103 predict_refactor hands to the model (l.72-76): the call is
    made with a stand-in model that records its input and returns zeros; torch.Tensor.cuda is the identity (no GPU in
    the build container).  Images are seeded random uint8, including sizes that need no padding and odd sizes.
  * coco_7x.npz — evaluate.py:182-209 append_result (the --run_refactor branch) on Human objects built exactly as
    evaluate.py:111-129 builds them from the getter protocol; evaluate.py parses the command line and loads a
    checkpoint at import time, so the function's source is executed in a namespace that holds its three globals
    (args.run_refactor, NUM_KEYPOINTS, ORDER_COCO taken from the module's own source text).
"""
import ast
import hashlib
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from oracle import ref_python  # noqa: E402

PREP_CASES = [("prep_61", 61, 333, 427), ("prep_62", 62, 512, 512), ("prep_63", 63, 480, 640), ("prep_64", 64, 65, 1)]
COCO_CASES = [("coco_71", 71, 7, 60), ("coco_72", 72, 1, 15), ("coco_73", 73, 0, 5)]


def prep_case(ref, seed, h, w):
    rng = np.random.default_rng(seed)
    image = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    seen = {}

    def model(x):
        seen["x"] = x.numpy().copy()
        return [[torch.zeros((2, 50, x.shape[1] // 4, x.shape[2] // 4))]]

    old = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        ref.parse_skeletons.predict_refactor(image, model, dict(ref.test_cfg), ref.model_cfg, "synthetic", flip_avg=True, config=ref.config)
    finally:
        torch.Tensor.cuda = old
    x = seen["x"]
    assert x.dtype == np.float32 and x.shape[0] == 2 and x.shape[3] == 3
    return dict(seed=seed, img_h=h, img_w=w, shape=np.array(x.shape), sha256=hashlib.sha256(np.ascontiguousarray(x).tobytes()).hexdigest(),
                probe=x[:, ::37, ::41, :].copy())


def load_append_result():
    src = open(os.path.join(ref_python.REF_ROOT, "evaluate.py")).read()
    tree = ast.parse(src)
    ns = {"np": np}
    for node in tree.body:
        if isinstance(node, ast.Assign) and any(isinstance(t, ast.Name) and t.id in ("ORDER_COCO", "NUM_KEYPOINTS", "NUM_COCO_KEYPOINTS") for t in node.targets):
            exec(compile(ast.Module([node], []), "evaluate.py", "exec"), ns)
        if isinstance(node, ast.FunctionDef) and node.name == "append_result":
            exec(compile(ast.Module([node], []), "evaluate.py", "exec"), ns)
    if "NUM_KEYPOINTS" not in ns:  # imported from utils.parse_skeletons in the reference
        ns["NUM_KEYPOINTS"] = 18
    if "NUM_COCO_KEYPOINTS" not in ns:
        ns["NUM_COCO_KEYPOINTS"] = 17
    ns["args"] = types.SimpleNamespace(run_refactor=True)
    return ns["append_result"]


def coco_case(ref, append_result, seed, n_humans, n_peaks):
    sys.path.insert(0, ref_python.REF_ROOT)
    from utils.common import BodyPart, Human  # the reference's own result types (common.py:39-51, 267-288)
    rng = np.random.default_rng(seed)
    xy = rng.integers(0, 640, (n_peaks, 2)).astype(np.int32)
    ids = np.full((n_humans, 18), -1, dtype=np.int32)
    scores = rng.uniform(0.45, 2.0, n_humans).astype(np.float32)
    humans = []
    for k in range(n_humans):
        human = Human([])
        for part in range(18):
            if rng.uniform() < 0.7:
                pid = int(rng.integers(0, n_peaks))
                ids[k, part] = pid
                human.body_parts[part] = BodyPart('%d-%d' % (k, part), part, float(xy[pid, 0]), float(xy[pid, 1]), 0.5)
        human.score = float(scores[k])
        humans.append(human)
    out = []
    append_result(123, humans, out)
    rows = np.zeros((n_humans, 52), dtype=np.float32)
    for k, r in enumerate(out):
        rows[k, :51] = np.asarray(r["keypoints"], dtype=np.float64)
        rows[k, 51] = r["score"]
    return dict(seed=seed, xy=xy, ids=ids, scores=scores, rows=rows)


def main():
    ref = ref_python.load()
    for name, seed, h, w in PREP_CASES:
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **prep_case(ref, seed, h, w))
        print("wrote", name)
    append_result = load_append_result()
    for name, seed, nh, npk in COCO_CASES:
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **coco_case(ref, append_result, seed, nh, npk))
        print("wrote", name)


if __name__ == "__main__":
    main()
