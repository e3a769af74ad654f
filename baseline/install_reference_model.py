"""Recipe for BASELINE.json configs[4] ("10k-image sweep ... fed in-stream by random-init IMHN (posenet) forward").

The IMHN network is OUTSIDE the product (north_star: "the random-init PyTorch IMHN forward stays on cuDNN, outside the
product"); it is the feeder of the sweep.  This script copies the reference's UNMODIFIED network definition
(/root/reference/models/{__init__,posenet,layers_transposed,loss_model,loss_model_parallel}.py — what
evaluate.py:24 / demo_image.py:19 import) into baseline/_ref/models/, which is git-ignored (never in the history) but
NOT gpurun-ignored, so it travels to the GPU box like the other reference artefact, oracle/_ref.  No file of the
reference is edited; `thop` (a FLOP counter the reference imports and never uses at inference) is stubbed by the loader.

    python baseline/install_reference_model.py        (build container only: /root/reference does not exist elsewhere)
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/models"
DST = os.path.join(HERE, "_ref", "models")
FILES = ["__init__.py", "posenet.py", "layers_transposed.py", "loss_model.py", "loss_model_parallel.py"]


def install():
    if not os.path.isdir(SRC):
        return None
    os.makedirs(DST, exist_ok=True)
    for f in FILES:
        shutil.copyfile(os.path.join(SRC, f), os.path.join(DST, f))
    return DST


if __name__ == "__main__":
    d = install()
    print(d or "reference tree not present: nothing installed")
    sys.exit(0 if d else 1)
