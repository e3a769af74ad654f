"""B200-native keypoint-assignment path of SimplePose (network output -> skeletons).

Host side: Python/PyTorch plumbing over the C ABI of libibp_b200.so (include/ibp.h); all compute is hand-written
sm_100a CUDA.  Sub-modules mirror the reference's call surface:

    pafprocess        utils/pafprocess (SWIG module): process_paf + six getters
    parse_skeletons   utils/parse_skeletons.py: predict_refactor, heatmap_nms, ...
    common            utils/common.py: Human, BodyPart, CocoPart
    config            config/config.py + utils/config_reader.py constants
    pipeline          the batched device-resident pipeline and the multi-GPU sharding helpers
    synth             synthetic network outputs for tests and benchmarks
"""
from . import config  # noqa: F401

__all__ = ["config", "pafprocess", "parse_skeletons", "common", "pipeline", "synth", "build_library"]


def build_library(force=False):
    """Compile libibp_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("_ibp_build", os.path.join(__path__[0], "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.build(force=force)
