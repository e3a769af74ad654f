"""TEST INFRASTRUCTURE ONLY — imports the reference's own Python modules from /root/reference (read-only).

Only usable in the build container (the GPU box has no /root/reference): golden-vector generation
(tests/golden/make_golden.py) and the oracle-vs-reference tests, which skip when the tree is absent.
Nothing here is copied from the reference; absent third-party imports are stubbed (matplotlib, thop) and the
configobj-based `config_reader()` is replaced by the literal values of /root/reference/utils/config:12-40.
"""
import contextlib
import io
import os
import sys
import types

REF_ROOT = "/root/reference"


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "utils"))


_cache = {}


def load():
    """Returns a namespace with parse_skeletons, util, config (Canonical), Heatmapper, test_cfg, model_cfg."""
    if _cache:
        return _cache["ns"]
    if not available():
        raise RuntimeError("reference tree not present")
    for name in ("matplotlib", "matplotlib.pyplot", "thop"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    if "matplotlib" in sys.modules and "matplotlib.pyplot" in sys.modules:
        setattr(sys.modules["matplotlib"], "pyplot", sys.modules["matplotlib.pyplot"])
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import warnings
    with warnings.catch_warnings(), contextlib.redirect_stdout(io.StringIO()):
        warnings.simplefilter("ignore")
        from utils import parse_skeletons, util  # noqa
        from config.config import GetConfig
        from py_cocodata_server.py_data_heatmapper import Heatmapper
        config = GetConfig("Canonical")
    ns = types.SimpleNamespace()
    ns.parse_skeletons = parse_skeletons
    ns.util = util
    ns.config = config
    ns.Heatmapper = Heatmapper
    # /root/reference/utils/config:12-40 through config_reader.py:53-76
    ns.test_cfg = dict(scale_search=[0.5, 1.0, 1.5, 2.0, 3.0], rotation_search=[0.0], thre1=0.1, thre2=0.1,
                       connect_ration=0.8, min_num=4, mid_num=20, len_rate=16, connection_tole=0.7, crop_ratio=2.5,
                       bbox_ratio=0.25, offset_radius=2, remove_recon=0, starting_range=0.8, ending_range=2.0)
    ns.model_cfg = dict(boxsize=512, padValue=128, np=12, stride=4, max_downsample=64)
    _cache["ns"] = ns
    return ns
