"""CPU tests of everything around the kernels: the introsort emulation against the real std::sort, the C ABI
(library loads, every declared symbol is exported, no compute without a GPU), argument validation of the pafprocess
mirror, sharding/merging and a world-size-2 gloo gather."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SORT_HARNESS = r"""
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>
#include "stl_sort_emul.h"
struct Cand { int a, b; float limb, rank, len; };
static bool rank_ge(Cand p, Cand q) { return p.rank >= q.rank; }
// returns 1 when the emulation reproduces std::sort's permutation (or reports the case as undefined)
extern "C" int check(const float* keys, int n, int* undefined_out) {
    std::vector<Cand> v(n);
    std::vector<uint64_t> e(n);
    for (int i = 0; i < n; i++) {
        v[i].a = i; v[i].b = 0; v[i].limb = 0; v[i].len = 0; v[i].rank = keys[i];
        uint32_t bits; std::memcpy(&bits, &keys[i], 4);
        e[i] = ((uint64_t)bits << 32) | (uint32_t)i;
    }
    bool defined = ibp_sortemu::sort_like_libstdcxx(e.data(), n);
    *undefined_out = defined ? 0 : 1;
    if (!defined) return 1;            // the real sort would leave the array: never call it
    std::sort(v.begin(), v.end(), rank_ge);
    for (int i = 0; i < n; i++) if ((int)(uint32_t)e[i] != v[i].a) return 0;
    return 1;
}
"""


@pytest.fixture(scope="module")
def sort_lib(tmp_path_factory):
    d = tmp_path_factory.mktemp("sortemu")
    src = d / "harness.cpp"
    src.write_text(SORT_HARNESS)
    so = d / "harness.so"
    subprocess.check_call(["g++", "-O2", "-std=c++14", "-fPIC", "-shared", "-I",
                           os.path.join(ROOT, "improved-body-parts_b200", "csrc"), str(src), "-o", str(so)])
    lib = ctypes.CDLL(str(so))
    lib.check.restype = ctypes.c_int
    lib.check.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    return lib


def test_sort_emulation_equals_libstdcxx(sort_lib):
    """Duplicate-heavy keys under the reference's non-strict comparator (pafprocess.cpp:109, 333-335)."""
    rng = np.random.default_rng(0)
    n_defined = n_undefined = 0
    for trial in range(3000):
        n = int(rng.integers(1, 400))
        distinct = int(rng.integers(1, max(2, n)))
        keys = rng.integers(0, distinct, size=n).astype(np.float32) / 7.0 + 0.1
        if trial % 3 == 0:
            keys = rng.random(n).astype(np.float32)  # no duplicates
        und = ctypes.c_int(0)
        ok = sort_lib.check(keys.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), n, ctypes.byref(und))
        assert ok == 1, "emulation diverged from std::sort at trial %d (n=%d)" % (trial, n)
        n_undefined += und.value
        n_defined += 1 - und.value
    assert n_defined > 1500  # the check is not vacuous


def test_small_n_ties_come_out_reversed(sort_lib):
    """n <= 16: insertion sort with >= puts equal keys in reverse input order — the GPU's (score, index desc) key."""
    keys = np.array([0.5, 0.7, 0.5, 0.7, 0.5], dtype=np.float32)
    und = ctypes.c_int(0)
    assert sort_lib.check(keys.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), 5, ctypes.byref(und)) == 1


def test_library_exports_every_declared_symbol():
    from improved_body_parts_b200 import _lib
    header = open(os.path.join(ROOT, "include", "ibp.h")).read()
    declared = set(re.findall(r"\b(ibp_[a-z_0-9]+)\s*\(", header))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    L = _lib.lib()
    for name in declared:
        assert hasattr(L, name), name


def test_no_cpu_fallback_without_gpu():
    import torch
    from improved_body_parts_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_lib.IbpError) as e:
        _lib.Handle(0)
    assert e.value.code == -6
    from improved_body_parts_b200 import pafprocess
    with pytest.raises(_lib.IbpError):
        pafprocess.process_paf(np.zeros((1, 1, 5), np.float32), np.zeros((8, 8, 30), np.float32), 8)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "improved-body-parts_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("the oracle", "").replace("CPU oracle", "") or f == "build.py", f


def test_pafprocess_typemap_errors():
    """numpy.i IN_ARRAY3 behaviour (pafprocess.i:14): wrong rank -> ValueError, unconvertible -> TypeError."""
    from improved_body_parts_b200 import pafprocess
    with pytest.raises(ValueError):
        pafprocess.process_paf(np.zeros((3, 5), np.float32), np.zeros((8, 8, 30), np.float32), 8)
    with pytest.raises(TypeError):
        pafprocess.process_paf([["a"]], np.zeros((8, 8, 30), np.float32), 8)


def test_defaults_equal_reference_constants():
    """pafprocess.h:6-18 and utils/config:16-27."""
    from improved_body_parts_b200 import _lib
    p = _lib.default_params()
    assert (p.limb_sample_thresh, p.step_ratio, p.min_parts, p.max_steps, p.len_rate) == \
        (np.float32(0.1), np.float32(0.8), 2, 20, 16)
    assert (p.min_human_score, p.merge_tolerance) == (np.float32(0.45), np.float32(0.7))
    assert (p.weight_limb, p.weight_a, p.weight_b, p.upsample, p.refine_window) == (0.5, 0.25, 0.25, 4, 2)
    from improved_body_parts_b200.config import config_reader
    param, model = config_reader()
    assert param["thre2"] == 0.1 and param["mid_num"] == 20 and param["len_rate"] == 16
    assert model["stride"] == 4 and model["max_downsample"] == 64 and model["padValue"] == 128


def test_shard_and_merge_any_world_size():
    from improved_body_parts_b200.pipeline import merge_shards, shard_indices
    n = 37
    for world in (1, 2, 3, 8):
        shards = [[(i, "img%d" % i) for i in shard_indices(n, r, world)] for r in range(world)]
        assert merge_shards(shards, n) == ["img%d" % i for i in range(n)]
    with pytest.raises(ValueError):
        merge_shards([[(0, "a")]], 2)


GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, %r)
import torch.distributed as dist
from improved_body_parts_b200.pipeline import shard_indices, gather_results
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
local = [(i, {"image": i, "rank": r}) for i in shard_indices(11, r, w)]
full = gather_results(local, 11)
assert [f["image"] for f in full] == list(range(11))
assert all(f["rank"] == f["image"] %% w for f in full)
dist.destroy_process_group()
print("ok", r)
"""


def test_gather_world_size_2_gloo(tmp_path):
    """The N>1 host path: images[r::W] per rank, object gather, order independent of W."""
    script = tmp_path / "worker.py"
    script.write_text(GLOO_WORKER % ROOT)
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29611", str(script)],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    assert res.stdout.count("ok") == 2


def test_no_contracted_packed_multiply_add():
    """ptxas contracts a packed float32x2 multiply feeding a packed add into ONE FFMA2 (a single
    rounding), which would break parity with OpenCV's separate multiply and add.  The kernels therefore issue every
    packed add as fma(x, one, y) with `one` a run-time 1.0f, which shows in the SASS as a scalar uniform-register
    multiplier (URn.F32) — any other FFMA2 would be a fused multiply-add that slipped in."""
    import re
    import shutil
    import subprocess
    from improved_body_parts_b200 import _lib
    tool = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(tool) or not os.path.exists(_lib.LIB_PATH):
        pytest.skip("cuobjdump or the library is not available")
    sass = subprocess.run([tool, "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    # the second operator mode (fuse_linear_kernel, gauss_smooth_kernel) is held to a tolerance and fuses on purpose
    exact = "".join(f for f in sass.split("Function :") if not re.match(r"\s*\S*(fuse_linear|gauss_smooth)", f))
    packed = re.findall(r"\bFFMA2\s+([^;]*);", exact)
    assert len(packed) > 50, "the multi-scale kernel is expected to use packed float32x2 arithmetic"
    bad = [p for p in packed if not re.match(r"R\d+, R\d+(\.reuse)?\.F32x2\.HI_LO, UR\d+\.F32, R\d+(\.reuse)?\.F32x2\.HI_LO\s*$", p.strip())]
    assert not bad, "fused packed multiply-add in the library: %s" % bad[:3]
