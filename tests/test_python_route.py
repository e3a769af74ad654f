"""The reference's PYTHON assignment route (find_connections / find_humans, utils/parse_skeletons.py:324-599): oracle
against golden vectors made by the reference's own functions (tests/golden/make_golden_py.py); CUDA kernels (through
the C ABI) against both.  Everything is float64 and must be bit-exact."""
import glob
import hashlib
import os
import sys

import numpy as np
import pytest

from oracle import np_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLD)
CASES = sorted(glob.glob(os.path.join(GOLD, "py_*.npz")))
TEST_CFG = dict(thre1=0.1, thre2=0.1, connect_ration=0.8, mid_num=20, len_rate=16, connection_tole=0.7, offset_radius=2,
                remove_recon=0)


def load_py(path):
    from make_golden_py import inputs
    g = np.load(path)
    heat, paf, peaks = inputs(int(g["seed"]), int(g["people"]), int(g["img_h"]), int(g["img_w"]))
    assert hashlib.sha256(np.ascontiguousarray(paf).tobytes()).hexdigest() == str(g["paf_sha256"]), "generator drifted"
    return heat, paf, peaks, g


def check_connections(connected, special, g):
    from make_golden_py import pack_connections
    rows, counts = pack_connections(connected)
    assert list(special) == list(g["special"])
    assert np.array_equal(counts, g["conn_counts"])
    assert np.array_equal(rows, g["conn_rows"])


def test_golden_present():
    assert len(CASES) >= 5


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_oracle_matches_reference(path):
    _, paf, peaks, g = load_py(path)
    cl, sp = O.find_connections(peaks, paf, int(g["img_h"]))
    check_connections(cl, sp, g)
    assoc, cand = O.find_humans(cl, sp, peaks)
    assert np.array_equal(assoc, g["people_assoc"])
    assert np.array_equal(np.asarray(cand, dtype=np.float64).reshape(-1, 4), g["candidates"])


@pytest.mark.gpu
@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_gpu_connections_and_humans_exact(path):
    from improved_body_parts_b200 import parse_skeletons as PS
    _, paf, peaks, g = load_py(path)
    cl, sp = PS.find_connections(peaks, paf, int(g["img_h"]), TEST_CFG)
    check_connections(cl, sp, g)
    assoc, cand = PS.find_humans(cl, sp, peaks, TEST_CFG)
    assert np.array_equal(assoc, g["people_assoc"])
    assert np.array_equal(np.asarray(cand, dtype=np.float64).reshape(-1, 4), g["candidates"])


@pytest.mark.gpu
def test_gpu_python_route_edge_cases():
    """coincident joints (length 0), a one-sample limb, equal scores (stable order), no peaks at all."""
    from improved_body_parts_b200 import parse_skeletons as PS
    H, W = 64, 80
    paf = np.full((H, W, 30), 0.5)
    peaks = [[] for _ in range(18)]
    peaks[1] = [(20.0, 20.0, np.float32(0.9), 0), (40.25, 20.5, np.float32(0.9), 1)]       # necks
    peaks[0] = [(20.0, 20.0, np.float32(0.8), 2), (20.4, 20.3, np.float32(0.8), 3), (40, 12, np.float32(0.8), 4)]  # noses
    peaks[14] = [(22.5, 10.5, np.float32(0.7), 5)]
    for fn_name in ("find_connections",):
        cl_o, sp_o = O.find_connections(peaks, paf, H)
        cl_g, sp_g = PS.find_connections(peaks, paf, H, TEST_CFG)
        assert sp_o == sp_g
        for a, b in zip(cl_o, cl_g):
            assert (len(a) == 0 and len(b) == 0) or np.array_equal(np.asarray(a).reshape(-1, 6), b)
    a_o, _ = O.find_humans(cl_o, sp_o, peaks)
    a_g, _ = PS.find_humans(cl_g, sp_g, peaks, TEST_CFG)
    assert np.array_equal(a_o, a_g)
    empty = [[] for _ in range(18)]
    cl_g, sp_g = PS.find_connections(empty, paf, H, TEST_CFG)
    assert sp_g == list(range(30)) and all(len(c) == 0 for c in cl_g)
    a_g, c_g = PS.find_humans(cl_g, sp_g, empty, TEST_CFG)
    assert a_g.shape == (0, 20, 2)
