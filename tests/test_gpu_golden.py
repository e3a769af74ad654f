"""GPU tests against the committed golden vectors (outputs of the reference itself, tests/golden/make_golden.py)
and the edge cases the domain has: empty inputs, ragged map sizes, unsorted joint lists, whole batches."""
import glob
import os

import numpy as np
import pytest

from test_oracle_golden import CASES, load_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pp():
    from improved_body_parts_b200.pipeline import PostProcessor
    return PostProcessor(device=0, max_batch=8, max_bodies=1024, max_humans=256, max_candidates_per_limb=4096)


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p) for p in CASES])
def test_golden_through_stages(pp, path):
    """peaks + assignment from the fused maps of every golden case == what the reference produced."""
    import torch
    _, heat, paf, g = load_case(path)
    planes = np.concatenate([paf.transpose(2, 0, 1), heat.transpose(2, 0, 1)])[None]
    t = torch.from_numpy(np.ascontiguousarray(planes)).cuda()
    pp.find_peaks(t, 30)
    pp.assign_feat(t, 512)
    res = pp.fetch(1)
    jl = g["joint_list"].reshape(-1, 5)
    got = res.peaks[0]
    assert np.array_equal(got["x"], jl[:, 0].astype(np.int32)) and np.array_equal(got["y"], jl[:, 1].astype(np.int32))
    assert np.array_equal(got["score"], jl[:, 2]) and np.array_equal(got["part"], jl[:, 4].astype(np.int32))
    assert res.n_humans[0] == g["human_ids"].shape[0]
    for rec, gid, gsc in zip(res.humans[0], g["human_ids"], g["human_scores"]):
        assert np.array_equal(rec["peak_id"], gid) and rec["score"] == gsc


@pytest.mark.parametrize("path", [p for p in CASES if "crowd" in p or "synth_13" in p],
                         ids=lambda p: os.path.basename(p))
def test_golden_full_pipeline_from_network_output(pp, path):
    out, _, _, g = load_case(path)
    res = pp(out[None], 512)
    assert res.n_humans[0] == g["human_ids"].shape[0]
    for rec, gid, gsc in zip(res.humans[0], g["human_ids"], g["human_scores"]):
        assert np.array_equal(rec["peak_id"], gid) and rec["score"] == gsc


def test_legacy_unsorted_joint_list(oracle):
    from improved_body_parts_b200 import pafprocess, synth
    from oracle import np_oracle as O
    out, _ = synth.make_network_output(401, 6)
    heat, paf = O.fuse_flip(out)
    jl = O.joint_list(O.heatmap_nms(heat, 4))
    jl = np.ascontiguousarray(jl[np.random.default_rng(1).permutation(jl.shape[0])])
    up = O.limb_upsample(paf, 4)
    for img_h in (512, 64):
        oracle.process_paf(jl[None], up, img_h)
        pafprocess.process_paf(jl[None], up, img_h)
        assert pafprocess.get_num_humans() == oracle.get_num_humans()
        for h in range(oracle.get_num_humans()):
            assert [pafprocess.get_part_peak_id(h, p) for p in range(18)] == \
                   [oracle.get_part_peak_id(h, p) for p in range(18)]
            assert np.float32(pafprocess.get_score(h)) == np.float32(oracle.get_score(h))


def test_empty_image_and_bad_input():
    from improved_body_parts_b200 import _lib, pafprocess
    up = np.zeros((32, 32, 30), dtype=np.float32)
    assert pafprocess.process_paf(np.zeros((1, 0, 5), np.float32), up, 32) == 0
    assert pafprocess.get_num_humans() == 0
    with pytest.raises(_lib.IbpError) as e:
        pafprocess.process_paf(np.array([[[1, 1, 0.5, 0, 18]]], np.float32), up, 32)  # part id out of range
    assert e.value.code == -2
    with pytest.raises(_lib.IbpError) as e:
        pafprocess.process_paf(np.zeros((1, 2, 4), np.float32), up, 32)  # rows need 5 columns
    assert e.value.code == -1


def test_all_zero_maps_give_nothing(pp):
    res = pp(np.zeros((2, 2, 50, 128, 128), np.float32), 512)
    assert res.n_peaks.tolist() == [0, 0] and res.n_humans.tolist() == [0, 0]


@pytest.mark.parametrize("h,w", [(120, 160), (46, 62), (33, 47)])
def test_ragged_map_sizes(pp, oracle, h, w):
    """Maps that are not the 128x128 canonical size, including widths that defeat every vector/bulk fast path."""
    import torch
    from improved_body_parts_b200 import synth
    from oracle import np_oracle as O
    out, _ = synth.make_network_output(500 + h, 4, height=4 * h, width=4 * w)
    assert out.shape == (2, 50, h, w)
    heat, paf = O.fuse_flip(out)
    jl = O.joint_list(O.heatmap_nms(heat, 4))
    oracle.process_paf_feat(jl[None], np.ascontiguousarray(paf.transpose(2, 0, 1)), 4, 4 * h)
    want = oracle.humans()
    res = pp(out[None], 4 * h)
    assert np.array_equal(res.peaks[0]["x"], jl[:, 0].astype(np.int32))
    assert np.array_equal(res.peaks[0]["score"], jl[:, 2])
    assert res.n_humans[0] == len(want)
    for rec, (ids, score) in zip(res.humans[0], want):
        assert np.array_equal(rec["peak_id"], ids) and rec["score"] == score
    got_up = pp.upsample_limbs(torch.from_numpy(np.ascontiguousarray(paf.transpose(2, 0, 1))[None]).cuda(), 4)
    if (w * 4 * 30) % 4 == 0:
        assert np.array_equal(got_up[0].cpu().numpy(), O.limb_upsample(paf, 4))


def test_whole_batch_equals_per_image_and_is_idempotent(pp):
    """Size-independent properties at batch scale: batch result == image-by-image result; running twice is identical."""
    from improved_body_parts_b200 import synth
    batch = synth.make_batch(2000, 8, people=(1, 10))
    a = pp(batch, 512)
    b = pp(batch, 512)
    for i in range(8):
        single = pp(batch[i:i + 1], 512)
        assert np.array_equal(a.peaks[i], single.peaks[0]) and np.array_equal(a.humans[i], single.humans[0])
        assert np.array_equal(a.peaks[i], b.peaks[i]) and np.array_equal(a.humans[i], b.humans[i])


def test_fp16_network_output_pipeline(pp, oracle):
    from improved_body_parts_b200 import synth
    from oracle import np_oracle as O
    out, _ = synth.make_network_output(611, 5)
    out16 = out.astype(np.float16)
    heat, paf = O.fuse_flip(out16)
    jl = O.joint_list(O.heatmap_nms(heat, 4))
    oracle.process_paf_feat(jl[None], np.ascontiguousarray(paf.transpose(2, 0, 1)), 4, 512)
    want = oracle.humans()
    res = pp(out16[None], 512)
    assert np.array_equal(res.peaks[0]["score"], jl[:, 2])
    assert [tuple(r["peak_id"]) for r in res.humans[0]] == [tuple(ids) for ids, _ in want]
    assert [r["score"] for r in res.humans[0]] == [s for _, s in want]
