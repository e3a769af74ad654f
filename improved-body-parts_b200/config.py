"""Read-only constants of the keypoint-assignment path, restated from the reference's three config mechanisms
(SURVEY.md §5): config/config.py:51-154 (channel layout), utils/config:12-40 via utils/config_reader.py:53-76
(test-time thresholds) and utils/pafprocess/pafprocess.h:6-27 (the C++ duplicates of the same numbers)."""
import numpy as np

NUM_PARTS = 18            # config.py:64 / pafprocess.h:11
NUM_HEAT = NUM_PARTS + 2  # + 2 background channels (config.py:67-70)
NUM_LIMBS = 30            # config.py:125 / pafprocess.h:19
NUM_LAYERS = NUM_LIMBS + NUM_HEAT  # channel order: [0,30) limb, [30,48) keypoint, [48,50) background (config.py:129)

PARTS = ["nose", "neck", "Rsho", "Relb", "Rwri", "Lsho", "Lelb", "Lwri", "Rhip", "Rkne", "Rank", "Lhip", "Lkne",
         "Lank", "Reye", "Leye", "Rear", "Lear"]

# config.py:114-123 (limb_from/limb_to) == pafprocess.h:20-27 (COCOPAIRS)
LIMBS = np.array([[1, 0], [1, 14], [1, 15], [1, 16], [1, 17], [0, 14], [0, 15], [14, 16], [15, 17], [1, 2],
                  [2, 3], [3, 4], [1, 5], [5, 6], [6, 7], [1, 8], [8, 9], [9, 10], [1, 11], [11, 12],
                  [12, 13], [0, 2], [0, 5], [2, 8], [8, 12], [5, 11], [11, 9], [16, 2], [17, 5], [8, 11]])

# config.py:150-152
FLIP_HEAT_ORD = np.array([0, 1, 5, 6, 7, 2, 3, 4, 11, 12, 13, 8, 9, 10, 15, 14, 17, 16, 18, 19])
FLIP_PAF_ORD = np.array([0, 2, 1, 4, 3, 6, 5, 8, 7, 12, 13, 14, 9, 10, 11, 18, 19, 20, 15, 16, 17,
                         22, 21, 25, 26, 23, 24, 28, 27, 29])

ORDER_COCO = [0, 15, 14, 17, 16, 5, 2, 6, 3, 7, 4, 11, 8, 12, 9, 13, 10]  # evaluate.py:40


class CanonicalConfig:
    """The attributes of config.config.GetConfig('Canonical') that the path reads (parse_skeletons.py:41-42,
    83-88; evaluate.py:61)."""
    num_parts = NUM_PARTS
    paf_layers = NUM_LIMBS
    heat_layers = NUM_PARTS
    num_layers = NUM_LAYERS
    stride = 4
    width = 512
    height = 512
    limbs_conn = LIMBS
    flip_heat_ord = FLIP_HEAT_ORD
    flip_paf_ord = FLIP_PAF_ORD
    parts = PARTS + ["background", "reverseKeypoint"]


def GetConfig(name="Canonical"):
    if name != "Canonical":
        raise KeyError(name)
    return CanonicalConfig()


def config_reader():
    """-> (param, model) with the keys and values of utils/config_reader.py:47-78 reading utils/config."""
    param = dict(use_gpu=1, GPUdeviceNumber=0, modelID="1", starting_range=0.8, ending_range=2.0,
                 scale_search=[0.5, 1.0, 1.5, 2.0, 3.0], rotation_search=[0.0], thre1=0.1, thre2=0.1,
                 connect_ration=0.8, min_num=4, mid_num=20, len_rate=16, connection_tole=0.7, crop_ratio=2.5,
                 bbox_ratio=0.25, offset_radius=2, remove_recon=0)
    model = dict(boxsize=512, padValue=128, np=12, stride=4, max_downsample=64,
                 part_str=["nose", "neck", "Rsho", "Relb", "Rwri", "Lsho", "Lelb", "Lwri", "Rhip", "Rkne", "Rank",
                           "Lhip", "Lkne", "Lank", "Leye", "Reye", "Lear", "Rear", "pt19"])
    return param, model
