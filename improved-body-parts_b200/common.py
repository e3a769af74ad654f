"""Result containers with the attribute names the reference's callers use (utils/common.py:17-51, 267-288):
`Human.body_parts[part_idx] = BodyPart(uidx, part_idx, x, y, score)`, `human.score`.  Drawing helpers are out of
scope."""
from enum import Enum


class CocoPart(Enum):  # utils/common.py:17-36
    Nose = 0
    Neck = 1
    RShoulder = 2
    RElbow = 3
    RWrist = 4
    LShoulder = 5
    LElbow = 6
    LWrist = 7
    RHip = 8
    RKnee = 9
    RAnkle = 10
    LHip = 11
    LKnee = 12
    LAnkle = 13
    REye = 14
    LEye = 15
    REar = 16
    LEar = 17
    Background = 18


class BodyPart:  # utils/common.py:267-288
    __slots__ = ("uidx", "part_idx", "x", "y", "score")

    def __init__(self, uidx, part_idx, x, y, score):
        self.uidx = uidx
        self.part_idx = part_idx
        self.x, self.y = x, y
        self.score = score

    def get_part_name(self):
        return CocoPart(self.part_idx)

    def __repr__(self):
        return "BodyPart:%d-(%.2f, %.2f) score=%.2f" % (self.part_idx, self.x, self.y, self.score)

    __str__ = __repr__


class Human:  # utils/common.py:39-51
    __slots__ = ("body_parts", "pairs", "uidx_list", "score")

    def __init__(self, pairs=()):
        self.pairs = list(pairs)
        self.uidx_list = set()
        self.body_parts = {}
        self.score = 0.0

    def part_count(self):
        return len(self.body_parts)

    def get_max_score(self):
        return max(p.score for p in self.body_parts.values())

    def __repr__(self):
        return " ".join(str(p) for p in self.body_parts.values())


def humans_from_result(peaks, humans):
    """evaluate.py:111-129 for one image of a BatchResult: structured arrays -> [Human]."""
    out = []
    for hid, rec in enumerate(humans):
        human = Human()
        for part_idx in range(18):
            pid = int(rec["peak_id"][part_idx])
            if pid < 0:
                continue
            pk = peaks[pid]
            human.body_parts[part_idx] = BodyPart("%d-%d" % (hid, part_idx), part_idx, int(pk["x"]), int(pk["y"]),
                                                  float(pk["score"]))
        if human.body_parts:
            human.score = float(rec["score"])
            out.append(human)
    return out
