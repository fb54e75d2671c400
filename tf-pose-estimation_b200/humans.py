"""Result objects of the decode path: ``Human`` and ``BodyPart``.

Same public contract as the reference's result classes (tf_pose/estimator.py:43-265):
slots, attribute names, ``uidx`` format, ``__str__`` format, ``part_count`` /
``get_max_score`` / ``get_face_box`` / ``get_upper_body_box`` results -- so code written
against ``TfPoseEstimator.inference`` keeps working.  The implementation is this repo's own.
"""
import math

from .common import CocoPart

_FACE_MIN_SCORE = 0.2     # estimator.py:100
_BODY_MIN_SCORE = 0.3     # estimator.py:178
_UPPER_BODY_PARTS = (0, 1, 2, 5, 8, 11, 14, 15, 16, 17)   # estimator.py:181


def _iround(v):
    return int(round(v))


class BodyPart(object):
    """One detected key point; x / y are fractions of the map width / height."""
    __slots__ = ('uidx', 'part_idx', 'x', 'y', 'score')

    def __init__(self, uidx, part_idx, x, y, score):
        self.uidx = uidx
        self.part_idx = part_idx
        self.x, self.y = x, y
        self.score = score

    def get_part_name(self):
        return CocoPart(self.part_idx)

    def __str__(self):
        return 'BodyPart:%d-(%.2f, %.2f) score=%.2f' % (self.part_idx, self.x, self.y, self.score)

    __repr__ = __str__


class Human(object):
    """A skeleton: ``body_parts`` maps part index -> BodyPart (ascending insertion order)."""
    __slots__ = ('body_parts', 'pairs', 'uidx_list', 'score')

    def __init__(self, pairs):
        self.pairs = []
        self.uidx_list = set()
        self.body_parts = {}
        for pair in pairs:
            self.add_pair(pair)
        self.score = 0.0

    @staticmethod
    def _get_uidx(part_idx, idx):
        return '%d-%d' % (part_idx, idx)

    def add_pair(self, pair):
        # legacy (pre-pafprocess) construction path of the reference, estimator.py:61-71
        self.pairs.append(pair)
        for part_idx, idx, coord in ((pair.part_idx1, pair.idx1, pair.coord1),
                                     (pair.part_idx2, pair.idx2, pair.coord2)):
            uidx = Human._get_uidx(part_idx, idx)
            self.body_parts[part_idx] = BodyPart(uidx, part_idx, coord[0], coord[1], pair.score)
            self.uidx_list.add(uidx)

    def is_connected(self, other):
        return len(self.uidx_list & other.uidx_list) > 0

    def merge(self, other):
        for pair in other.pairs:
            self.add_pair(pair)

    def part_count(self):
        return len(self.body_parts)

    def get_max_score(self):
        return max(part.score for part in self.body_parts.values())

    # -- boxes ---------------------------------------------------------------------------
    def _confident(self, min_score):
        return {p.part_idx: p for p in self.body_parts.values() if p.score > min_score}

    def get_face_box(self, img_w, img_h, mode=0):
        """Face box in pixels of an img_w x img_h image, or None (estimator.py:84-159).
        mode 0 returns centre x/y, mode 1 returns the top-left corner and needs an eye."""
        found = self._confident(_FACE_MIN_SCORE)
        nose = found.get(CocoPart.Nose.value)
        if nose is None:
            return None
        neck = found.get(CocoPart.Neck.value)
        reye, leye = found.get(CocoPart.REye.value), found.get(CocoPart.LEye.value)
        rear, lear = found.get(CocoPart.REar.value), found.get(CocoPart.LEar.value)

        size = 0
        if neck is not None:
            size = max(size, img_h * (neck.y - nose.y) * 0.8)
        if reye is not None and leye is not None:
            size = max(size, img_w * (reye.x - leye.x) * 2.0)
            size = max(size, img_w * math.sqrt((reye.x - leye.x) ** 2 + (reye.y - leye.y) ** 2) * 2.0)
        if mode == 1 and reye is None and leye is None:
            return None
        if rear is not None and lear is not None:
            size = max(size, img_w * (rear.x - lear.x) * 1.6)
        if size <= 0:
            return None

        if reye is None and leye is not None:
            left = nose.x * img_w - (size // 3 * 2)
        elif reye is not None and leye is None:
            left = nose.x * img_w - (size // 3)
        else:
            left = nose.x * img_w - size // 2
        right = left + size
        top = nose.y * img_h - size // 3 if mode == 0 else nose.y * img_h - _iround(size / 2 * 1.2)
        bottom = top + size
        return _clip_box(left, top, right, bottom, img_w, img_h, centre=(mode == 0))

    def get_upper_body_box(self, img_w, img_h):
        """Upper-body box (centre x/y, w, h) in pixels, or None (estimator.py:161-235)."""
        if not (img_w > 0 and img_h > 0):
            raise Exception("img size should be positive")
        found = self._confident(_BODY_MIN_SCORE)
        pts = [(img_w * p.x, img_h * p.y) for p in found.values() if p.part_idx in _UPPER_BODY_PARTS]
        if len(pts) < 5:
            return None
        left, right = min(p[0] for p in pts), max(p[0] for p in pts)
        top, bottom = min(p[1] for p in pts), max(p[1] for p in pts)

        nose, neck = found.get(CocoPart.Nose.value), found.get(CocoPart.Neck.value)
        if nose is not None and neck is not None:
            top -= (neck.y * img_h - top) * 0.8
        rsho, lsho = found.get(CocoPart.RShoulder.value), found.get(CocoPart.LShoulder.value)
        if rsho is not None and lsho is not None:
            grow = (right - left) * 0.15
            left -= grow
            right += grow
        elif neck is not None and (rsho is None) != (lsho is None):
            one = lsho if lsho is not None else rsho
            half = abs(one.x - neck.x) * img_w * 1.15
            left = min(neck.x * img_w - half, left)
            right = max(neck.x * img_w + half, right)
        return _clip_box(left, top, right, bottom, img_w, img_h, centre=True)

    def __str__(self):
        return ' '.join(str(p) for p in self.body_parts.values())

    __repr__ = __str__


def _clip_box(left, top, right, bottom, img_w, img_h, centre):
    """Fit the box into the frame (estimator.py:139-158,223-235) and round."""
    left = max(0, left)
    top = max(0, top)
    right = min(img_w - left, right - left) + left
    bottom = min(img_h - top, bottom - top) + top
    if _iround(right - left) == 0.0 or _iround(bottom - top) == 0.0:
        return None
    if centre:
        return {"x": _iround((left + right) / 2), "y": _iround((top + bottom) / 2),
                "w": _iround(right - left), "h": _iround(bottom - top)}
    return {"x": _iround(left), "y": _iround(top), "w": _iround(right - left), "h": _iround(bottom - top)}
