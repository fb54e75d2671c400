"""Output formats next to the decode path: COCO keypoint records.

* ``write_coco_json(human, image_w, image_h)`` -- same name, arguments and result as
  tf_pose/eval.py:35-44: 17 COCO keypoints as ``[x, y, 2]`` integer triples (``[0, 0, 0]`` for a missing
  part) in COCO order, Python ``round`` (banker's rounding) like the reference's ``round_int`` (:31-32).
* ``coco_result_item`` -- the dict eval.py:109-114 appends per human.
* ``openpose_keypoints`` / ``translate_op_to_coco`` / ``ariel_result_item`` -- the 18x3 float list and its
  17-keypoint permutation of ariel_run.py:22-28,73-86 (``translate_op_top_coco`` there).
* ``coco_items_from_table`` -- the same records built straight from the ``tfpd_human`` tables a decode
  returns (no ``Human`` objects in between), for whole batches.
* ``humans_to_msg`` -- scripts/broadcaster_ros.py:19-36: ``Persons`` / ``Person`` / ``BodyPartElm`` objects
  (msg/*.msg).  rospy's generated classes are passed in when ROS is there; plain stand-ins with the same
  fields otherwise.  ``persons_records_from_table`` gives the same content as one packed array per frame
  (``BODY_PART_ELM_DTYPE`` = the 16-byte little-endian wire layout of ``BodyPartElm``).
"""
import numpy as np

# OpenPose part index of each COCO keypoint (eval.py:37, ariel_run.py:23)
COCO_IDS = [0, 15, 14, 17, 16, 5, 2, 6, 3, 7, 4, 11, 8, 12, 9, 13, 10]


def round_int(val):
    return int(round(val))


def write_coco_json(human, image_w, image_h):
    keypoints = []
    for coco_id in COCO_IDS:
        if coco_id not in human.body_parts.keys():
            keypoints.extend([0, 0, 0])
            continue
        body_part = human.body_parts[coco_id]
        keypoints.extend([round_int(body_part.x * image_w), round_int(body_part.y * image_h), 2])
    return keypoints


def coco_result_item(human, image_id, image_w, image_h):
    """eval.py:109-114."""
    return {'image_id': image_id, 'category_id': 1, 'keypoints': write_coco_json(human, image_w, image_h),
            'score': human.score}


def openpose_keypoints(human, image_w, image_h):
    """ariel_run.py:76-80: 18 x (x*w, y*h, score) floats, zeros for a missing part."""
    kps = []
    for i in range(18):
        if i in human.body_parts.keys():
            kps += [human.body_parts[i].x * image_w, human.body_parts[i].y * image_h, human.body_parts[i].score]
        else:
            kps += [0, 0, 0]
    return kps


def translate_op_to_coco(kps):
    """ariel_run.py:22-28 (``translate_op_top_coco``)."""
    assert len(kps) == 54
    result = []
    for i in COCO_IDS:
        result += kps[3 * i:3 * i + 3]
    return result


def ariel_result_item(human, filename, image_id, image_w, image_h):
    """ariel_run.py:73-85."""
    return {"filename": filename, 'category_id': 1, 'image_id': image_id,
            'keypoints': translate_op_to_coco(openpose_keypoints(human, image_w, image_h))}


def coco_items_from_table(result, image_ids, image_sizes):
    """COCO result items of a whole ``DecodeResult`` without building ``Human`` objects.

    image_ids[i], image_sizes[i] = (w, h) of frame i.  Equal, item for item, to
    ``[coco_result_item(h, ...) for h in result.humans(i)]`` (humans with no part are skipped like
    estimate_paf's ``is_added``, estimator.py:293)."""
    items = []
    H, W = result.H, result.W
    for i in range(len(result)):
        n = int(result.info["n_humans"][i])
        iw, ih = image_sizes[i]
        rows = result.table[i][:n]
        for r in rows:
            cid = r["cid"]
            if not (cid >= 0).any():
                continue
            kp = []
            for p in COCO_IDS:
                if cid[p] < 0:
                    kp.extend([0, 0, 0])
                else:
                    kp.extend([round_int(float(int(r["x"][p])) / W * iw), round_int(float(int(r["y"][p])) / H * ih), 2])
            items.append({'image_id': image_ids[i], 'category_id': 1, 'keypoints': kp, 'score': float(r["score"])})
    return items


def table_to_keypoint_array(result, frame):
    """[n_humans, 18, 3] float32 (x_px, y_px, score) on the upsampled map, NaN-free: missing parts are 0."""
    n = int(result.info["n_humans"][frame])
    out = np.zeros((n, 18, 3), np.float32)
    rows = result.table[frame][:n]
    ok = rows["cid"] >= 0
    out[..., 0] = np.where(ok, rows["x"], 0)
    out[..., 1] = np.where(ok, rows["y"], 0)
    out[..., 2] = np.where(ok, rows["part_score"], 0)
    return out


class _Msg:
    """Field bag shaped like a genpy message (used when the ROS message classes are not importable)."""
    __slots__ = ()

    def __repr__(self):
        return "%s(%s)" % (type(self).__name__, ", ".join("%s=%r" % (n, getattr(self, n)) for n in self.__slots__))


class BodyPartElm(_Msg):                 # msg/BodyPartElm.msg
    __slots__ = ("part_id", "x", "y", "confidence")

    def __init__(self):
        self.part_id, self.x, self.y, self.confidence = 0, 0.0, 0.0, 0.0


class Person(_Msg):                      # msg/Person.msg
    __slots__ = ("body_part",)

    def __init__(self):
        self.body_part = []


class Persons(_Msg):                     # msg/Persons.msg
    __slots__ = ("persons", "image_w", "image_h", "header")

    def __init__(self):
        self.persons, self.image_w, self.image_h, self.header = [], 0, 0, None


def humans_to_msg(humans, msg_types=None):
    """scripts/broadcaster_ros.py:19-36.  msg_types = (Persons, Person, BodyPartElm) classes of
    ``tfpose_ros.msg`` when running under ROS; the caller sets image_w / image_h / header (:57-59)."""
    P, Q, B = msg_types if msg_types is not None else (Persons, Person, BodyPartElm)
    persons = P()
    for human in humans:
        person = Q()
        for k in human.body_parts:
            body_part = human.body_parts[k]
            m = B()
            m.part_id = body_part.part_idx
            m.x = body_part.x
            m.y = body_part.y
            m.confidence = body_part.score
            person.body_part.append(m)
        persons.persons.append(person)
    return persons


# uint32 part_id, float32 x, float32 y, float32 confidence: the serialised form of one BodyPartElm
BODY_PART_ELM_DTYPE = np.dtype([("part_id", "<u4"), ("x", "<f4"), ("y", "<f4"), ("confidence", "<f4")])


def persons_records_from_table(result, frame):
    """(records, counts) of one frame straight from the ``tfpd_human`` table: ``records`` is a packed
    BODY_PART_ELM_DTYPE array with the body parts of all persons in ``humans_to_msg`` order (person by
    person, ascending part index), ``counts[i]`` the number of parts of person i.  x, y are the float32
    roundings of the Python floats ``px / W``, ``py / H`` the message fields get when serialised."""
    n = int(result.info["n_humans"][frame])
    rows = result.table[frame][:n]
    ok = rows["cid"] >= 0
    keep = ok.any(axis=1)                                  # estimate_paf's is_added (estimator.py:293)
    rows, ok = rows[keep], ok[keep]
    counts = ok.sum(axis=1).astype(np.int64)
    rec = np.empty(int(counts.sum()), BODY_PART_ELM_DTYPE)
    hi, pi = np.nonzero(ok)
    rec["part_id"] = pi
    rec["x"] = (rows["x"][hi, pi].astype(np.float64) / result.W).astype(np.float32)
    rec["y"] = (rows["y"][hi, pi].astype(np.float64) / result.H).astype(np.float32)
    rec["confidence"] = rows["part_score"][hi, pi]
    return rec, counts
