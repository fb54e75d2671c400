"""Generates the committed golden vectors.  Run in the BUILD container (needs /root/reference,
scipy and oracle/_ref built from the reference sources); the vectors then travel with the repo.

  gauss_kernel_25_3.npy   the literal reference Smoother.gauss_kernel(25, 3.0, 1)
                          (tf_pose/tensblur/smoother.py:62-71, executed from the reference file with
                          a stub `tensorflow` module -- the function itself is pure numpy/scipy).
  toy_dense.npz           a small dense (peaks, heat, paf) triple with two hand-placed skeletons and
                          the humans the reference's own PoseEstimator.estimate_paf
                          (tf_pose/estimator.py:272-299, executed from the reference file) returns for
                          it on top of the reference's own compiled pafprocess.cpp (oracle/_ref).
  human_boxes.json        what the reference's Human helpers (part_count, get_max_score, get_face_box,
                          get_upper_body_box, __str__; tf_pose/estimator.py:73-241) return for random humans
                          built with the reference's own Human / BodyPart classes.
  synth_humans.json       for seeded synthetic frames: the humans the same reference code returns when
                          fed the oracle's dense maps, plus a sha256 of those maps so a test can tell
                          "inputs differ" from "decode differs".
"""
import hashlib
import importlib.util
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)

from oracle import oracle as O          # noqa: E402
import tfpose_b200                      # noqa: E402,F401
from tfpose_b200 import synth           # noqa: E402


class _Stub(types.ModuleType):
    """Stand-in for an absent third-party module (tensorflow, cv2, slidingwindow): any attribute is
    another stub, calling one returns a stub.  The decode code paths exercised here never touch it."""

    def __init__(self, name="stub"):
        types.ModuleType.__init__(self, name)

    def __getattr__(self, item):
        if item.startswith("__"):
            raise AttributeError(item)
        return _Stub(self.__name__ + "." + item)

    def __call__(self, *a, **k):
        return _Stub(self.__name__ + "()")


def load_reference_module(path, name, stubs):
    saved = {}
    for k, v in stubs.items():
        saved[k] = sys.modules.get(k)
        sys.modules[k] = v
    try:
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def reference_gauss_kernel():
    tf = _Stub("tensorflow")
    sm = load_reference_module(os.path.join(REF, "tf_pose/tensblur/smoother.py"), "ref_smoother", {"tensorflow": tf})
    k = sm.Smoother.gauss_kernel(None, 25, 3.0, 1)      # `self` is unused by the reference function
    return np.ascontiguousarray(k[:, :, 0, 0])


def reference_estimator(variant="asbuilt"):
    """The reference's estimator module with its compiled pafprocess bound through ctypes."""
    ref = O.ReferencePafprocess(variant)
    paf_mod = types.ModuleType("tf_pose.pafprocess.pafprocess")
    for name in ("process_paf", "get_num_humans", "get_part_cid", "get_score", "get_part_x", "get_part_y",
                 "get_part_score"):
        setattr(paf_mod, name, getattr(ref, name))
    pkg = types.ModuleType("tf_pose")
    pkg.__path__ = []
    pafpkg = types.ModuleType("tf_pose.pafprocess")
    pafpkg.pafprocess = paf_mod
    pafpkg.__path__ = []
    tf = _Stub("tensorflow")
    sw = _Stub("slidingwindow")
    cv2 = _Stub("cv2")
    common = load_reference_module(os.path.join(REF, "tf_pose/common.py"), "tf_pose.common",
                                   {"tensorflow": tf, "cv2": cv2})
    pkg.common = common
    tens = types.ModuleType("tf_pose.tensblur")
    tens.__path__ = []
    smo = types.ModuleType("tf_pose.tensblur.smoother")
    smo.Smoother = object
    stubs = {"tensorflow": tf, "slidingwindow": sw, "cv2": cv2, "tf_pose": pkg, "tf_pose.common": common,
             "tf_pose.pafprocess": pafpkg, "tf_pose.pafprocess.pafprocess": paf_mod, "tf_pose.tensblur": tens,
             "tf_pose.tensblur.smoother": smo}
    return load_reference_module(os.path.join(REF, "tf_pose/estimator.py"), "ref_estimator", stubs)


def humans_to_json(humans):
    out = []
    for h in humans:
        parts = {str(k): [bp.uidx, bp.x, bp.y, float(np.float32(bp.score))] for k, bp in h.body_parts.items()}
        out.append({"score": float(np.float32(h.score)), "parts": parts, "str": str(h)})
    return out


def toy_dense():
    """Two skeletons drawn directly at 'upsampled' resolution 48x40 (no TF ops involved)."""
    H, W = 48, 40
    heat = np.zeros((H, W, 19), np.float32)
    peaks = np.zeros((H, W, 19), np.float32)
    paf = np.zeros((H, W, 38), np.float32)
    rng = np.random.default_rng(7)
    from tfpose_b200.common import CocoPairs, CocoPairsNetwork
    people = [(10.0, 6.0, 30.0), (28.0, 8.0, 34.0)]
    tmpl = synth._TEMPLATE
    for (nx, ny, size) in people:
        joints = np.round(tmpl * size + np.array([nx, ny])).astype(int)
        joints[:, 0] = np.clip(joints[:, 0], 0, W - 1)
        joints[:, 1] = np.clip(joints[:, 1], 0, H - 1)
        for k in range(18):
            x, y = joints[k]
            v = np.float32(0.6 + 0.4 * rng.random())
            heat[y, x, k] = v
            peaks[y, x, k] = v
        for limb, (pa, pb) in enumerate(CocoPairs):
            ax, ay = joints[pa]
            bx, by = joints[pb]
            d = np.hypot(bx - ax, by - ay)
            if d < 1e-6:
                continue
            ux, uy = (bx - ax) / d, (by - ay) / d
            chx, chy = CocoPairsNetwork[limb]
            for t in np.linspace(0, 1, int(4 * d) + 2):
                px, py = int(round(ax + t * (bx - ax))), int(round(ay + t * (by - ay)))
                for dy in (-1, 0, 1):
                    for dx in (-1, 0, 1):
                        yy, xx = min(max(py + dy, 0), H - 1), min(max(px + dx, 0), W - 1)
                        paf[yy, xx, chx] = ux
                        paf[yy, xx, chy] = uy
    return peaks, heat, paf


def random_humans(est_mod, n=60, seed=11):
    """Humans with random subsets of parts, built with the REFERENCE's Human/BodyPart classes."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        hm = est_mod.Human([])
        keep = rng.random(18) < rng.uniform(0.3, 1.0)
        cx, cy, size = rng.uniform(0.2, 0.8), rng.uniform(0.1, 0.5), rng.uniform(0.2, 0.9)
        for k in range(18):
            if not keep[k]:
                continue
            px = float(np.clip(cx + synth._TEMPLATE[k, 0] * size + rng.normal(0, 0.01), 0, 1))
            py = float(np.clip(cy + synth._TEMPLATE[k, 1] * size + rng.normal(0, 0.01), 0, 1))
            hm.body_parts[k] = est_mod.BodyPart('%d-%d' % (i, k), k, px, py, float(np.float32(rng.uniform(0.05, 1.0))))
        hm.score = float(rng.uniform(0.3, 2.0))
        out.append(hm)
    return out


def boxes_golden(est_mod):
    """Results of the reference's Human helpers (estimator.py:73-241) for random humans."""
    recs = []
    for hm in random_humans(est_mod):
        if not hm.body_parts:
            continue
        rec = {"parts": {str(k): [bp.uidx, bp.x, bp.y, bp.score] for k, bp in hm.body_parts.items()},
               "score": hm.score, "part_count": hm.part_count(), "max_score": hm.get_max_score(), "str": str(hm),
               "names": [str(bp.get_part_name()) for bp in hm.body_parts.values()], "boxes": []}
        for (iw, ih) in ((432, 368), (656, 368), (1312, 736), (100, 37)):
            rec["boxes"].append({"w": iw, "h": ih, "face0": hm.get_face_box(iw, ih, mode=0),
                                 "face1": hm.get_face_box(iw, ih, mode=1), "upper": hm.get_upper_body_box(iw, ih)})
        recs.append(rec)
    with open(os.path.join(HERE, "human_boxes.json"), "w") as fh:
        json.dump(recs, fh)
    print("human boxes:", len(recs), "humans;",
          sum(1 for r in recs for b in r["boxes"] if b["face0"]), "face boxes,",
          sum(1 for r in recs for b in r["boxes"] if b["upper"]), "upper-body boxes")


def coco_golden(est_mod):
    """Results of the reference's own output writers -- write_coco_json (tf_pose/eval.py:35-44) and
    translate_op_top_coco (ariel_run.py:22-28) with the keypoint list of ariel_run.py:73-80 -- for the
    same random humans, everything else those scripts import stubbed."""
    names = ["common", "estimator", "networks", "pycocotools", "pycocotools.coco", "pycocotools.cocoeval", "tqdm",
             "cv2", "tf_pose", "tf_pose.common", "tf_pose.estimator", "tf_pose.networks"]
    stubs = {n: _Stub(n) for n in names}
    ev = load_reference_module(os.path.join(REF, "tf_pose/eval.py"), "ref_eval", stubs)
    ar = load_reference_module(os.path.join(REF, "ariel_run.py"), "ref_ariel_run", stubs)
    recs = []
    for hm in random_humans(est_mod, n=24, seed=23):
        rec = {"parts": {str(k): [bp.x, bp.y, bp.score] for k, bp in hm.body_parts.items()}, "score": hm.score,
               "sizes": []}
        for (iw, ih) in ((432, 368), (1312, 736), (101, 37)):
            kps = []
            for i in range(18):        # ariel_run.py:76-80, verbatim semantics
                if i in hm.body_parts.keys():
                    kps += [hm.body_parts[i].x * iw, hm.body_parts[i].y * ih, hm.body_parts[i].score]
                else:
                    kps += [0, 0, 0]
            rec["sizes"].append({"w": iw, "h": ih, "write_coco_json": ev.write_coco_json(hm, iw, ih),
                                 "translate_op_top_coco": ar.translate_op_top_coco(kps)})
        recs.append(rec)
    with open(os.path.join(HERE, "coco_writers.json"), "w") as fh:
        json.dump(recs, fh)
    print("coco writers:", len(recs), "humans x 3 image sizes")


class _Msg:
    """Stand-in for a genpy message class: attributes are whatever the caller sets, list fields start empty."""
    _lists = ()

    def __init__(self):
        for n in self._lists:
            setattr(self, n, [])


def ros_golden(est_mod):
    """Result of the reference's own humans_to_msg (scripts/broadcaster_ros.py:19-36) for random humans, with
    rospy / cv_bridge / the generated tfpose_ros.msg classes stubbed (plain attribute bags)."""
    class Persons(_Msg):
        _lists = ("persons",)

    class Person(_Msg):
        _lists = ("body_part",)

    class BodyPartElm(_Msg):
        pass

    names = ["rospy", "rospkg", "cv_bridge", "std_msgs", "std_msgs.msg", "sensor_msgs", "sensor_msgs.msg",
             "tfpose_ros", "tfpose_ros.msg", "tf_pose", "tf_pose.estimator", "tf_pose.networks"]
    stubs = {n: _Stub(n) for n in names}
    stubs["tfpose_ros.msg"].Persons, stubs["tfpose_ros.msg"].Person = Persons, Person
    stubs["tfpose_ros.msg"].BodyPartElm = BodyPartElm
    ros = load_reference_module(os.path.join(REF, "scripts/broadcaster_ros.py"), "ref_broadcaster_ros", stubs)
    humans = random_humans(est_mod, n=16, seed=31)
    msg = ros.humans_to_msg(humans)
    rec = {"humans": [{"parts": {str(k): [bp.x, bp.y, bp.score] for k, bp in hm.body_parts.items()},
                       "score": hm.score} for hm in humans],
           "persons": [[[b.part_id, b.x, b.y, b.confidence] for b in p.body_part] for p in msg.persons]}
    with open(os.path.join(HERE, "ros_persons.json"), "w") as fh:
        json.dump(rec, fh)
    print("ros persons:", len(rec["persons"]), "persons,", sum(len(p) for p in rec["persons"]), "body parts")


PREPROC_CASES = [  # (seed, src_h, src_w, cn, dst_w, dst_h)
    (0, 480, 640, 3, 432, 368), (1, 720, 1280, 3, 432, 368), (2, 368, 432, 3, 656, 368), (3, 100, 75, 3, 432, 368),
    (4, 333, 517, 3, 431, 367), (5, 50, 61, 1, 43, 37), (6, 64, 64, 4, 33, 129), (7, 7, 9, 3, 20, 20),
    (8, 2, 2, 1, 9, 9), (9, 1, 5, 3, 4, 3), (10, 200, 300, 2, 101, 55)]


def preproc_image(seed, sh, sw, cn):
    return np.random.default_rng(9000 + seed).integers(0, 256, (sh, sw, cn), dtype=np.uint8)


def preproc_golden():
    """cv2.resize(INTER_CUBIC) itself (the call of estimator.py:423), IPP off = OpenCV's own code.  Stores the
    sha256 of every output and one small output in full."""
    import cv2
    cv2.ipp.setUseIPP(False)
    shas = []
    for (seed, sh, sw, cn, dw, dh) in PREPROC_CASES:
        out = cv2.resize(preproc_image(seed, sh, sw, cn), (dw, dh), interpolation=cv2.INTER_CUBIC)
        shas.append(hashlib.sha256(np.ascontiguousarray(out).tobytes()).hexdigest())
    small = cv2.resize(preproc_image(7, 7, 9, 3), (20, 20), interpolation=cv2.INTER_CUBIC)
    np.savez_compressed(os.path.join(HERE, "preproc_cubic.npz"), cases=np.array(PREPROC_CASES, np.int64),
                        sha256=np.array(shas), small=small, cv2_version=np.array(cv2.__version__))
    print("preproc cubic", len(shas), "cases, cv2", cv2.__version__)


SCALED_CASES = [  # (seed, img_h, img_w, target (w, h), scale)
    (20, 300, 400, (216, 184), None), (21, 184, 216, (216, 184), None), (22, 300, 400, (216, 184), 1.0),
    (23, 300, 400, (216, 184), 1.5), (24, 200, 150, (216, 184), 0.7), (25, 300, 400, (216, 184), (1.2, 0.3)),
    (26, 300, 400, (216, 184), (0.5, 0.5, 1.5)), (27, 300, 400, (216, 184), (0.05, 0.95, 2.0)),
    (28, 300, 400, (216, 184), (0.5, 0.5, 0.8))]


def reference_slidingwindow():
    """The reference's VENDORED copy of the third-party ``slidingwindow`` package
    (tf_pose/slidingwindow/SlidingWindow.py:87-110 ``generate`` / ``generateForSize``), imported from where it lies
    under the name estimator.py:4 imports it by."""
    pkg_dir = os.path.join(REF, "tf_pose", "slidingwindow")
    spec = importlib.util.spec_from_file_location("slidingwindow", os.path.join(pkg_dir, "__init__.py"),
                                                  submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    saved = sys.modules.get("slidingwindow")
    sys.modules["slidingwindow"] = mod
    try:
        spec.loader.exec_module(mod)
    finally:
        if saved is None:
            sys.modules.pop("slidingwindow", None)
        else:
            sys.modules["slidingwindow"] = saved
    return mod


SLIDING_CASES = [  # (width, height, win_w, win_h, overlap): data smaller / equal / larger than the window, exact and
    (216, 184, 216, 184, 0.2), (400, 300, 216, 184, 0.2), (600, 450, 216, 184, 0.2), (600, 450, 216, 184, 0.3),   # ragged steps
    (150, 100, 216, 184, 0.2), (432, 368, 216, 184, 0.5), (433, 369, 216, 184, 0.5), (1000, 184, 216, 184, 0.0),
    (217, 185, 216, 184, 0.9), (656, 368, 432, 368, 0.2), (1312, 736, 432, 368, 0.25), (5, 5, 3, 2, 0.34)]


def scaled_img_golden():
    """The reference's OWN ``TfPoseEstimator._get_scaled_img`` / ``_crop_roi`` (estimator.py:417-516) run on real
    cv2 (IPP off) over the reference's own vendored ``slidingwindow`` package: every piece of the scaled forms,
    window enumeration included, is what the reference computes.  Also stores the bare window lists of
    ``slidingwindow.generateForSize`` for a set of sizes (``sliding_windows``)."""
    import cv2
    cv2.ipp.setUseIPP(False)
    est = reference_estimator("asbuilt")
    est.cv2 = cv2
    sw = reference_slidingwindow()
    est.sw = sw
    out = []
    for (seed, ih, iw, target, scale) in SCALED_CASES:
        me = types.SimpleNamespace(target_size=target)
        me._crop_roi = lambda img, rx, ry, me=me: est.TfPoseEstimator._crop_roi(me, img, rx, ry)
        rois, ratios = est.TfPoseEstimator._get_scaled_img(me, preproc_image(seed, ih, iw, 3), scale)
        out.append({"seed": seed, "h": ih, "w": iw, "target": list(target),
                    "scale": list(scale) if isinstance(scale, tuple) else scale,
                    "rois": [None if r is None else [list(r.shape), hashlib.sha256(np.ascontiguousarray(r).tobytes()).hexdigest()]
                             for r in rois],
                    "ratios": [list(map(float, r)) for r in ratios]})
        print("scaled", scale, len(rois), [None if r is None else r.shape for r in rois][:3])
    wins = []
    for (width, height, ww, wh, ov) in SLIDING_CASES:
        got = sw.generateForSize(width, height, sw.DimOrder.HeightWidthChannel, ww, wh, ov)
        wins.append({"width": width, "height": height, "win_w": ww, "win_h": wh, "overlap": ov,
                     "windows": [list(map(int, w_.getRect())) for w_ in got]})
        print("windows", (width, height, ww, wh, ov), len(got))
    with open(os.path.join(HERE, "scaled_img.json"), "w") as fh:
        json.dump({"cases": out, "sliding_windows": wins,
                   "source": "reference estimator.py:417-516 over the vendored tf_pose/slidingwindow package, cv2 %s IPP off" % cv2.__version__}, fh)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "scaled":
        scaled_img_golden()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "preproc":
        preproc_golden()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "boxes":
        boxes_golden(reference_estimator("asbuilt"))
        return
    if len(sys.argv) > 1 and sys.argv[1] == "ros":
        ros_golden(reference_estimator("asbuilt"))
        return
    if len(sys.argv) > 1 and sys.argv[1] == "coco":
        coco_golden(reference_estimator("asbuilt"))
        return
    k = reference_gauss_kernel()
    np.save(os.path.join(HERE, "gauss_kernel_25_3.npy"), k)
    print("gauss kernel", k.shape, k.dtype, k[12, 12], k[0, 0])

    est = {v: reference_estimator(v) for v in ("asbuilt", "upstream")}
    peaks, heat, paf = toy_dense()
    toy = {}
    for v in est:
        toy[v] = humans_to_json(est[v].PoseEstimator.estimate_paf(peaks, heat, paf))
    np.savez_compressed(os.path.join(HERE, "toy_dense.npz"), peaks=peaks, heat=heat, paf=paf)
    print("toy humans", {v: len(t) for v, t in toy.items()})

    cases = []
    for (cfg, frame, h, w, people) in [(2, 0, 46, 54, None), (2, 1, 46, 54, None), (2, 2, 46, 54, None),
                                       (2, 5, 46, 54, None), (3, 0, 46, 82, 5), (4, 0, 46, 54, 30),
                                       (4, 1, 46, 54, 30), (5, 0, 92, 164, 12)]:
        x = synth.synth_frame(h, w, synth.frame_seed(cfg, frame), people=people)
        H, W = 4 * h, 4 * w
        pk, hu, pu, g = O.dense_maps(x, H, W)
        sha = hashlib.sha256(pk.tobytes() + hu.tobytes() + pu.tobytes()).hexdigest()
        rec = {"config": cfg, "frame": frame, "h": h, "w": w, "people": people, "maps_sha256": sha,
               "x_sha256": hashlib.sha256(x.tobytes()).hexdigest(), "fragile_pixels": int(O.fragile_pixels(g).sum())}
        for v in est:
            rec[v] = humans_to_json(est[v].PoseEstimator.estimate_paf(pk, hu, pu))
        cases.append(rec)
        print(cfg, frame, h, w, people, {v: len(rec[v]) for v in est}, "fragile", rec["fragile_pixels"])
    with open(os.path.join(HERE, "synth_humans.json"), "w") as fh:
        json.dump({"toy": toy, "cases": cases}, fh)
    boxes_golden(est["asbuilt"])
    coco_golden(est["asbuilt"])


if __name__ == "__main__":
    main()
