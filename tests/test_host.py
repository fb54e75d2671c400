"""CPU tests of the host side: the C-ABI library loads and exports every symbol the header
declares (no compute calls: there is no GPU here), the result objects honour the reference's
contract (checked against golden vectors made by the reference's own classes), table -> Human
conversion, and the by-image sharding helpers over a 2-rank gloo group."""
import ctypes as C
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import tfpose_b200 as tp
from tfpose_b200 import _lib, shard

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


# ------------------------------------------------------------------------------------ C-ABI
def header_symbols():
    src = open(os.path.join(ROOT, "include", "tfpose_decode.h")).read()
    return re.findall(r"TFPD_API\s+[\w\s\*]+?\b(tfpd_\w+)\s*\(", src)


def test_library_exports_every_declared_symbol():
    names = header_symbols()
    assert len(names) >= 22 and len(set(names)) == len(names)
    L = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), "libtfpose_decode.so does not export %s" % n
    assert sorted(names) == sorted(_lib.SIGNATURES), "ctypes prototypes out of step with the header"
    # nothing but the C-ABI is visible
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = [l.split()[-1] for l in out.splitlines() if " T " in l]
    assert sorted(e for e in exported if not e.startswith("_")) == sorted(names)


def test_struct_layouts_match_the_header():
    assert C.sizeof(_lib.Config) == 16 * 4
    assert C.sizeof(_lib.HumanRec) == 4 + 3 * 18 * 2 + 18 * 4 and _lib.HUMAN_DTYPE.itemsize == C.sizeof(_lib.HumanRec)
    assert C.sizeof(_lib.FrameInfo) == 24
    cfg = _lib.default_config()
    # decode constants of pafprocess.h:6-13 and estimator.py:332
    assert (round(cfg.thresh_heat, 6), round(cfg.thresh_vector_score, 6), cfg.thresh_vector_cnt1, cfg.thresh_part_cnt,
            round(cfg.thresh_human_score, 6), cfg.blur_ksize, cfg.blur_sigma) == (0.05, 0.05, 6, 4, 0.3, 25, 3.0)
    with pytest.raises(TypeError):
        _lib.default_config(no_such_field=1)


def test_smoother_weights_are_the_reference_kernel():
    """tfpd_gauss_kernel is host code: the 25x25 float32 weights must be bit-identical to what the
    reference's Smoother.gauss_kernel produced (golden made by executing smoother.py:62-71)."""
    gold = np.load(os.path.join(GOLD, "gauss_kernel_25_3.npy"))
    k = _lib.gauss_kernel(25, 3.0)
    assert np.array_equal(k, gold)
    k7 = _lib.gauss_kernel(7, 1.5)
    assert abs(float(k7.astype(np.float64).sum()) - 1.0) < 1e-6 and np.allclose(k7, k7.T)
    with pytest.raises(tp.DecodeError):
        _lib.gauss_kernel(24, 3.0)


def test_no_cpu_fallback_without_a_device():
    """On a box without CUDA the product must fail loudly, not compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    with pytest.raises(tp.DecodeError) as ei:
        tp.PoseDecoder(max_batch=1)
    assert ei.value.code == _lib.ERR_CUDA and "no CPU path" in str(ei.value)
    from tfpose_b200 import pafprocess
    with pytest.raises(tp.DecodeError):
        pafprocess.process_paf(np.zeros((4, 4, 19), np.float32), np.zeros((4, 4, 19), np.float32),
                               np.zeros((4, 4, 38), np.float32))


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "tf-pose-estimation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert "liboracle" not in src and "/root/reference" not in src, f


def test_missing_library_raises(tmp_path, monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "libtfpose_decode.so"))
    with pytest.raises(tp.DecodeLibraryMissing):
        _lib.lib()


# ------------------------------------------------------------------------------------ result objects
def _human_from_json(rec):
    hm = tp.Human([])
    for k, (uidx, x, y, s) in rec["parts"].items():
        hm.body_parts[int(k)] = tp.BodyPart(uidx, int(k), x, y, s)
    hm.score = rec["score"]
    return hm


def test_human_helpers_match_reference_golden():
    recs = json.load(open(os.path.join(GOLD, "human_boxes.json")))
    n_face = n_upper = 0
    for rec in recs:
        hm = _human_from_json(rec)
        assert hm.part_count() == rec["part_count"] and hm.get_max_score() == rec["max_score"]
        assert str(hm) == rec["str"]
        assert [str(bp.get_part_name()) for bp in hm.body_parts.values()] == rec["names"]
        for b in rec["boxes"]:
            assert hm.get_face_box(b["w"], b["h"], mode=0) == b["face0"]
            assert hm.get_face_box(b["w"], b["h"], mode=1) == b["face1"]
            assert hm.get_upper_body_box(b["w"], b["h"]) == b["upper"]
            n_face += b["face0"] is not None
            n_upper += b["upper"] is not None
    assert n_face > 20 and n_upper > 20


def test_result_object_contract():
    assert tp.Human.__slots__ == ('body_parts', 'pairs', 'uidx_list', 'score')
    assert tp.BodyPart.__slots__ == ('uidx', 'part_idx', 'x', 'y', 'score')
    bp = tp.BodyPart('3-7', 7, 0.25, 0.5, 0.875)
    assert str(bp) == 'BodyPart:7-(0.25, 0.50) score=0.88' and repr(bp) == str(bp)
    assert bp.get_part_name() == tp.CocoPart.LWrist
    hm = tp.Human([])
    assert hm.pairs == [] and hm.uidx_list == set() and hm.body_parts == {} and hm.score == 0.0
    with pytest.raises(Exception):
        hm.get_upper_body_box(0, 10)
    assert len(tp.CocoPairs) == 19 and len(tp.CocoPairsRender) == 17 and len(tp.CocoPairsNetwork) == 19
    assert [p.value for p in tp.CocoPart] == list(range(19))


def test_humans_from_table_follows_estimate_paf():
    """Golden humans -> tfpd_human records -> Human objects must reproduce the reference's objects
    (uidx '%d-%d' % (human_id, part), x = px / W as Python floats, ascending part order)."""
    gold = json.load(open(os.path.join(GOLD, "synth_humans.json")))
    case = gold["cases"][0]
    H, W = 4 * case["h"], 4 * case["w"]
    want = case["asbuilt"]
    table = np.zeros(8, _lib.HUMAN_DTYPE)
    table["cid"] = -1
    for i, hm in enumerate(want):
        table[i]["score"] = hm["score"]
        for k, (uidx, x, y, s) in hm["parts"].items():
            k = int(k)
            table[i]["cid"][k] = k
            table[i]["x"][k] = int(round(x * W))
            table[i]["y"][k] = int(round(y * H))
            table[i]["part_score"][k] = s
    humans = tp.humans_from_table(table, len(want), H, W)
    assert len(humans) == len(want)
    for hm, w_ in zip(humans, want):
        assert str(hm) == w_["str"]
        assert {str(k): [bp.uidx, bp.x, bp.y, bp.score] for k, bp in hm.body_parts.items()} == w_["parts"]
        assert hm.score == w_["score"]


def test_upsample_size_rule():
    # estimator.py:522-523 with the CLI defaults (run.py:30: 4.0 ; eval.py:50: 8.0)
    assert tp.upsample_size_for(432, 368, 4.0) == [184, 216]
    assert tp.upsample_size_for(656, 368, 4.0) == [184, 328]
    assert tp.upsample_size_for(1312, 736, 4.0) == [368, 656]
    assert tp.upsample_size_for(432, 368, 8.0) == [368, 432]
    assert tp.upsample_size_for(432, 368, 2.5) == [115, 135]


# ------------------------------------------------------------------------------------ sharding
def test_shard_bounds_cover_the_batch():
    for n in (0, 1, 7, 64, 128, 512, 513):
        for world in (1, 2, 3, 4, 8):
            spans = [shard.shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    assert shard.shard_bounds(128, 1, 2) == (64, 128) and shard.shard_bounds(512, 7, 8) == (448, 512)
    with pytest.raises(ValueError):
        shard.shard_bounds(8, 2, 2)


def test_pack_unpack_roundtrip():
    rng = np.random.default_rng(3)
    counts = np.array([2, 0, 3], np.int32)
    recs = np.zeros(5, _lib.HUMAN_DTYPE)
    recs["score"] = rng.random(5).astype(np.float32)
    recs["cid"] = rng.integers(-1, 50, (5, 18))
    c2, r2 = shard.unpack_result(shard.pack_result(counts, recs))
    assert np.array_equal(c2, counts) and r2.tobytes() == recs.tobytes()
    parts = shard.split_records(c2, r2)
    assert [len(p) for p in parts] == [2, 0, 3]
    c0, r0 = shard.unpack_result(shard.pack_result(np.zeros(0, np.int32), np.zeros(0, _lib.HUMAN_DTYPE)))
    assert len(c0) == 0 and len(r0) == 0


_GLOO_WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r)
import numpy as np, torch, torch.distributed as dist
import tfpose_b200
from tfpose_b200 import shard, _lib
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%(port)d", rank=rank, world_size=world)
N = 13
lo, hi = shard.shard_bounds(N, rank, world)
# frame f "decodes" to (f %% 3) humans whose score encodes (frame, human)
counts = np.array([f %% 3 for f in range(lo, hi)], np.int32)
recs = np.zeros(int(counts.sum()), _lib.HUMAN_DTYPE)
k = 0
for f in range(lo, hi):
    for j in range(f %% 3):
        recs[k]["score"] = f + j / 8.0
        recs[k]["cid"][:] = f
        k += 1
out = shard.gather_results(counts, recs, dst=0)
if rank == 0:
    c, r = out
    assert c.tolist() == [f %% 3 for f in range(N)], c
    per = shard.split_records(c, r)
    for f in range(N):
        assert [float(x) for x in per[f]["score"]] == [f + j / 8.0 for j in range(f %% 3)]
        assert all((row["cid"] == f).all() for row in per[f])
    print("GATHER_OK")
else:
    assert out is None
dist.barrier()
dist.destroy_process_group()
'''


@pytest.mark.parametrize("world", [2, 3])
def test_gather_results_over_gloo(world, tmp_path):
    """The N>1 path on CPU: every rank 'decodes' its contiguous shard, rank 0 receives the tables in
    frame order (the only communication the multi-GPU path has)."""
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER % {"root": ROOT, "port": port})
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=180)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert "GATHER_OK" in outs[0]


def test_coco_writers_match_reference_golden():
    """tests/golden/coco_writers.json holds what the reference's own write_coco_json (tf_pose/eval.py:35-44)
    and translate_op_top_coco (ariel_run.py:22-28) returned for random humans (make_golden.py coco)."""
    import json
    from tfpose_b200 import writers
    recs = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "coco_writers.json")))
    assert len(recs) >= 20
    for i, rec in enumerate(recs):
        hm = tp.Human([])
        for k, (x, y, s) in rec["parts"].items():
            hm.body_parts[int(k)] = tp.BodyPart('%d-%s' % (i, k), int(k), x, y, s)
        hm.score = rec["score"]
        for sz in rec["sizes"]:
            assert writers.write_coco_json(hm, sz["w"], sz["h"]) == sz["write_coco_json"]
            kps = writers.openpose_keypoints(hm, sz["w"], sz["h"])
            assert writers.translate_op_to_coco(kps) == sz["translate_op_top_coco"]
            item = writers.coco_result_item(hm, 7, sz["w"], sz["h"])
            assert item == {"image_id": 7, "category_id": 1, "keypoints": sz["write_coco_json"], "score": rec["score"]}
            a = writers.ariel_result_item(hm, "a.jpg", 3, sz["w"], sz["h"])
            assert a["keypoints"] == sz["translate_op_top_coco"] and a["filename"] == "a.jpg" and a["image_id"] == 3


def test_coco_items_from_table_equal_the_per_human_writer():
    """The batch writer that reads the tfpd_human tables directly gives the same records as
    write_coco_json over the Human objects built from the same tables."""
    from tfpose_b200 import writers, _lib
    from tfpose_b200.decoder import DecodeResult
    rng = np.random.default_rng(5)
    N, MH, H, W = 3, 6, 184, 216
    info = np.zeros(N, _lib.INFO_DTYPE)
    table = np.zeros((N, MH), _lib.HUMAN_DTYPE)
    for i in range(N):
        info["n_humans"][i] = rng.integers(0, MH + 1)
        for j in range(int(info["n_humans"][i])):
            present = rng.random(18) < 0.7
            if j == 0:
                present[:] = False                      # a human without parts is skipped (estimator.py:293)
            table["cid"][i, j] = np.where(present, rng.integers(0, 300, 18), -1)
            table["x"][i, j] = np.where(present, rng.integers(0, W, 18), -1)
            table["y"][i, j] = np.where(present, rng.integers(0, H, 18), -1)
            table["part_score"][i, j] = np.where(present, rng.random(18), 0).astype(np.float32)
            table["score"][i, j] = rng.random()
    res = DecodeResult(info, table, H, W)
    sizes = [(432, 368), (640, 427), (101, 37)]
    got = writers.coco_items_from_table(res, [10, 11, 12], sizes)
    want = []
    for i in range(N):
        want += [writers.coco_result_item(h, 10 + i, *sizes[i]) for h in res.humans(i)]
    assert got == want
    kp = writers.table_to_keypoint_array(res, 1)
    assert kp.shape == (int(info["n_humans"][1]), 18, 3)


def test_humans_to_msg_matches_reference_golden():
    """tests/golden/ros_persons.json holds what the reference's own humans_to_msg
    (scripts/broadcaster_ros.py:19-36) produced for random humans (make_golden.py ros)."""
    import json
    from tfpose_b200 import writers
    rec = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ros_persons.json")))
    humans = []
    for i, h in enumerate(rec["humans"]):
        hm = tp.Human([])
        for k, (x, y, s) in h["parts"].items():
            hm.body_parts[int(k)] = tp.BodyPart('%d-%s' % (i, k), int(k), x, y, s)
        hm.score = h["score"]
        humans.append(hm)
    msg = writers.humans_to_msg(humans)
    got = [[[b.part_id, b.x, b.y, b.confidence] for b in p.body_part] for p in msg.persons]
    assert got == rec["persons"]
    assert msg.image_w == 0 and msg.image_h == 0 and msg.header is None

    class P:                                                 # caller-supplied message classes are used as given
        def __init__(self):
            self.persons = []

    class Q:
        def __init__(self):
            self.body_part = []

    class B:
        pass
    msg2 = writers.humans_to_msg(humans, (P, Q, B))
    assert isinstance(msg2, P) and isinstance(msg2.persons[0], Q) and isinstance(msg2.persons[0].body_part[0], B)
    assert [[[b.part_id, b.x, b.y, b.confidence] for b in p.body_part] for p in msg2.persons] == rec["persons"]


def test_persons_records_from_table_equal_humans_to_msg():
    """The packed BodyPartElm records built from the tfpd_human table hold what humans_to_msg over the Human
    objects of the same table holds, after the float32 narrowing ROS serialisation applies."""
    from tfpose_b200 import writers, _lib
    from tfpose_b200.decoder import DecodeResult
    rng = np.random.default_rng(9)
    N, MH, H, W = 2, 5, 184, 216
    info = np.zeros(N, _lib.INFO_DTYPE)
    table = np.zeros((N, MH), _lib.HUMAN_DTYPE)
    for i in range(N):
        info["n_humans"][i] = MH - i
        for j in range(int(info["n_humans"][i])):
            present = rng.random(18) < 0.6
            if j == 1:
                present[:] = False
            table["cid"][i, j] = np.where(present, rng.integers(0, 300, 18), -1)
            table["x"][i, j] = np.where(present, rng.integers(0, W, 18), -1)
            table["y"][i, j] = np.where(present, rng.integers(0, H, 18), -1)
            table["part_score"][i, j] = np.where(present, rng.random(18), 0).astype(np.float32)
            table["score"][i, j] = rng.random()
    res = DecodeResult(info, table, H, W)
    assert writers.BODY_PART_ELM_DTYPE.itemsize == 16
    for i in range(N):
        rec, counts = writers.persons_records_from_table(res, i)
        msg = writers.humans_to_msg(res.humans(i))
        assert counts.tolist() == [len(p.body_part) for p in msg.persons]
        flat = [b for p in msg.persons for b in p.body_part]
        assert rec["part_id"].tolist() == [b.part_id for b in flat]
        assert rec["x"].tolist() == [float(np.float32(b.x)) for b in flat]
        assert rec["y"].tolist() == [float(np.float32(b.y)) for b in flat]
        assert rec["confidence"].tolist() == [float(np.float32(b.confidence)) for b in flat]


def test_every_option_the_library_accepts_is_documented():
    """tfpd_set_option names in api.cu == names described in include/tfpose_decode.h (and mentioned in
    INTEGRATION.md), so that the header stays the reference for the tuning knobs."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    api = open(os.path.join(root, "tf-pose-estimation_b200", "csrc", "api.cu")).read()
    body = api[api.index("int tfpd_set_option("):]
    body = body[:body.index("\n}\n")]
    accepted = set(re.findall(r'strcmp\(name, "([a-z0-9_]+)"\)', body))
    header = open(os.path.join(root, "include", "tfpose_decode.h")).read()
    # the comment block in front of the declaration, from its first line (not a fixed distance: it grows with the options)
    doc = header[header.index("/* Tuning options, by name"):header.index("TFPD_API int tfpd_set_option")]
    documented = set(re.findall(r'\*\s+"([a-z0-9_]+)"\s', doc))
    assert accepted and accepted == documented, (sorted(accepted - documented), sorted(documented - accepted))
    integ = open(os.path.join(root, "INTEGRATION.md")).read()
    assert all(("`%s`" % n) in integ for n in accepted), [n for n in accepted if ("`%s`" % n) not in integ]


def test_gpu_numa_node_lookup_and_unbind(tmp_path):
    """shard.bind_host_to_gpu_node reads the GPU's NUMA node from sysfs; best effort, never raises."""
    from tfpose_b200 import shard
    dev = tmp_path / "bus" / "pci" / "devices" / "0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("1\n")
    assert shard.gpu_numa_node(0, 0x1b, 0, sysfs=str(tmp_path)) == 1
    assert shard.gpu_numa_node(0, 0x2c, 0, sysfs=str(tmp_path)) == -1
    assert shard._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    done = shard.bind_host_to_gpu_node(0, sysfs=str(tmp_path))       # no CUDA device here: nothing done, no raise
    assert done["node"] == -1 and not done["mempolicy"]
    shard.unbind_host(done)


def test_objbuild_extension_builds_the_same_objects_as_the_python_loop():
    """csrc/objbuild.c (DecodeResult.all_humans) against decoder.humans_from_table, the Python statement of
    estimate_paf's object loop (estimator.py:276-299): same humans, same insertion order, same attribute VALUES AND
    TYPES (Python float / int / str), Human([])'s empty pairs / uidx_list, a human without parts dropped, human ids
    beyond the cached uidx range."""
    from tfpose_b200 import _lib
    from tfpose_b200.decoder import DecodeResult, _objbuild, humans_from_table
    from tfpose_b200.humans import BodyPart, Human
    ob = _objbuild()
    if ob is None:          # a build artefact (not in git): build it like __graft_entry__.build() does, then look again
        import subprocess
        from tfpose_b200 import decoder as _decoder
        subprocess.run(["make", "-C", os.path.join(ROOT, "tf-pose-estimation_b200", "csrc"), "../_objbuild.so"], check=True)
        _decoder._OBJBUILD = False
        ob = _objbuild()
    assert ob is not None, "tf-pose-estimation_b200/_objbuild.so could not be built (make -C tf-pose-estimation_b200/csrc)"
    assert ob.HUMAN_RECORD_BYTES == _lib.HUMAN_DTYPE.itemsize == 184
    rng = np.random.default_rng(3)
    N, MH = 9, 600
    table = np.zeros((N, MH), _lib.HUMAN_DTYPE)
    counts = np.array([0, 1, 3, 6, 2, 600, 1, 4, 5], np.int32)
    for i in range(N):
        for j in range(counts[i]):
            present = rng.random(18) < 0.8
            table[i, j]["cid"] = np.where(present, rng.integers(0, 4000, 18), -1)
            table[i, j]["x"] = rng.integers(0, 656, 18)
            table[i, j]["y"] = rng.integers(0, 368, 18)
            table[i, j]["part_score"] = rng.random(18).astype(np.float32)
            table[i, j]["score"] = np.float32(rng.random() * 2)
    table[3, 1]["cid"] = -1                      # a record without any part: estimate_paf never appends it
    info = np.zeros(N, _lib.INFO_DTYPE)
    info["n_humans"] = counts
    res = DecodeResult(info, table, 368, 656)
    got = res.all_humans()
    assert isinstance(got, list) and len(got) == N
    for i in range(N):
        want = humans_from_table(table[i], counts[i], 368, 656)
        assert isinstance(got[i], list) and len(got[i]) == len(want)
        for a, b in zip(want, got[i]):
            assert type(b) is Human and b.pairs == [] and b.uidx_list == set() and type(b.body_parts) is dict
            assert type(b.score) is float and b.score == a.score
            assert list(b.body_parts) == list(a.body_parts)
            for k, p in a.body_parts.items():
                q = b.body_parts[k]
                assert type(q) is BodyPart
                assert (q.uidx, q.part_idx, q.x, q.y, q.score) == (p.uidx, p.part_idx, p.x, p.y, p.score)
                assert (type(q.uidx), type(q.part_idx), type(q.x), type(q.y), type(q.score)) == (str, int, float, float, float)
                assert str(q) == str(p)
    assert len(got[3]) == 5 and len(got[5]) == 600 and got[5][599].body_parts
    assert next(iter(got[5][599].body_parts.values())).uidx.startswith("599-")
    with pytest.raises(ValueError):
        ob.build_humans(BodyPart, Human, table[:1].tobytes(), counts.tobytes(), MH, 368, 656)    # table too small
    with pytest.raises(TypeError):
        ob.build_humans(int, Human, table.tobytes(), counts.tobytes(), MH, 368, 656)             # no such slots


def test_bench_config_table_matches_the_baseline_configs():
    """bench.py's workload table against BASELINE.json / SURVEY.md 8(d): shapes, frames per GPU at 1/2/4/8 GPUs
    (C3 is STRONG-scaled: 128 frames in total; C2/C4/C5 keep 64 per GPU, C5 = 512 over 8), and the algorithmic
    bytes per frame that roofline.achieved is computed from."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    C = bench.CONFIGS
    assert sorted(C) == ["c2", "c3", "c4", "c5"]
    assert [(C[k]["h"], C[k]["w"]) for k in ("c2", "c3", "c4", "c5")] == [(46, 54), (46, 82), (46, 54), (92, 164)]
    assert [bench.frames_per_rank(C["c3"], n) for n in (1, 2, 4, 8)] == [128, 64, 32, 16]
    assert bench.scaling_of(C["c3"]) == "strong"
    for k in ("c2", "c4", "c5"):
        assert [bench.frames_per_rank(C[k], n) for n in (1, 2, 4, 8)] == [64] * 4 and bench.scaling_of(C[k]) == "weak"
    assert bench.frames_per_rank(C["c5"], 8) * 8 == 512
    assert bench.frames_per_rank(C["c2"], 4, 16) == 16                      # --batch override
    assert bench.dense_bytes_per_frame(46, 54) == 12648528
    assert bench.dense_bytes_per_frame(46, 82) == 19207024
    assert bench.dense_bytes_per_frame(92, 164) == 76828096
    assert abs(bench.polyphase_flops_per_frame(46, 54) / 13.2e6 - 1) < 0.01   # "polyphase: ~13.2 MFLOP" per frame
    assert C["c4"]["people"] == 30


def test_literal_std_sort_matches_libstdcxx(tmp_path):
    """csrc/std_sort_literal.h (the assembly kernel's order for limb candidates with bit-equal scores) against the REAL
    std::sort with the reference's comparator (pafprocess.cpp:97,239-241), compiled together into a small harness
    (tests/native/std_sort_harness.cpp): identical element order on tie-heavy arrays in generation order, on sorted /
    reversed / organ-pipe / all-equal inputs around the 16-element insertion-sort threshold, and on Musser's
    median-of-3 killer sequences, which must reach the depth-limit heap-sort fallback."""
    import ctypes as C
    import subprocess
    so = str(tmp_path / "std_sort_harness.so")
    subprocess.run(["g++", "-O2", "-std=c++14", "-fPIC", "-shared", "-o", so,
                    os.path.join(ROOT, "tests", "native", "std_sort_harness.cpp")], check=True)
    L = C.CDLL(so)
    P = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
    L.harness_literal_sort.argtypes = [P, C.c_int]
    L.harness_std_sort.argtypes = [P, C.c_int]
    rng = np.random.default_rng(1)

    def keys_of(scores, limb=7):
        n = len(scores)
        bits = np.asarray(scores, np.float32).view(np.uint32).astype(np.uint64)
        nb = int(np.ceil(np.sqrt(max(n, 1)))) + 1
        idx = np.arange(n, dtype=np.uint64)
        return (np.uint64(limb) << np.uint64(56)) | ((np.uint64(0xffffffff) - bits) << np.uint64(24)) | \
               ((idx // np.uint64(nb)) << np.uint64(12)) | (idx % np.uint64(nb))

    def same(k):
        a, b = k.copy(), k.copy()
        L.harness_literal_sort(a, len(k))
        L.harness_std_sort(b, len(k))
        return np.array_equal(a, b)

    checked = 0
    for n in list(range(0, 40)) + [63, 64, 65, 129, 500, 1024, 4096]:
        for levels in (1, 2, 5, 1000):
            r = rng.integers(1, levels + 1, n)
            half = np.arange(n // 2)
            organ = (np.concatenate([half, half[::-1], np.zeros(n - 2 * (n // 2), int)])[:n] + 1) if n else r
            for sc in (r, np.sort(r), np.sort(r)[::-1], organ):
                assert same(keys_of(np.asarray(sc, np.float32) / np.float32(max(levels, n, 1)))), (n, levels)
                checked += 1
    before = L.harness_heap_sort_calls()
    for n in (256, 1024, 4096):
        k = n // 2
        v = np.zeros(n)
        for i in range(1, k + 1):                       # Musser 1997
            v[i - 1] = i if i % 2 == 1 else k + i - 1
            v[k + i - 1] = 2 * i
        for seq in (v, v.max() + 1 - v):
            assert same(keys_of((seq / (n + 2.0)).astype(np.float32), limb=3)), n
    assert L.harness_heap_sort_calls() - before >= 6, "the killer sequences did not reach the heap-sort fallback"
    assert checked > 500
