"""CPU tests: the oracle is pinned against the reference's own code and golden vectors.
(oracle/ is test infrastructure; nothing here touches the CUDA library's compute paths.)"""
import hashlib
import json
import os

import numpy as np
import pytest

import tfpose_b200  # noqa: F401
from tfpose_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HAVE_REF = os.path.exists(os.path.join(os.path.dirname(GOLD), "..", "oracle", "_ref", "libpafprocess_asbuilt.so")) \
    or os.path.isdir("/root/reference")


def golden_cases():
    with open(os.path.join(GOLD, "synth_humans.json")) as fh:
        return json.load(fh)


def tuples_from_json(humans):
    return [(np.float32(h["score"]), {int(k): (v[1], v[2], np.float32(v[3])) for k, v in h["parts"].items()})
            for h in humans]


def tuples_from_result(res_tuples, H, W):
    return [(np.float32(s), {k: (x / W, y / H, np.float32(ps)) for k, (x, y, ps) in parts.items()})
            for s, parts in res_tuples]


def test_gauss_kernel_matches_reference_golden(oracle):
    gold = np.load(os.path.join(GOLD, "gauss_kernel_25_3.npy"))
    k, s = oracle.gauss_kernel(25, 3.0)
    assert k.dtype == np.float32 and np.array_equal(k, gold)
    # separable: K = s s^T within float32 rounding, s sums to 1
    assert abs(s.sum() - 1.0) < 1e-12
    assert np.max(np.abs(np.outer(s, s) - gold.astype(np.float64)) / gold) < 2e-7


def test_resize_area_integer_factors_replicate(oracle):
    rng = np.random.default_rng(0)
    img = rng.standard_normal((5, 7, 3)).astype(np.float32)
    for f in (1, 2, 4, 8):
        up = oracle.resize_area(img, 5 * f, 7 * f)
        assert np.array_equal(up, np.repeat(np.repeat(img, f, 0), f, 1)), f
    up3 = oracle.resize_area(img, 15, 21)       # 1/3 is inexact: close, not bit-exact
    assert np.allclose(up3, np.repeat(np.repeat(img, 3, 0), 3, 1), rtol=0, atol=5e-6)


def test_resize_area_general_properties(oracle):
    rng = np.random.default_rng(1)
    const = np.full((6, 9, 2), 0.37, np.float32)
    assert np.allclose(oracle.resize_area(const, 11, 13), 0.37, atol=1e-6)
    img = rng.random((8, 8, 1)).astype(np.float32)
    down = oracle.resize_area(img, 4, 4)        # exact 2x2 box average
    box = img.reshape(4, 2, 4, 2, 1).mean(axis=(1, 3))
    assert np.allclose(down, box, atol=1e-6)


def test_nms_plateau_and_border(oracle):
    g = np.zeros((4, 5, 1), np.float32)
    g[0, 0, 0] = 1.0            # corner maximum: window is clipped, still a peak
    g[2, 2, 0] = g[2, 3, 0] = 0.5   # plateau: both kept
    p = oracle.nms(g)
    assert p[0, 0, 0] == 1.0 and p[2, 2, 0] == 0.5 and p[2, 3, 0] == 0.5
    assert p[1, 1, 0] == 0.0    # neighbour of the corner maximum
    assert p[3, 0, 0] == 0.0 and oracle.nms(-g)[0, 0, 0] == 0.0     # where(eq, g, 0): zeros stay zeros


def test_blur_is_the_direct_convolution(oracle):
    rng = np.random.default_rng(2)
    up = rng.random((30, 33, 2)).astype(np.float32)
    k, _ = oracle.gauss_kernel()
    g32, g64 = oracle.blur(up, k), oracle.blur(up, k, f64=True)
    assert np.max(np.abs(g32 - g64)) / np.max(np.abs(g64)) < 5e-6     # 625 sequential float32 taps
    Y, X = 3, 29     # taps falling outside are dropped (SAME zero padding)
    acc = 0.0
    for r in range(25):
        for q in range(25):
            yy, xx = Y + r - 12, X + q - 12
            if 0 <= yy < 30 and 0 <= xx < 33:
                acc += float(k[r, q]) * float(up[yy, xx, 1])
    assert abs(acc - g64[Y, X, 1]) < 1e-12


def test_maps_restatement_agrees_with_independent_torch_ops(oracle):
    """TensorFlow is unavailable (parity of this half stays unpinned), but the op definitions the graph uses
    (estimator.py:326-337, smoother.py:91) also exist as independent CPU code in PyTorch: adaptive average pooling
    is the area resize for integer factors, conv2d(groups=C, padding=12) is the SAME-padded depthwise convolution,
    max_pool2d(3, 1, 1) pads with -inf = the clipped 3x3 window.  The C restatement must agree with them."""
    import torch
    import torch.nn.functional as F
    torch.set_num_threads(2)
    k2d, _ = oracle.gauss_kernel()
    x = synth.synth_frame(46, 54, synth.frame_seed(2, 3))
    heat, paf = oracle.split(x)
    assert np.array_equal(heat, x[:, :, :19]) and np.array_equal(paf, x[:, :, 19:])
    for f in (1, 2, 4):
        H, W = 46 * f, 54 * f
        th = torch.from_numpy(heat).permute(2, 0, 1)[None]                         # NCHW
        up_t = F.interpolate(th, size=(H, W), mode="area")
        up_o = oracle.resize_area(heat, H, W)
        assert np.array_equal(up_t[0].permute(1, 2, 0).numpy(), up_o), f          # replication: bit-exact
        w = torch.from_numpy(k2d)[None, None].repeat(19, 1, 1, 1)
        g_t = F.conv2d(up_t, w, padding=12, groups=19)
        g64 = F.conv2d(up_t.double(), w.double(), padding=12, groups=19)[0].permute(1, 2, 0).numpy()
        g_o = oracle.blur(up_o, k2d)
        scale = np.abs(g64).max()
        assert np.abs(g_o - g64).max() / scale < 5e-6, f                           # the oracle against float64 truth
        assert np.abs(g_t[0].permute(1, 2, 0).numpy() - g64).max() / scale < 5e-6  # and so is torch's float32
        # NMS on the SAME smoothed map: equality with the pooled maximum, bit for bit
        g_in = torch.from_numpy(g_o).permute(2, 0, 1)[None]
        m = F.max_pool2d(g_in, 3, 1, 1)
        pk_t = torch.where(g_in == m, g_in, torch.zeros_like(g_in))[0].permute(1, 2, 0).numpy()
        assert np.array_equal(pk_t, oracle.nms(g_o)), f
    # integer down-scale: the area resize is the block mean
    small = oracle.resize_area(heat[:44, :54], 22, 27)
    ref = F.interpolate(torch.from_numpy(heat[:44, :54].copy()).permute(2, 0, 1)[None], size=(22, 27), mode="area")
    assert np.abs(small - ref[0].permute(1, 2, 0).numpy()).max() < 1e-6


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref (reference pafprocess.cpp) not built")
@pytest.mark.parametrize("mode", ["asbuilt", "upstream"])
def test_paf_restatement_matches_reference_binary(oracle, mode):
    ref = oracle.ReferencePafprocess(mode)
    for cfg, frame, h, w, people in [(2, 0, 46, 54, None), (2, 2, 46, 54, None), (4, 0, 46, 54, 30), (3, 1, 46, 82, 6)]:
        x = synth.synth_frame(h, w, synth.frame_seed(cfg, frame), people=people)
        pk, hu, pu, _ = oracle.dense_maps(x, 4 * h, 4 * w)
        ref.process_paf(pk, hu, pu)
        R = oracle.process_paf(pk, hu, pu, mode=mode, sort="std")
        assert ref.humans_as_tuples() == R.humans_as_tuples(), (mode, cfg, frame)
        if not (R.flags & 1):       # without exact score ties the total order gives the same result
            assert oracle.process_paf(pk, hu, pu, mode=mode, sort="total").humans_as_tuples() == R.humans_as_tuples()


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref (reference pafprocess.cpp) not built")
@pytest.mark.parametrize("mode", ["asbuilt", "upstream"])
def test_paf_restatement_matches_reference_on_random_dense_with_ties(oracle, mode):
    """Sparse random dense maps with quantised values: many exact ties, > 16 candidates per limb,
    found == 2 / found > 2 situations -- std::sort instability included (same libstdc++)."""
    ref = oracle.ReferencePafprocess(mode)
    for seed in range(12):
        rng = np.random.default_rng(100 + seed)
        H, W = 40, 48
        peaks = np.zeros((H, W, 19), np.float32)
        n = int(rng.integers(20, 160))
        ys, xs, cs = rng.integers(0, H, n), rng.integers(0, W, n), rng.integers(0, 18, n)
        peaks[ys, xs, cs] = (rng.integers(1, 5, n) / 4.0).astype(np.float32)
        heat = peaks + np.float32(0.125)
        paf = (rng.integers(-2, 3, (H, W, 38)) / 2.0).astype(np.float32)
        ref.process_paf(peaks, heat, paf)
        R = oracle.process_paf(peaks, heat, paf, mode=mode, sort="std")
        assert ref.humans_as_tuples() == R.humans_as_tuples(), (mode, seed)


def test_oracle_matches_golden_humans(oracle):
    gold = golden_cases()
    for case in gold["cases"]:
        x = synth.synth_frame(case["h"], case["w"], synth.frame_seed(case["config"], case["frame"]),
                              people=case["people"])
        assert hashlib.sha256(x.tobytes()).hexdigest() == case["x_sha256"], "synthetic generator changed"
        H, W = 4 * case["h"], 4 * case["w"]
        pk, hu, pu, _ = oracle.dense_maps(x, H, W)
        sha = hashlib.sha256(pk.tobytes() + hu.tobytes() + pu.tobytes()).hexdigest()
        assert sha == case["maps_sha256"], "oracle dense maps differ from the ones the golden was made with"
        for mode in ("asbuilt", "upstream"):
            R = oracle.process_paf(pk, hu, pu, mode=mode, sort="std")
            got = tuples_from_result(R.humans_as_tuples(), H, W)
            assert got == tuples_from_json(case[mode]), (case["config"], case["frame"], mode)


def test_oracle_matches_golden_toy(oracle):
    gold = golden_cases()["toy"]
    d = np.load(os.path.join(GOLD, "toy_dense.npz"))
    H, W = d["heat"].shape[:2]
    for mode in ("asbuilt", "upstream"):
        R = oracle.process_paf(d["peaks"], d["heat"], d["paf"], mode=mode)
        assert tuples_from_result(R.humans_as_tuples(), H, W) == tuples_from_json(gold[mode])
        assert len(gold[mode]) == 2


@pytest.mark.parametrize("mode", ["asbuilt", "upstream"])
def test_tie_order_matters_and_the_std_mode_is_the_reference(oracle, mode):
    """Premise of the GPU test of "std_sort_ties": on helpers.chained_tie_case inputs (dozens of bit-equal candidates
    per limb competing for the same peaks) the reference's own compiled pafprocess.cpp returns what the restatement
    returns with the real std::sort, and NOT what the total order (score desc, idx1, idx2) returns."""
    from helpers import chained_tie_case
    ref = oracle.ReferencePafprocess(mode)
    differ = 0
    for seed in range(10):
        pk, hu, pu = chained_tie_case(seed)
        Rs = oracle.process_paf(pk, hu, pu, mode=mode, sort="std")
        Rt = oracle.process_paf(pk, hu, pu, mode=mode, sort="total")
        ref.process_paf(pk, hu, pu)
        assert ref.humans_as_tuples() == Rs.humans_as_tuples()
        assert (Rs.flags & 1) and (Rt.flags & 1) and len(Rs.humans_as_tuples()) > 0
        assert max(len(c[0]) for c in Rs.cands) > 16
        differ += int(Rs.humans_as_tuples() != Rt.humans_as_tuples())
    assert differ >= 8
