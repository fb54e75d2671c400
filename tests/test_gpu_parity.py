"""GPU parity tests: the CUDA decode path, called through the C-ABI (ctypes), against the CPU
oracle on the same seeded inputs -- the configurations of BASELINE.json at sizes the oracle finishes
in seconds.  Bars: heat_up / paf_up bit-exact (x4 area resize == replication),
peaks tensor <= 1e-5 relative to max|g|, peak table / connections / humans bit-exact; frames whose
blurred map has a near-tie pixel (decision within 2e-5 relative of flipping) are exempt from the
exactness of the peak set only if the CUDA path flagged them too."""
import numpy as np
import pytest

from tfpose_b200 import synth
import tfpose_b200 as tp

from helpers import MAP_TOL, check_peaks_tensor, conns_equal, humans_equal, peaks_equal, rel_err

pytestmark = pytest.mark.gpu



def _dense(torch, N, H, W, fill=7.0):
    return (torch.full((N, H, W, 19), fill, device="cuda"), torch.full((N, H, W, 38), fill, device="cuda"),
            torch.full((N, H, W, 19), fill, device="cuda"))


def check_frames(torch, oracle, x, factor, mode="asbuilt", dense_mode=True, frames=None, **cfg):
    """Decode the batch x[N,h,w,57] on the GPU and compare every listed frame with the oracle.

    exact      frames without a near-tie: peak table, connections, humans, flags bit-exact vs the oracle
               run on its own maps.
    decomposed frames WITH a near-tie pixel (x4 replication makes them structural): the
               GPU must have flagged the frame, its peak set may differ from the oracle's only AT the
               near-tie pixels, and (dense mode) the oracle's process_paf fed with the GPU's own three
               dense tensors must reproduce the GPU's peaks, connections and humans bit-exactly.
    Returns (#exact, #decomposed)."""
    N, h, w, _ = x.shape
    H, W = factor * h, factor * w
    dec = tp.PoseDecoder(max_batch=N, max_upsample=(H, W), device=0, mode=mode, **cfg)
    xd = torch.from_numpy(x).cuda()
    dense = _dense(torch, N, H, W) if dense_mode else None
    res = dec.decode(xd, (H, W), dense=dense)
    exact = decomposed = 0
    for i in (range(N) if frames is None else frames):
        R, (peaks, heat_up, paf_up, g) = oracle.decode_frame(x[i], H, W, mode=mode, sort="total")
        frag_map = oracle.fragile_pixels(g)
        fragile = int(frag_map.sum())
        if dense_mode:
            hu_g, pu_g, gp = (t[i].cpu().numpy() for t in dense)
            assert np.array_equal(hu_g, heat_up), "frame %d heat_up not bit-exact" % i
            assert np.array_equal(pu_g, paf_up), "frame %d paf_up not bit-exact" % i
            check_peaks_tensor(oracle, gp, peaks, g, fragile, "frame %d" % i)
        pk = dec.peaks(i)
        if fragile:
            assert int(res.info["near_ties"][i]) > 0, "oracle sees a near-tie in frame %d, the GPU did not flag it" % i
            got = set(zip(pk["part"].tolist(), pk["y"].tolist(), pk["x"].tolist()))
            want = set(zip(R.peaks["part"].tolist(), R.peaks["y"].tolist(), R.peaks["x"].tolist()))
            for (c, yy, xx) in got ^ want:
                assert frag_map[yy, xx, c], "frame %d: peak (%d,%d,%d) differs off a near-tie pixel" % (i, c, yy, xx)
            if dense_mode:
                R = oracle.process_paf(gp, hu_g, pu_g, mode=mode, sort="total")     # oracle on the GPU's maps
            else:
                decomposed += 1
                continue
        assert int(res.info["n_peaks"][i]) == len(R.peaks["x"])
        assert peaks_equal(pk, R), "frame %d peak table differs" % i
        assert conns_equal(dec.connections(i), R), "frame %d connections differ" % i
        ok, why = humans_equal(res.humans(i), R.humans_as_tuples(), H, W)
        assert ok, "frame %d: %s" % (i, why)
        assert bool(res.info["flags"][i] & 1) == bool(R.flags & 1), "score-tie flag differs"
        assert int(res.info["n_rows"][i]) == R.rows_unfiltered
        if fragile:
            decomposed += 1
        else:
            exact += 1
    dec.close()
    return exact, decomposed


@pytest.mark.parametrize("mode", ["asbuilt", "upstream"])
def test_c1_single_frame_thin(cuda, oracle, mode):
    """configs[0]: one 46x54x57 tensor, upsample x4."""
    for frame in range(4):
        x = synth.synth_batch(1, 46, 54, config_id=1, first_frame=frame)
        e, f = check_frames(cuda, oracle, x, 4, mode=mode)
        assert e + f == 1


@pytest.mark.parametrize("dense_mode", [True, False])
def test_c2_thin_batch(cuda, oracle, dense_mode):
    """configs[1] shape, 12 different frames in one call (the oracle takes ~0.1 s per frame)."""
    x = synth.synth_batch(12, 46, 54, config_id=2)
    e, f = check_frames(cuda, oracle, x, 4, dense_mode=dense_mode)
    assert e + f == 12 and e >= 6, "too many near-tie frames (%d) for the exactness check to mean anything" % f


def test_c3_cmu_shape(cuda, oracle):
    """configs[2]: 656x368 CMU -> 46x82 maps (two column chunks per row in K1)."""
    x = synth.synth_batch(6, 46, 82, config_id=3, people=5)
    e, f = check_frames(cuda, oracle, x, 4)
    assert e + f == 6 and e >= 2


@pytest.mark.parametrize("mode", ["asbuilt", "upstream"])
def test_c4_crowded(cuda, oracle, mode):
    """configs[3]: ~30 people per frame: hundreds of peaks, > 16 candidates per limb, found == 2."""
    x = synth.synth_batch(4, 46, 54, config_id=4, people=30)
    e, f = check_frames(cuda, oracle, x, 4, mode=mode, max_peaks=4096, max_cands=8192, max_humans=128)
    assert e + f == 4
    x = synth.synth_batch(2, 46, 82, config_id=4, people=30, first_frame=10)
    check_frames(cuda, oracle, x, 4, mode=mode, max_peaks=4096, max_cands=8192, max_humans=128)


def test_c5_highres_shape(cuda, oracle):
    """configs[4]: 1312x736 -> 92x164 maps, 368x656 upsampled."""
    x = synth.synth_batch(2, 92, 164, config_id=5, people=12)
    e, f = check_frames(cuda, oracle, x, 4, max_peaks=4096, max_cands=8192, max_humans=128)
    assert e + f == 2


@pytest.mark.parametrize("hw", [(1, 1), (2, 3), (3, 57), (5, 113), (7, 8), (9, 56), (13, 29)])
def test_odd_shapes_random_maps(cuda, oracle, hw):
    """Ragged sizes: single cell, widths around the 56-cell chunk / 28-cell copy boundaries, odd
    heights around the band split.  Uniform random maps: every pixel above the heat threshold."""
    h, w = hw
    rng = np.random.default_rng(h * 1000 + w)
    x = rng.random((3, h, w, 57), dtype=np.float32)
    x[..., 19:] = x[..., 19:] * 2 - 1
    check_frames(cuda, oracle, x, 4, max_peaks=4096, max_cands=8192, max_rows=1024, max_humans=512)


def test_empty_maps(cuda, oracle):
    """All-zero network output: g == 0 everywhere is a plateau of 'peaks' with value 0 -> nothing
    passes the 0.05 threshold, no humans; negative maps likewise."""
    import torch
    for val in (0.0, -1.0):
        x = np.full((2, 46, 54, 57), val, np.float32)
        dec = tp.PoseDecoder(max_batch=2, max_upsample=(184, 216))
        dense = _dense(torch, 2, 184, 216)
        res = dec.decode(torch.from_numpy(x).cuda(), (184, 216), dense=dense)
        assert res.info["n_peaks"].tolist() == [0, 0] and res.info["n_humans"].tolist() == [0, 0]
        assert res.all_humans() == [[], []]
        pk, hu, pu, g = oracle.dense_maps(x[0], 184, 216)
        assert np.array_equal(dense[0][0].cpu().numpy(), hu) and np.array_equal(dense[1][1].cpu().numpy(), pu)
        gp = dense[2][0].cpu().numpy()
        assert rel_err(gp, pk, np.maximum(np.abs(g), 1e-6)) <= MAP_TOL
        dec.close()


def test_table_overflow_is_reported_not_truncated(cuda):
    """More peaks than max_peaks: the call must fail loudly (TFPD_ERR_CAPACITY) and name the table;
    same for the candidate table."""
    import torch
    rng = np.random.default_rng(5)
    x = rng.random((1, 13, 29, 57), dtype=np.float32)
    xd = torch.from_numpy(x).cuda()
    dec = tp.PoseDecoder(max_batch=1, max_upsample=(52, 116), max_peaks=32)
    with pytest.raises(tp.CapacityError) as ei:
        dec.decode(xd, (52, 116))
    assert "flags=0x" in str(ei.value) and (int(str(ei.value).split("flags=0x")[1].split(":")[0], 16) & 4)
    dec.close()
    dec = tp.PoseDecoder(max_batch=1, max_upsample=(52, 116), max_peaks=4096, max_cands=32)
    with pytest.raises(tp.CapacityError):
        dec.decode(xd, (52, 116))
    dec.close()


def test_minimal_and_dense_mode_agree(cuda):
    import torch
    from helpers import table_tuples
    x = synth.synth_batch(8, 46, 54, config_id=2, first_frame=100)
    xd = torch.from_numpy(x).cuda()
    dec = tp.PoseDecoder(max_batch=8, max_upsample=(184, 216))
    r_min = dec.decode(xd, (184, 216))
    pk_min = [dec.peaks(i) for i in range(8)]
    r_den = dec.decode(xd, (184, 216), dense=_dense(torch, 8, 184, 216))
    for i in range(8):
        assert table_tuples(r_min, i) == table_tuples(r_den, i)
        pk = dec.peaks(i)
        assert all(np.array_equal(pk[k], pk_min[i][k]) for k in pk)
    assert np.array_equal(r_min.info, r_den.info)
    dec.close()


def test_golden_humans_from_reference_code(cuda):
    """tests/golden/synth_humans.json holds what the REFERENCE's own estimate_paf + pafprocess.cpp
    returned for these seeded frames (made by tests/golden/make_golden.py); the CUDA path must
    return the same humans from the low-res network output alone."""
    import json
    import os
    import torch
    from test_oracle import tuples_from_json
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "synth_humans.json")))
    compared = 0
    for case in gold["cases"]:
        h, w = case["h"], case["w"]
        H, W = 4 * h, 4 * w
        x = synth.synth_frame(h, w, synth.frame_seed(case["config"], case["frame"]), people=case["people"])
        for mode in ("asbuilt", "upstream"):
            dec = tp.PoseDecoder(max_batch=1, max_upsample=(H, W), mode=mode, max_peaks=4096, max_cands=8192,
                                 max_humans=128)
            res = dec.decode(torch.from_numpy(x[None]).cuda(), (H, W))
            if int(res.info["near_ties"][0]) or (int(res.info["flags"][0]) & 1):
                dec.close()
                continue        # flagged frame: exactness not required
            got = [(np.float32(hm.score), {k: (bp.x, bp.y, np.float32(bp.score)) for k, bp in hm.body_parts.items()})
                   for hm in res.humans(0)]
            assert got == tuples_from_json(case[mode]), (case["config"], case["frame"], mode)
            compared += 1
            dec.close()
    assert compared >= 10
