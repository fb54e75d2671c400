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



_REF = {}


def reference_binary(oracle, mode):
    """The reference's own pafprocess.cpp (oracle/_ref, built by oracle/Makefile), one instance per variant: the
    library keeps its results in module globals (pafprocess.cpp:12-13)."""
    if mode not in _REF:
        _REF[mode] = oracle.ReferencePafprocess(mode)
    return _REF[mode]


def peaks_tensor_from_table(pk, H, W):
    """A dense peaks tensor that lists exactly the GPU's peak table when pafprocess.cpp:24-37 scans it: 1.0 (> the
    0.05 heat threshold) at every (y, x, part) of the table, 0 elsewhere.  The peak SCORE comes from the heat map
    (pafprocess.cpp:32), so the value stored here does not matter beyond passing the threshold."""
    t = np.zeros((H, W, 19), np.float32)
    t[pk["y"], pk["x"], pk["part"]] = 1.0
    return t


def _dense(torch, N, H, W, fill=7.0):
    return (torch.full((N, H, W, 19), fill, device="cuda"), torch.full((N, H, W, 38), fill, device="cuda"),
            torch.full((N, H, W, 19), fill, device="cuda"))


def check_frames(torch, oracle, x, factor, mode="asbuilt", dense_mode=True, frames=None, **cfg):
    """Decode the batch x[N,h,w,57] on the GPU and compare every listed frame with the oracle.

    exact      frames without a near-tie: peak table, connections, humans, flags bit-exact vs the oracle
               run on its own maps.
    decomposed frames WITH a near-tie pixel (x4 replication makes them structural): the
               GPU must have flagged the frame, its peak set may differ from the oracle's only AT the
               near-tie pixels, and the stages behind the peak set must then agree bit-exactly on the
               GPU's own peak set: in dense mode the oracle's process_paf is fed the GPU's three dense
               tensors; in minimal mode (no dense tensors exist) it is fed the oracle's heat_up / paf_up
               (bit-exact replicas either way) and a peaks tensor rebuilt from the GPU's peak table.
    Every frame is ALSO decoded by the reference's own compiled pafprocess.cpp (oracle/_ref) from the same three
    maps, and the GPU's humans must equal its getters' answers one for one -- bit-equal score ties included: the
    library orders them with a literal std::sort ("std_sort_ties"), and the oracle runs in its std::sort mode.
    Returns (#exact, #decomposed); ``check_frames.ref_compared`` counts the frames the reference binary saw."""
    N, h, w, _ = x.shape
    H, W = factor * h, factor * w
    dec = tp.PoseDecoder(max_batch=N, max_upsample=(H, W), device=0, mode=mode, **cfg)
    xd = torch.from_numpy(x).cuda()
    dense = _dense(torch, N, H, W) if dense_mode else None
    res = dec.decode(xd, (H, W), dense=dense)
    exact = decomposed = 0
    for i in (range(N) if frames is None else frames):
        R, (peaks, heat_up, paf_up, g) = oracle.decode_frame(x[i], H, W, mode=mode, sort="std")
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
                ref_maps = (gp, hu_g, pu_g)                                          # the GPU's own maps
            else:
                ref_maps = (peaks_tensor_from_table(pk, H, W), heat_up, paf_up)      # the GPU's own peak set
            R = oracle.process_paf(*ref_maps, mode=mode, sort="std")
        else:
            ref_maps = (peaks, heat_up, paf_up)
        assert int(res.info["n_peaks"][i]) == len(R.peaks["x"])
        assert peaks_equal(pk, R), "frame %d peak table differs" % i
        assert conns_equal(dec.connections(i), R), "frame %d connections differ" % i
        ok, why = humans_equal(res.humans(i), R.humans_as_tuples(), H, W)
        assert ok, "frame %d: %s" % (i, why)
        assert bool(res.info["flags"][i] & 1) == bool(R.flags & 1), "score-tie flag differs"
        assert int(res.info["n_rows"][i]) == R.rows_unfiltered
        # the reference's own compiled pafprocess.cpp, on every frame: bit-equal score ties are ordered the way its
        # std::sort orders them ("std_sort_ties", on by default)
        ref = reference_binary(oracle, mode)
        ref.process_paf(*ref_maps)
        ok, why = humans_equal(res.humans(i), ref.humans_as_tuples(), H, W)
        assert ok, "frame %d vs the reference binary (%s): %s" % (i, mode, why)
        check_frames.ref_compared += 1
        if fragile:
            decomposed += 1
        else:
            exact += 1
    dec.close()
    return exact, decomposed


check_frames.ref_compared = 0


@pytest.mark.parametrize("mode", ["asbuilt", "upstream"])
def test_c1_single_frame_thin(cuda, oracle, mode):
    """configs[0]: one 46x54x57 tensor, upsample x4."""
    seen = check_frames.ref_compared
    for frame in range(4):
        x = synth.synth_batch(1, 46, 54, config_id=1, first_frame=frame)
        e, f = check_frames(cuda, oracle, x, 4, mode=mode)
        assert e + f == 1
    assert check_frames.ref_compared - seen == 4, "the reference binary must have checked all four frames"


@pytest.mark.parametrize("dense_mode", [True, False])
def test_c2_thin_batch(cuda, oracle, dense_mode):
    """configs[1] shape, 12 different frames in one call (the oracle takes ~0.1 s per frame)."""
    x = synth.synth_batch(12, 46, 54, config_id=2)
    seen = check_frames.ref_compared
    e, f = check_frames(cuda, oracle, x, 4, dense_mode=dense_mode)
    assert e + f == 12 and e >= 6, "too many near-tie frames (%d) for the exactness check to mean anything" % f
    # near-tie frames are verified too now (both modes), so all 12 reach the reference binary unless a limb has
    # bit-equal scores, which these noisy synthetic maps do not produce
    assert check_frames.ref_compared - seen == 12


def test_c3_cmu_shape(cuda, oracle):
    """configs[2]: 656x368 CMU -> 46x82 maps (two column chunks per row in K1)."""
    x = synth.synth_batch(6, 46, 82, config_id=3, people=5)
    seen = check_frames.ref_compared
    e, f = check_frames(cuda, oracle, x, 4)
    assert e + f == 6 and e >= 2
    assert check_frames.ref_compared - seen == 6
    e, f = check_frames(cuda, oracle, x[:3], 4, dense_mode=False)
    assert e + f == 3


@pytest.mark.parametrize("mode", ["asbuilt", "upstream"])
def test_c4_crowded(cuda, oracle, mode):
    """configs[3]: ~30 people per frame: hundreds of peaks, > 16 candidates per limb, found == 2."""
    x = synth.synth_batch(4, 46, 54, config_id=4, people=30)
    seen = check_frames.ref_compared
    e, f = check_frames(cuda, oracle, x, 4, mode=mode, max_peaks=4096, max_cands=8192, max_humans=128)
    assert e + f == 4
    x = synth.synth_batch(2, 46, 82, config_id=4, people=30, first_frame=10)
    check_frames(cuda, oracle, x, 4, mode=mode, max_peaks=4096, max_cands=8192, max_humans=128)
    check_frames(cuda, oracle, x, 4, mode=mode, dense_mode=False, max_peaks=4096, max_cands=8192, max_humans=128)
    assert check_frames.ref_compared - seen >= 6, "crowded frames must reach the reference binary (> 16 candidates per limb)"


def test_c5_highres_shape(cuda, oracle):
    """configs[4]: 1312x736 -> 92x164 maps, 368x656 upsampled."""
    x = synth.synth_batch(6, 92, 164, config_id=5, people=12)
    seen = check_frames.ref_compared
    e, f = check_frames(cuda, oracle, x, 4, max_peaks=4096, max_cands=8192, max_humans=128)
    assert e + f == 6
    assert check_frames.ref_compared - seen == 6
    e, f = check_frames(cuda, oracle, x[:2], 4, dense_mode=False, max_peaks=4096, max_cands=8192, max_humans=128)
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
    return the same humans from the low-res network output alone.

    Which cases are compared is fixed by the golden file, not by what the GPU flags: a case whose oracle maps have
    no near-tie pixel (``fragile_pixels == 0``) MUST match in both modes, flagged or not.  A case with near-tie
    pixels may differ only if the GPU flagged the frame as well; it is still compared and counted."""
    import json
    import os
    import torch
    from test_oracle import tuples_from_json
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "synth_humans.json")))
    must = [(c["config"], c["frame"]) for c in gold["cases"] if c["fragile_pixels"] == 0]
    assert must == [(2, 0), (2, 1), (2, 2), (2, 5), (5, 0)], "golden file changed: update the expected case list"
    matched, near_tie_matched, near_tie_differs = [], [], []
    for case in gold["cases"]:
        h, w = case["h"], case["w"]
        H, W = 4 * h, 4 * w
        key = (case["config"], case["frame"])
        x = synth.synth_frame(h, w, synth.frame_seed(case["config"], case["frame"]), people=case["people"])
        for mode in ("asbuilt", "upstream"):
            dec = tp.PoseDecoder(max_batch=1, max_upsample=(H, W), mode=mode, max_peaks=4096, max_cands=8192,
                                 max_humans=128)
            res = dec.decode(torch.from_numpy(x[None]).cuda(), (H, W))
            got = [(np.float32(hm.score), {k: (bp.x, bp.y, np.float32(bp.score)) for k, bp in hm.body_parts.items()})
                   for hm in res.humans(0)]
            same = got == tuples_from_json(case[mode])
            flagged = int(res.info["near_ties"][0]) > 0
            dec.close()
            if case["fragile_pixels"] == 0:
                assert same, "golden case %r mode %s differs from the reference's humans" % (key, mode)
                matched.append((key, mode))
            elif same:
                near_tie_matched.append((key, mode))
            else:
                assert flagged, "golden case %r mode %s differs and the GPU did not flag a near-tie" % (key, mode)
                near_tie_differs.append((key, mode))
    assert matched == [(k, m) for k in must for m in ("asbuilt", "upstream")]
    assert len(near_tie_matched) + len(near_tie_differs) == 6      # (3,0), (4,0), (4,1) in both modes
    print("golden humans: %d exact cases matched; near-tie cases matched %r, differ (flagged) %r"
          % (len(matched), near_tie_matched, near_tie_differs))


def test_second_decode_before_fetch_is_ordered_behind_the_first(cuda):
    """One call in flight per ctx (include/tfpose_decode.h): a decode issued on ANOTHER stream before the previous
    one was fetched must wait for it -- both use the ctx's workspace -- and then replace its results, as the next
    process_paf replaces the reference's module globals (pafprocess.cpp:39,128).  The second call's tables must be
    those of a clean decode of the second input, whatever the first call was still doing."""
    import torch
    from helpers import table_tuples
    xa = torch.from_numpy(synth.synth_batch(16, 46, 54, config_id=2, first_frame=300)).cuda()
    xb = torch.from_numpy(synth.synth_batch(16, 46, 54, config_id=2, first_frame=400)).cuda()
    dec = tp.PoseDecoder(max_batch=16, max_upsample=(184, 216))
    want = dec.decode(xb, (184, 216))
    want_pk = [dec.peaks(i) for i in range(16)]
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    for _ in range(20):
        dec.decode_async(xa, (184, 216), dense=_dense(torch, 16, 184, 216), stream=s1)    # long (dense) call, never fetched
        dec.decode_async(xb, (184, 216), stream=s2)
        got = dec.fetch()
        assert np.array_equal(got.info, want.info)
        for i in range(16):
            assert table_tuples(got, i) == table_tuples(want, i)
    pk = [dec.peaks(i) for i in range(16)]
    assert all(np.array_equal(pk[i][k], want_pk[i][k]) for i in range(16) for k in pk[i])
    dec.close()


def test_overflowing_frame_does_not_cost_the_batch_its_other_frames(cuda):
    """A crowded frame inside a batch of sparse ones overflows max_humans: fetch() raises CapacityError (nothing is
    truncated silently), and the error carries the batch's DecodeResult -- the frames whose flags show no overflow
    are complete and equal to what a roomy decoder returns for them."""
    import torch
    from helpers import table_tuples
    x = synth.synth_batch(4, 46, 54, config_id=2, first_frame=500)
    x[2] = synth.synth_frame(46, 54, synth.frame_seed(4, 77), people=30)
    xd = torch.from_numpy(x).cuda()
    roomy = tp.PoseDecoder(max_batch=4, max_upsample=(184, 216), max_peaks=4096, max_cands=8192, max_humans=128)
    want = roomy.decode(xd, (184, 216))
    assert int(want.info["n_humans"][2]) > 8
    tight = tp.PoseDecoder(max_batch=4, max_upsample=(184, 216), max_peaks=4096, max_cands=8192, max_humans=8)
    with pytest.raises(tp.CapacityError) as ei:
        tight.decode(xd, (184, 216))
    res = ei.value.result
    assert res is not None and len(res) == 4
    ovf = 4 | 8 | 16 | 32
    assert [bool(f & ovf) for f in res.info["flags"].tolist()] == [False, False, True, False]
    for i in (0, 1, 3):
        assert table_tuples(res, i) == table_tuples(want, i)
    roomy.close()
    tight.close()
    # TfPosePostProcessor's defaults are the pafprocess drop-in's roomy capacities: the crowded frame just decodes
    post = tp.TfPosePostProcessor(target_size=(432, 368), keep_maps=False)
    assert len(post.inference_from_output(x[2], upsample_size=4.0)) == int(want.info["n_humans"][2])
    post.close()


def test_minimal_mode_early_out_is_exact(cuda):
    """Minimal mode skips low-res row groups whose heat window cannot reach the threshold ("k1_early_out",
    include/tfpose_decode.h).  The skip must change nothing: tables, peak lists and frame info -- near-tie counts
    included -- are identical with the option off, on sparse, crowded and empty frames, on frames whose blobs sit
    just below / just above the bound (the decisive case: a window maximum within 1e-3 of thresh_heat), and with a
    raised heat threshold (more of the map is dead)."""
    import torch
    from helpers import table_tuples

    def both_ways(x, H, W, **cfg):
        xd = torch.from_numpy(x).cuda()
        out = []
        for early in (1, 0):
            dec = tp.PoseDecoder(max_batch=len(x), max_upsample=(H, W), **cfg)
            dec.set_option("k1_early_out", early)
            res = dec.decode(xd, (H, W))
            out.append((res, [dec.peaks(i) for i in range(len(x))]))
            dec.close()
        (ra, pa), (rb, pb) = out
        assert np.array_equal(ra.info, rb.info), "frame info (peaks / near-ties / humans) differs with the early-out"
        for i in range(len(x)):
            assert table_tuples(ra, i) == table_tuples(rb, i)
            assert all(np.array_equal(pa[i][k], pb[i][k]) for k in pa[i])
        return ra

    caps = dict(max_peaks=4096, max_cands=8192, max_humans=128)
    r = both_ways(synth.synth_batch(16, 46, 54, config_id=2, first_frame=700), 184, 216)
    assert int(r.info["n_humans"].sum()) > 0
    both_ways(synth.synth_batch(4, 46, 54, config_id=4, people=30, first_frame=20), 184, 216, **caps)
    both_ways(synth.synth_batch(3, 46, 82, config_id=3, first_frame=40), 184, 328)
    both_ways(np.zeros((2, 46, 54, 57), np.float32), 184, 216)
    # isolated blobs scaled so that the blurred peak / the window maximum straddle the threshold and the bound
    rng = np.random.default_rng(11)
    x = np.zeros((12, 20, 30, 57), np.float32)
    for i, amp in enumerate(np.linspace(0.0495, 0.0515, 12)):
        x[i, :, :, :19] = rng.normal(0, 1e-4, (20, 30, 19))
        for c in range(18):
            cy, cx = rng.integers(2, 18), rng.integers(2, 28)
            x[i, cy, cx, c] = amp                       # single hot cell: window max == amp, blurred peak far lower
            x[i, (cy + 9) % 20, (cx + 11) % 30, c] = amp * 21.0   # and one whose BLURRED value is near the threshold
    both_ways(x, 80, 120, **caps)
    r = both_ways(synth.synth_batch(8, 46, 54, config_id=2, first_frame=800), 184, 216, thresh_heat=0.3)
    both_ways(synth.synth_batch(8, 46, 54, config_id=2, first_frame=800), 184, 216, thresh_heat=1e-3, **caps)


def test_minimal_mode_work_queue_band_sizes_agree(cuda):
    """Minimal mode hands out (frame, column chunk, band of low-res rows) items dynamically and never computes the
    background channel ("k1_min_band", include/tfpose_decode.h).  Whatever the band size -- one row, a size that
    does not divide the height, the whole frame, more than the frame -- tables, peak lists and frame info must equal
    those of a dense-mode decode of the same batch (which still runs the static row spans and all 19 channels),
    on a narrow map, a two-chunk map and a batch larger than the grid's CTA count would need."""
    import torch
    from helpers import table_tuples
    for (n, h, w, cfg) in ((5, 46, 54, 2), (3, 46, 82, 3), (2, 7, 8, 2), (1, 1, 1, 2)):
        x = synth.synth_batch(n, h, w, config_id=cfg, first_frame=900)
        if h < 10:
            rng = np.random.default_rng(h * 31 + w)
            x = rng.random((n, h, w, 57), dtype=np.float32)
            x[..., 19:] = x[..., 19:] * 2 - 1
        H, W = 4 * h, 4 * w
        xd = torch.from_numpy(x).cuda()
        caps = dict(max_peaks=4096, max_cands=8192, max_rows=1024, max_humans=512)
        dec = tp.PoseDecoder(max_batch=n, max_upsample=(H, W), **caps)
        want = dec.decode(xd, (H, W), dense=_dense(torch, n, H, W))
        want_pk = [dec.peaks(i) for i in range(n)]
        for band in (1, 3, 4, 5, 46, 200):
            dec.set_option("k1_min_band", band)
            got = dec.decode(xd, (H, W))
            assert np.array_equal(got.info, want.info), (h, w, band)
            for i in range(n):
                assert table_tuples(got, i) == table_tuples(want, i), (h, w, band, i)
                pk = dec.peaks(i)
                assert all(np.array_equal(pk[k], want_pk[i][k]) for k in pk), (h, w, band, i)
        dec.close()


def test_limb_scoring_early_rejection_is_exact(cuda, oracle):
    """K2 drops a pair after its first 10 - thresh_vector_cnt1 samples when none of them passes ("k2_prefilter").
    Candidates, connections, humans and frame info (n_cands included) must be identical with the option off -- on
    sparse and crowded frames, on random maps (every pair has some passing samples), with other vector thresholds
    (cnt1 = 4: six leading samples; 9: one; 3 and 10: the rejection is off or nothing can qualify) -- and equal to
    the oracle's with the reference's constants."""
    import torch
    from helpers import table_tuples
    caps = dict(max_peaks=4096, max_cands=8192, max_rows=1024, max_humans=512)     # 195 KB of K3 shared memory: the most that fits
    rng = np.random.default_rng(77)
    xr = rng.random((2, 9, 13, 57), dtype=np.float32)
    xr[..., 19:] = xr[..., 19:] * 2 - 1
    batches = [(synth.synth_batch(6, 46, 54, config_id=2, first_frame=1000), 4),
               (synth.synth_batch(2, 46, 54, config_id=4, people=30, first_frame=30), 4), (xr, 4)]
    for x, f in batches:
        H, W = f * x.shape[1], f * x.shape[2]
        xd = torch.from_numpy(x).cuda()
        for cnt1, vs in ((6, 0.05), (4, 0.05), (9, 0.02), (3, 0.05), (10, 0.05), (6, 0.3)):
            res = []
            for pre in (1, 0):
                dec = tp.PoseDecoder(max_batch=len(x), max_upsample=(H, W), thresh_vector_cnt1=cnt1,
                                     thresh_vector_score=vs, **caps)
                dec.set_option("k2_prefilter", pre)
                r = dec.decode(xd, (H, W))
                res.append((r, [dec.connections(i) for i in range(len(x))]))
                dec.close()
            (ra, ca), (rb, cb) = res
            assert np.array_equal(ra.info, rb.info), (cnt1, vs)
            for i in range(len(x)):
                assert table_tuples(ra, i) == table_tuples(rb, i), (cnt1, vs, i)
                assert all(np.array_equal(ca[i][k], cb[i][k]) for k in ca[i]), (cnt1, vs, i)
            if cnt1 == 10:
                assert int(ra.info["n_cands"].sum()) == 0
    # and against the oracle with the reference's constants (check_frames runs with the rejection on by default)
    e, f_ = check_frames(cuda, oracle, batches[1][0], 4, max_peaks=4096, max_cands=8192, max_humans=128)
    assert e + f_ == 2


def test_repeated_decodes_are_bit_identical(cuda):
    """Determinism stress (stands in for compute-sanitizer racecheck, which the GPU pool does not allow): a data race
    between the map kernel's roles (TMA staging slots, mbarrier phases, lazily cleared peak slots), in the fused peak
    ranking or in the assembly kernel's in-place packing would show as run-to-run differences.  The same batch is
    decoded 60 times per mode on alternating streams and contexts; dense tensors (whole buffers, pre-filled with a
    sentinel), tables and frame info must be bit-identical every time."""
    import torch
    from helpers import table_tuples
    x = synth.synth_batch(8, 46, 54, config_id=2, first_frame=1200)
    x[5] = synth.synth_frame(46, 54, synth.frame_seed(4, 1200), people=30)          # one crowded frame in the batch
    xd = torch.from_numpy(x).cuda()
    decs = [tp.PoseDecoder(max_batch=8, max_upsample=(184, 216), max_peaks=4096, max_cands=8192, max_humans=128)
            for _ in range(2)]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    first = {}
    for rep in range(60):
        for dense_mode in (True, False):
            d, st = decs[rep & 1], streams[(rep >> 1) & 1]
            dense = _dense(torch, 8, 184, 216, fill=float(rep + 3)) if dense_mode else None
            res = d.decode(xd, (184, 216), dense=dense, stream=st)
            st.synchronize()
            got = ([table_tuples(res, i) for i in range(8)], res.info.tobytes(),
                   [t.cpu().numpy().tobytes() for t in dense] if dense_mode else None)
            if dense_mode not in first:
                first[dense_mode] = got
            else:
                assert got[0] == first[dense_mode][0] and got[1] == first[dense_mode][1], (rep, dense_mode)
                if dense_mode:
                    assert got[2] == first[True][2], "dense tensors differ between repeats (rep %d)" % rep
    assert first[True][0] == first[False][0]
    for d in decs:
        d.close()
