"""GPU tests at BASELINE.json's full sizes, where the CPU oracle would take minutes: parity is
carried by size-independent properties plus an oracle spot check of a few frames.

  * batch independence: frame i of a big batch == the same frame decoded alone (bit-exact tables and maps)
  * replication: heat_up / paf_up == the network output repeated x4 (a plain torch index op)
  * table <-> tensor consistency: the pixels of the dense peaks tensor above the threshold ARE the peak
    table (count, coordinates in (part, y, x) order), and every peak score is heat_up at that pixel
  * humans only reference existing peaks; scores are sum/count of what they reference
  * determinism: two runs give identical bytes
"""
import numpy as np
import pytest

import tfpose_b200 as tp
from tfpose_b200 import synth

from helpers import humans_equal, table_tuples

pytestmark = pytest.mark.gpu


def _batch(n_unique, N, h, w, cfg, people=None):
    base = synth.synth_batch(n_unique, h, w, config_id=cfg, people=people)
    reps = (N + n_unique - 1) // n_unique
    return np.concatenate([base] * reps)[:N].copy(), n_unique


def _run_properties(torch, oracle, N, h, w, cfg, people, n_unique, spot, **caps):
    H, W = 4 * h, 4 * w
    x, nu = _batch(n_unique, N, h, w, cfg, people)
    xd = torch.from_numpy(x).cuda()
    dec = tp.PoseDecoder(max_batch=N, max_upsample=(H, W), **caps)
    dense = (torch.empty((N, H, W, 19), device="cuda"), torch.empty((N, H, W, 38), device="cuda"),
             torch.empty((N, H, W, 19), device="cuda"))
    res = dec.decode(xd, (H, W), dense=dense)
    thr = 0.05

    # replication (bit-exact), checked on the device with a plain index op
    up = xd.repeat_interleave(4, dim=1).repeat_interleave(4, dim=2)
    assert torch.equal(dense[0], up[..., :19]) and torch.equal(dense[1], up[..., 19:])
    del up

    # dense peaks tensor <-> peak table
    pk_t = dense[2][..., :18]
    counts = (pk_t > thr).sum(dim=(1, 2, 3)).cpu().numpy()
    assert np.array_equal(counts, res.info["n_peaks"])
    for i in list(range(0, N, max(1, N // 6)))[:6]:
        tab = dec.peaks(i)
        idx = torch.nonzero(pk_t[i].permute(2, 0, 1) > thr).cpu().numpy()       # (part, y, x) row-major order
        assert np.array_equal(idx[:, 0], tab["part"]) and np.array_equal(idx[:, 1], tab["y"]) \
            and np.array_equal(idx[:, 2], tab["x"])
        hv = dense[0][i].cpu().numpy()[tab["y"], tab["x"], tab["part"]]
        assert np.array_equal(hv, tab["score"])
        # peaks tensor values are local maxima: nonzero support == (g == maxpool(g)) can not be checked without g,
        # but every table peak must be >= its 8 neighbours' peak values (they are 0 unless plateau)
        # humans reference existing peaks and their score is consistent
        n_h = int(res.info["n_humans"][i])
        for r in res.table[i][:n_h]:
            cids = r["cid"]
            present = cids >= 0
            assert present.sum() >= 4 and (cids[present] < len(tab["x"])).all()
            assert np.array_equal(r["x"][present], tab["x"][cids[present]])
            assert np.array_equal(r["y"][present], tab["y"][cids[present]])
            assert np.array_equal(r["part_score"][present], tab["score"][cids[present]])
            assert np.array_equal(tab["part"][cids[present]], np.nonzero(present)[0])
            assert r["score"] >= 0.3 - 1e-6

    # repeated frames inside the batch decode identically (batch independence across CTAs / bands)
    for i in range(nu, N):
        assert table_tuples(res, i) == table_tuples(res, i % nu), "frame %d != its twin %d" % (i, i % nu)
    assert torch.equal(dense[2][nu:2 * nu], dense[2][:nu]) if N >= 2 * nu else True

    # a frame decoded alone == the same frame inside the batch
    solo = tp.PoseDecoder(max_batch=1, max_upsample=(H, W), **caps)
    d1 = (torch.empty((1, H, W, 19), device="cuda"), torch.empty((1, H, W, 38), device="cuda"),
          torch.empty((1, H, W, 19), device="cuda"))
    for i in (0, nu - 1):
        r1 = solo.decode(xd[i:i + 1].contiguous(), (H, W), dense=d1)
        assert table_tuples(r1, 0) == table_tuples(res, i)
        assert torch.equal(d1[2][0], dense[2][i])
    solo.close()

    # determinism
    first = [table_tuples(res, i) for i in range(N)]
    keep = dense[2].clone()
    res2 = dec.decode(xd, (H, W), dense=dense)
    assert [table_tuples(res2, i) for i in range(N)] == first and torch.equal(keep, dense[2])
    assert np.array_equal(res.info, res2.info)

    # oracle spot check
    for i in range(spot):
        R, (_, _, _, g) = oracle.decode_frame(x[i], H, W, sort="std")
        if int(oracle.fragile_pixels(g).sum()) or int(res.info["near_ties"][i]):
            continue
        ok, why = humans_equal(res.humans(i), R.humans_as_tuples(), H, W)
        assert ok, why
    dec.close()


def test_c2_full_batch_64(cuda, oracle):
    """configs[1]: mobilenet_thin 432x368, batch 64, one GPU."""
    _run_properties(cuda, oracle, 64, 46, 54, 2, None, 16, spot=2)


def test_c3_per_gpu_share_64(cuda, oracle):
    """configs[2]: CMU 656x368, batch 128 over 2 GPUs -> 64 frames per GPU."""
    _run_properties(cuda, oracle, 64, 46, 82, 3, 5, 8, spot=1)


def test_c4_crowded_batch_64(cuda, oracle):
    """configs[3]: ~30 people per frame at the C2 batch size."""
    _run_properties(cuda, oracle, 64, 46, 54, 4, 30, 4, spot=1, max_peaks=4096, max_cands=8192, max_humans=128)


def test_c5_per_gpu_share_64(cuda, oracle):
    """configs[4]: 1312x736 (92x164 -> 368x656), batch 512 over 8 GPUs -> 64 frames per GPU
    (4.7 GB of dense outputs)."""
    _run_properties(cuda, oracle, 64, 92, 164, 5, 12, 2, spot=1, max_peaks=4096, max_cands=8192, max_humans=128)
