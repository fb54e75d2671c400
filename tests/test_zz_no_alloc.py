"""GPU test of tfpd_peak_table_noalloc (include/tfpose_decode.h): the peak table computed with the library's static
device scratch, no context and no allocation, against the existing entry points (tfpd_process_paf ->
tfpd_fetch_peaks) on the same dense maps.  Index / integer work: bit-exact, scores included (they are copied values).

Collected last on purpose; passes alone in a fresh process and after the other tests.
The entry point has no counterpart in the reference."""
import ctypes as C

import numpy as np
import pytest

import tfpose_b200 as tp
from tfpose_b200 import _lib, synth

pytestmark = pytest.mark.gpu

H_LOW, W_LOW = 46, 54
H, W = 4 * H_LOW, 4 * W_LOW


def _dense_maps(torch, dec, cfg, frame, people=None):
    """peaks / heat_up / paf_up of one synthetic frame, produced by the decoder's own dense mode."""
    x = synth.synth_frame(H_LOW, W_LOW, synth.frame_seed(cfg, frame), people=people)
    dense = tuple(torch.empty((1, H, W, c), device="cuda") for c in (19, 38, 19))
    dec.decode(torch.from_numpy(x[None]).cuda(), (H, W), dense=dense)
    torch.cuda.synchronize()
    heat, paf, peaks = dense           # decode's dense triple: heat_up, paf_up, peaks
    return peaks[0].contiguous(), heat[0].contiguous(), paf[0].contiguous()


def _existing(dec, peaks, heat, paf):
    assert dec.process_paf(peaks.cpu().numpy(), heat.cpu().numpy(), paf.cpu().numpy()) == 0
    return dec.peaks(0)


class _Out:
    def __init__(self, torch, cap):
        self.part, self.x, self.y = (torch.full((cap,), -7, dtype=torch.int32, device="cuda") for _ in range(3))
        self.score = torch.full((cap,), -7.0, dtype=torch.float32, device="cuda")
        self.count = torch.full((2,), -7, dtype=torch.int32, device="cuda")


def _call(torch, L, peaks, heat, thresh, cap, out, stream=None):
    err = C.create_string_buffer(512)
    s = stream.cuda_stream if stream is not None else torch.cuda.current_stream().cuda_stream
    rc = L.tfpd_peak_table_noalloc(peaks.data_ptr(), heat.data_ptr(), peaks.shape[0], peaks.shape[1], thresh, cap,
                                   out.part.data_ptr(), out.x.data_ptr(), out.y.data_ptr(), out.score.data_ptr(),
                                   out.count.data_ptr(), C.c_void_p(s), err, len(err))
    return rc, err.value.decode()


def _table(torch, out):
    torch.cuda.synchronize()
    n, found = (int(v) for v in out.count.cpu())
    return n, found, {"part": out.part[:n].cpu().numpy(), "x": out.x[:n].cpu().numpy(), "y": out.y[:n].cpu().numpy(),
                      "score": out.score[:n].cpu().numpy()}


def test_peak_table_noalloc_matches_process_paf_on_two_calls_in_a_row(cuda):
    torch = cuda
    L = _lib.lib()
    dec = tp.PoseDecoder(max_batch=1, max_upsample=(H, W), max_peaks=4096, max_cands=8192, max_humans=128)
    thresh = float(dec.cfg.thresh_heat)
    # a crowded frame first, a sparse one second: the second call finds the first call's records in the scratch
    crowded = _dense_maps(torch, dec, 4, 0, people=30)
    sparse = _dense_maps(torch, dec, 2, 3)
    want = [_existing(dec, *crowded), _existing(dec, *sparse)]
    assert len(want[0]["part"]) > 2 * len(want[1]["part"]) > 0

    cap = 4096
    outs = [_Out(torch, cap) for _ in range(4)]           # every buffer exists before the first call
    side = torch.cuda.Stream()
    torch.cuda.synchronize()
    free0 = None
    for rep, (maps, ref, stream) in enumerate([(crowded, want[0], None), (sparse, want[1], None),
                                               (crowded, want[0], side), (sparse, want[1], side)]):
        if rep == 1:
            free0 = torch.cuda.mem_get_info()[0]          # after the first call: module and symbol are resolved
        rc, msg = _call(torch, L, maps[0], maps[1], thresh, cap, outs[rep], stream)
        assert rc == 0, msg
        n, found, got = _table(torch, outs[rep])
        assert n == found == len(ref["part"])
        for k in ("part", "x", "y", "score"):
            assert got[k].dtype == ref[k].dtype and np.array_equal(got[k], ref[k]), k
        # nothing is written behind the table
        assert int(outs[rep].part[n:].max() if n < cap else -7) == -7
    assert torch.cuda.mem_get_info()[0] >= free0, "device memory was allocated between calls"
    dec.close()


def test_peak_table_noalloc_limits_and_overflow(cuda):
    torch = cuda
    L = _lib.lib()
    dec = tp.PoseDecoder(max_batch=1, max_upsample=(H, W))
    thresh = float(dec.cfg.thresh_heat)
    peaks, heat, paf = _dense_maps(torch, dec, 2, 0)
    full = _existing(dec, peaks, heat, paf)
    n_full = len(full["part"])
    assert n_full > 8
    out = _Out(torch, 4096)
    # beyond the static scratch: refused with a message, nothing launched
    rc, msg = _call(torch, L, peaks, heat, thresh, 4097, out)
    assert rc == 3 and "TFPD_NOALLOC_MAX_PEAKS" in msg and "4097" in msg
    rc, msg = _call(torch, L, peaks, heat, thresh, 0, out)
    assert rc == 3 and msg
    err = C.create_string_buffer(256)
    rc = L.tfpd_peak_table_noalloc(None, heat.data_ptr(), H, W, thresh, 16, out.part.data_ptr(), out.x.data_ptr(),
                                   out.y.data_ptr(), out.score.data_ptr(), out.count.data_ptr(), None, err, len(err))
    assert rc == 1 and b"null" in err.value
    torch.cuda.synchronize()
    assert int(out.count[0]) == -7                      # untouched by the refused calls
    # a table smaller than the frame's peak count: `cap` entries, in order, all of them real peaks; count[1] tells
    rc, msg = _call(torch, L, peaks, heat, thresh, 8, out)
    assert rc == 0, msg
    n, found, got = _table(torch, out)
    assert n == 8 and found == n_full
    keys = [(int(p), int(y), int(x)) for p, y, x in zip(got["part"], got["y"], got["x"])]
    assert keys == sorted(keys) and len(set(keys)) == 8
    allk = {(int(p), int(y), int(x)): float(s) for p, y, x, s in zip(full["part"], full["y"], full["x"], full["score"])}
    assert all(k in allk and allk[k] == float(s) for k, s in zip(keys, got["score"]))
    # and the next full-size call is not disturbed by the overflowed one
    out2 = _Out(torch, 4096)
    rc, msg = _call(torch, L, peaks, heat, thresh, 4096, out2)
    assert rc == 0, msg
    n, found, got = _table(torch, out2)
    assert n == found == n_full and all(np.array_equal(got[k], full[k]) for k in ("part", "x", "y", "score"))
    dec.close()
