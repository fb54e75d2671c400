#!/usr/bin/env python
"""bench.py -- decoded frames/s of the heatmap/PAF decode path (BASELINE.json metric).

One "step" = one batch of frames through the hot path:
  K1 fused upsample+smoother+NMS (dense outputs written: heat_up, paf_up, peaks -- the three
  tensors the reference's sess.run returns) -> K2 limb scoring (ranks the peaks itself) ->
  K3 assembly -> human tables copied to the host.

Workloads (BASELINE.json configs; --config picks the one the line's `value` is quoted on):
  c2  mobilenet_thin 432x368: X[64,46,54,57] per GPU, upsample x4 (default; weak scaling)
  c3  CMU 656x368: X[128,46,82,57] in total, sharded by image over the GPUs (STRONG scaling:
      128 / 64 / 32 / 16 frames per GPU at 1 / 2 / 4 / 8 GPUs)
  c4  crowded scene: ~30 people per frame on the c2 shape, 64 frames per GPU
  c5  1312x736: X[64,92,164,57] per GPU = 512 frames over 8 GPUs
The default line is c2 and carries the other three under `per_config`.

  value      frames/s with the inputs resident in HBM (dense mode), CUDA-event timed, max over ranks.
  e2e        frames/s through the host-buffer C-ABI call (tfpd_decode_host_async + tfpd_fetch_humans):
             pinned host input, H2D inside the timed region, human tables D2H.  Dense maps stay on
             the device (inference() returns humans); e2e_dense_maps also ships the three maps back,
             e2e_objects also builds the list[list[Human]] the reference returns.
  roofline   fused K1 kernel: algorithmic bytes 4*(57hw + 76HW) per frame / its CUDA-event time,
             against MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline / --impl reference: the same decode on the host cores (see cpu_arm below).

Multi-GPU: frames are independent, the batch is sharded by image; no collective on the data path,
only the timing reduction and a final gather of per-rank figures.
"""
import argparse
import gc
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FACTOR = 4
N_INPUT_SETS = 4          # rotating input batches: together with the outputs of a step they exceed the 126 MB L2

# BASELINE.json configs (C2..C5).  `total` = frames of the whole job (strong scaling: split over the
# ranks), `batch` = frames per GPU (weak scaling).  Capacities only where the defaults would overflow.
CONFIGS = {
    "c2": {"id": 2, "h": 46, "w": 54, "people": None, "batch": 64, "caps": {},
           "name": "mobilenet_thin 432x368 decode, batch %d per GPU, 46x54x57 -> 184x216 (upsample x4), dense maps "
                   "(heat_up+paf_up+peaks) produced, humans to host"},
    "c3": {"id": 3, "h": 46, "w": 82, "people": None, "total": 128, "caps": {},
           "name": "CMU 656x368 decode, 128 frames sharded by image (%d per GPU), 46x82x57 -> 184x328 (upsample x4), "
                   "dense maps produced, humans to host"},
    "c4": {"id": 4, "h": 46, "w": 54, "people": 30, "batch": 64,
           "caps": {"max_peaks": 4096, "max_cands": 8192, "max_humans": 128},
           "name": "crowded scene (~30 people per frame) on the mobilenet_thin shape, batch %d per GPU, 46x54x57 -> "
                   "184x216, dense maps produced, humans to host"},
    "c5": {"id": 5, "h": 92, "w": 164, "people": None, "batch": 64, "caps": {},
           "name": "high-res 1312x736 decode, batch %d per GPU (512 over 8 GPUs), 92x164x57 -> 368x656 (upsample "
                   "x4), dense maps produced, humans to host"},
}


def frames_per_rank(cfg, world, override=None):
    if override:
        return int(override)
    if "total" in cfg:
        return max(1, cfg["total"] // world)
    return cfg["batch"]


def scaling_of(cfg):
    return "strong" if "total" in cfg else "weak"


def dense_bytes_per_frame(h, w, f=FACTOR):
    return 4 * (57 * h * w + 76 * (f * h) * (f * w))


# polyphase form of the smoother (what K1 executes): 7-tap vertical pass per output pixel + 7-tap
# horizontal pass once per four output rows... counted per output pixel of the 19 heat channels as 2*(7 + 7/4) flop
def polyphase_flops_per_frame(h, w, f=FACTOR):
    return 2.0 * (7.0 + 7.0 / 4.0) * 19 * (f * h) * (f * w)


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json, burst copy figure: K1 is timed alone)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def profile_constant(name, key):
    """A per-launch figure that only a profiler can give (ncu `--set full`), kept under profiles/ with the
    capture it came from; None when absent."""
    try:
        with open(os.path.join(ROOT, "profiles", name)) as fh:
            return json.load(fh).get(key)
    except Exception:
        return None


def host_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


# ------------------------------------------------------------------------------------------ clocks
# NVML is polled from a separate PROCESS: a sampler thread inside the timed process has to take the GIL every
# period, and with eight ranks on one host those wake-ups landed inside the few-millisecond timed regions
# (driver-run N=8 efficiency 0.845 against 0.98 with identical kernels).
_SAMPLER_SRC = r'''
import json, select, sys, time
idx = [int(a) for a in sys.argv[2:]]
period = float(sys.argv[1])
out = {"t": [], "mhz": [], "bits": [], "max_mhz": None, "error": None}
try:
    import pynvml as nv
    nv.nvmlInit()
    hs = [nv.nvmlDeviceGetHandleByIndex(i) for i in idx]
    out["max_mhz"] = max(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM) for h in hs)
    while True:
        r, _, _ = select.select([sys.stdin], [], [], period)
        if r:
            break
        t = time.time()
        for h in hs:
            out["t"].append(t)
            out["mhz"].append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
            out["bits"].append(int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
except Exception as e:
    out["error"] = repr(e)
sys.stdout.write(json.dumps(out))
'''
_REASON_BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}


class ClockSampler(object):
    """SM clock and throttle reasons of the job's GPUs while the timed regions run (rank 0 starts it)."""

    def __init__(self, indices, period=0.002):
        self.windows = []
        try:
            self.proc = subprocess.Popen([sys.executable, "-c", _SAMPLER_SRC, str(period)] + [str(i) for i in indices],
                                         stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def window(self, t0, t1):
        self.windows.append((t0, t1))

    def result(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        try:
            raw, _ = self.proc.communicate(b"x", timeout=10)
            out = json.loads(raw.decode())
        except Exception:
            self.proc.kill()
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        if not out["mhz"]:
            return {"sm_mhz": None, "sm_max_mhz": out["max_mhz"], "reasons": ["unavailable"], "error": out["error"]}
        t, mhz, bits = np.array(out["t"]), np.array(out["mhz"]), np.array(out["bits"], dtype=np.int64)
        inside = np.zeros(len(t), bool)
        for (a, b) in self.windows:
            inside |= (t >= a) & (t <= b)
        sel = inside if inside.any() else np.ones(len(t), bool)
        seen = 0
        for b_ in bits[sel]:
            seen |= int(b_)
        return {"sm_mhz": float(np.median(mhz[sel])), "sm_max_mhz": out["max_mhz"],
                "reasons": sorted(n for b_, n in _REASON_BITS.items() if seen & b_),
                "samples": int(sel.sum()), "samples_in_timed_regions": int(inside.sum()), "samples_total": len(t)}


# ------------------------------------------------------------------------------------------ CPU arm
# The reference's CPU decode, stage by stage:
#   A  the TF graph ops of estimator.py:321-337 + smoother.py:62-95.  TensorFlow is unavailable on this box, so the
#      closest analogue of its CPU kernels stands in: PyTorch's CPU ops on the same op definitions --
#      interpolate(mode="area") for tf.image.resize_area, conv2d(groups=19, padding=12) (oneDNN) for
#      DepthwiseConv2dNative, max_pool2d(3,1,1) + where for tf.nn.pool / tf.equal / tf.where.  tests/test_oracle.py
#      checks these ops against the C restatement (upsample and NMS bit for bit, smoothed map to ~5e-6).
#      The plain-C restatement (oracle/maps_oracle.c, direct 625-tap loop) is timed beside it; it is several times
#      slower and is NOT the headline.
#   B  the reference's OWN pafprocess.cpp, compiled by oracle/Makefile into oracle/_ref (loaded in the parent, so
#      the driver's library hook sees it; no fallback: a missing library raises).
#   C  PoseEstimator.estimate_paf's object loop over the getters (estimator.py:276-299).
# Throughput forms: (i) one worker process per host core, one torch thread each (frames are independent);
# (ii) one process, torch's intra-op threads over all cores (what a single sess.run does).  The headline CPU figure
# is the FASTER of the two.
_W = {}


def _cpu_worker_init(x, threads, stage_a):
    import torch
    torch.set_num_threads(int(threads))
    from oracle import oracle as O
    k2d = O.gauss_kernel()[0]
    _W["x"], _W["torch"], _W["O"], _W["stage_a"] = x, torch, O, stage_a
    _W["ref"] = O.ReferencePafprocess("asbuilt")            # raises when oracle/_ref is absent: no silent port
    _W["K"] = torch.from_numpy(np.ascontiguousarray(k2d))[None, None].repeat(19, 1, 1, 1)


def _cpu_pool_init(x, threads, stage_a):
    """Pool initializer: a worker whose set-up raises is respawned by multiprocessing for ever (the bench would hang
    instead of failing), so the error is kept and re-raised by the first task, which reaches the parent."""
    try:
        _cpu_worker_init(x, threads, stage_a)
    except BaseException as e:      # noqa: B902
        _W["init_error"] = "%s: %s" % (type(e).__name__, e)


def _stage_a_torch(x):
    """x[h,w,57] -> (peaks, heat_up, paf_up) float32 NHWC arrays, the three tensors of estimator.py:536-539."""
    torch = _W["torch"]
    F = torch.nn.functional
    h, w, _ = x.shape
    H, W = FACTOR * h, FACTOR * w
    with torch.no_grad():
        t = torch.from_numpy(x)[None].permute(0, 3, 1, 2)               # NHWC memory seen as NCHW (channels_last)
        up = F.interpolate(t, size=(H, W), mode="area")
        heat_up, paf_up = up[:, :19], up[:, 19:]
        g = F.conv2d(heat_up, _W["K"], padding=12, groups=19)
        mx = F.max_pool2d(g, 3, 1, 1)                                   # pads with -inf == clipped border windows
        peaks = torch.where(g == mx, g, torch.zeros_like(g))
        to_nhwc = lambda a: np.ascontiguousarray(a[0].permute(1, 2, 0).numpy())
        return to_nhwc(peaks), to_nhwc(heat_up), to_nhwc(paf_up)


def _stage_a_c(x):
    h, w, _ = x.shape
    peaks, heat_up, paf_up, _g = _W["O"].dense_maps(x, FACTOR * h, FACTOR * w)
    return peaks, heat_up, paf_up


def _stage_c_objects(ref, H, W):
    """estimator.py:276-299 statement for statement, over the reference binary's getters."""
    from tfpose_b200.humans import BodyPart, Human
    humans = []
    for human_id in range(ref.get_num_humans()):
        human = Human([])
        is_added = False
        for part_idx in range(18):
            c_idx = int(ref.get_part_cid(human_id, part_idx))
            if c_idx < 0:
                continue
            is_added = True
            human.body_parts[part_idx] = BodyPart('%d-%d' % (human_id, part_idx), part_idx,
                                                  float(ref.get_part_x(c_idx)) / W, float(ref.get_part_y(c_idx)) / H,
                                                  ref.get_part_score(c_idx))
        if is_added:
            human.score = ref.get_score(human_id)
            humans.append(human)
    return humans


def _cpu_decode_one(i):
    """One frame, all three stages; returns (#humans, seconds in A, B, C)."""
    if "init_error" in _W:
        raise RuntimeError("CPU-arm worker failed to initialise: " + _W["init_error"])
    x = _W["x"][i % len(_W["x"])]
    t0 = time.perf_counter()
    peaks, heat_up, paf_up = (_stage_a_torch if _W["stage_a"] == "torch" else _stage_a_c)(x)
    t1 = time.perf_counter()
    _W["ref"].process_paf(peaks, heat_up, paf_up)
    t2 = time.perf_counter()
    humans = _stage_c_objects(_W["ref"], heat_up.shape[0], heat_up.shape[1])
    t3 = time.perf_counter()
    return len(humans), t1 - t0, t2 - t1, t3 - t2


def _run_steps(mapper, frames_per_step, steps, warmup):
    for s in range(warmup):
        mapper(range(frames_per_step))
    t0 = time.perf_counter()
    acc = np.zeros(4)
    for s in range(steps):
        for r in mapper(range(s * frames_per_step, (s + 1) * frames_per_step)):
            acc += np.array(r)
    dt = time.perf_counter() - t0
    return frames_per_step * steps / dt, dt, acc


def cpu_arm(x, frames_per_step, steps, warmup, stage_a="torch", forms=("procs", "intraop"), detail=True):
    """The CPU decode of the frames x[n,h,w,57] (cycled), `steps` steps of `frames_per_step` frames per form.
    Returns the dict that goes under `cpu_baseline` (headline = the fastest form)."""
    import multiprocessing as mp
    import torch
    from oracle import oracle as O
    O.build()
    O.ReferencePafprocess("asbuilt")        # in the parent too: fail here, loudly, if oracle/_ref is missing
    cores = host_cores()
    out = {"unit": "frames/s", "kind": "port",
           "parts": {"stage_a_maps": "port: PyTorch CPU ops on the TF op definitions (TensorFlow unavailable on this box)"
                                     if stage_a == "torch" else "port: C restatement oracle/maps_oracle.c",
                     "stage_b_pafprocess": "reference: tf_pose/pafprocess/pafprocess.cpp compiled into oracle/_ref",
                     "stage_c_objects": "reference statements (estimator.py:276-299) over the binary's getters"},
           "nproc": os.cpu_count(), "torch_num_threads_default": torch.get_num_threads()}
    best = None
    if "procs" in forms:
        ctx = mp.get_context("spawn")       # not fork: the parent may hold CUDA and OpenMP state
        with ctx.Pool(cores, initializer=_cpu_pool_init, initargs=(x, 1, stage_a)) as pool:
            rate, dt, acc = _run_steps(lambda rng: pool.map(_cpu_decode_one, rng, chunksize=1), frames_per_step, steps,
                                       warmup)
        n = frames_per_step * steps
        out["form_worker_per_core"] = {"value": rate, "processes": cores, "torch_threads_per_process": 1, "seconds": dt,
                                       "frames": n, "ms_per_frame_in_worker": {"stage_a": acc[1] / n * 1e3,
                                                                               "stage_b": acc[2] / n * 1e3,
                                                                               "stage_c": acc[3] / n * 1e3}}
        best = ("one worker process per core, one torch thread each", rate, cores, n, dt, int(acc[0]))
    if "intraop" in forms:
        _cpu_worker_init(x, cores, stage_a)
        fps = max(1, min(frames_per_step, 16))
        rate, dt, acc = _run_steps(lambda rng: [_cpu_decode_one(i) for i in rng], fps, max(1, steps), min(warmup, 1))
        n = fps * max(1, steps)
        out["form_one_process_all_cores"] = {"value": rate, "torch_threads": cores, "seconds": dt, "frames": n,
                                             "ms_per_frame": {"stage_a": acc[1] / n * 1e3, "stage_b": acc[2] / n * 1e3,
                                                              "stage_c": acc[3] / n * 1e3}}
        if best is None or rate > best[1]:
            best = ("one process, torch intra-op threads over all cores", rate, cores, n, dt, int(acc[0]))
    if detail:
        # per-stage cost of one frame on ONE thread, each stage-A implementation
        k = min(len(x), 4)
        _cpu_worker_init(x, 1, "torch")
        t_torch1 = min(_timeit(lambda i=i: _stage_a_torch(x[i])) for i in range(k))
        _cpu_worker_init(x, cores, "torch")
        t_torch_all = min(_timeit(lambda i=i: _stage_a_torch(x[i])) for i in range(k))
        t_c = _timeit(lambda: _stage_a_c(x[0]))
        out["stage_a_torch"] = {"threads_1_ms": t_torch1 * 1e3, "threads_all_ms": t_torch_all * 1e3, "threads_all": cores}
        out["stage_a_c_oracle_1_thread_ms"] = t_c * 1e3
        pk, hu, pu = _stage_a_torch(x[0])
        out["stage_b_pafprocess_ms"] = _timeit(lambda: _W["ref"].process_paf(pk, hu, pu)) * 1e3
        _W["ref"].process_paf(pk, hu, pu)
        out["stage_c_objects_ms"] = _timeit(lambda: _stage_c_objects(_W["ref"], hu.shape[0], hu.shape[1])) * 1e3
    form, rate, used, n, dt, humans = best
    out.update({"value": rate, "cores": used, "humans_decoded": humans,
                "sample": "%d frames of the workload's synthetic batch in %.1f s; fastest CPU form: %s" % (n, dt, form)})
    return out


def _timeit(fn, reps=3):
    fn()
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best
