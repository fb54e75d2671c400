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
    # ~30 people: <= 385 peaks, 1221 candidates, 41 subset rows, 25 humans per frame (oracle, 12 frames): the default
    # table capacities (2048 / 4096 / 512 / 64) hold them with a 3-5x margin; an overflow would fail the fetch loudly
    "c4": {"id": 4, "h": 46, "w": 54, "people": 30, "batch": 64, "caps": {},
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


def cpu_arm(x, frames_per_step, steps, warmup, stage_a="torch", forms=("procs", "intraop"), detail=True,
            intraop_cap=16):
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
        fps = max(1, min(frames_per_step, intraop_cap or frames_per_step))
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


# ------------------------------------------------------------------------------------------ GPU arm
class Job(object):
    """One rank of the job: device, process group, and the reductions the timing needs."""

    def __init__(self, rank, world, local_rank):
        import torch
        self.torch, self.rank, self.world, self.local_rank = torch, rank, world, local_rank
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        self.dist = None
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        self.stream = torch.cuda.current_stream(self.dev)
        self.ev0 = torch.cuda.Event(enable_timing=True)
        self.ev1 = torch.cuda.Event(enable_timing=True)

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, v):
        if self.dist is None:
            return float(v)
        t = self.torch.tensor([v], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, v):
        if self.dist is None:
            return float(v)
        t = self.torch.tensor([v], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def all_ranks(self, values):
        """[world][len(values)] float64 on every rank (a few numbers per rank: never on the data path)."""
        t = self.torch.tensor([list(values)], device=self.dev, dtype=self.torch.float64)
        if self.dist is None:
            return t.cpu().numpy()
        out = self.torch.zeros((self.world, len(values)), device=self.dev, dtype=self.torch.float64)
        self.dist.all_gather_into_tensor(out, t)
        return out.cpu().numpy()

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def make_inputs(cfg, B, rank, base_frames):
    """N_INPUT_SETS different batches of B frames for this rank: `base_frames` distinct synthetic frames per set
    (frame seeds never repeat across sets or ranks), tiled up to B."""
    from tfpose_b200 import synth
    sets = []
    for sidx in range(N_INPUT_SETS):
        nb = min(B, base_frames)
        kw = {"people": cfg["people"]} if cfg["people"] else {}
        base = synth.synth_batch(nb, cfg["h"], cfg["w"], config_id=cfg["id"],
                                 first_frame=(rank * N_INPUT_SETS + sidx) * base_frames, **kw)
        reps = (B + nb - 1) // nb
        sets.append(np.concatenate([base] * reps)[:B].copy())
    return sets


def timed_region(job, sampler, fn):
    """barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks (ms); also this
    rank's own time."""
    job.barrier()
    t0 = time.time()
    job.ev0.record(job.stream)
    fn()
    job.ev1.record(job.stream)
    job.barrier()
    if sampler is not None:
        sampler.window(t0, time.time())
    mine = job.ev0.elapsed_time(job.ev1)
    return job.max_over_ranks(mine), mine


def measure_device(job, key, args, sampler, steps, base_frames):
    """Device-resident throughput of one config on this job: pipelined dense (= value), one call at a time (per-kernel
    times, K1 roofline), and both again in minimal mode."""
    import tfpose_b200 as tp
    torch, dev = job.torch, job.dev
    cfg = CONFIGS[key]
    B = frames_per_rank(cfg, job.world, args.batch if key == args.config else None)
    h, w = cfg["h"], cfg["w"]
    H, W = FACTOR * h, FACTOR * w
    sets_host = make_inputs(cfg, B, job.rank, base_frames)
    x_dev = [torch.from_numpy(s).to(dev) for s in sets_host]
    n_pipe = max(1, args.pipeline)
    n_pipe_min = max(1, args.pipeline_minimal)          # minimal mode: the map kernel is short, a deeper queue pays
    mk_dense = lambda: (torch.empty((B, H, W, 19), device=dev), torch.empty((B, H, W, 38), device=dev),
                        torch.empty((B, H, W, 19), device=dev))
    decs = [tp.PoseDecoder(max_batch=B, max_upsample=(H, W), device=job.local_rank, **cfg["caps"])
            for _ in range(max(n_pipe, n_pipe_min))]
    denses = [mk_dense() for _ in range(n_pipe)]
    for d in decs:
        d.set_profiling(True)
        d.set_option("k1_ctas", args.k1_ctas)
    dec, dense = decs[0], denses[0]
    warm = max(args.warmup, 3)

    # ---- one call at a time (decode -> fetch): every kernel runs alone -> per-kernel times, K1 roofline ----
    def serial(with_dense, acc):
        for i in range(steps):
            dec.decode(x_dev[i % N_INPUT_SETS], (H, W), dense=dense if with_dense else None)
            if acc is not None:
                acc.append(dec.stage_ms())
    for i in range(warm):
        dec.decode(x_dev[i % N_INPUT_SETS], (H, W), dense=dense)
    st_dense = []
    ms_serial, _ = timed_region(job, sampler, lambda: serial(True, st_dense))
    for i in range(warm):
        dec.decode(x_dev[i % N_INPUT_SETS], (H, W))
    st_min = []
    ms_serial_min, _ = timed_region(job, sampler, lambda: serial(False, st_min))
    st_dense, st_min = np.array(st_dense), np.array(st_min)

    # ---- pipelined: n_pipe contexts round-robin; later steps are enqueued before step i's tables are fetched ----
    if int(os.environ.get("K1_TURNSTILE", "1")):
        for d in decs:
            d.set_option("k1_turnstile", 1)
    n_max = len(decs)
    streams = [job.stream] + [torch.cuda.Stream(dev) for _ in range(n_max - 1)] if args.pipe_streams > 1 \
        else [job.stream] * n_max
    for d in decs:
        d.set_profiling(False)      # the per-stage events are only needed in the one-call-at-a-time regions above
    stats = {"humans": 0, "near": 0, "launches": 0, "enqueue_s": 0.0, "fetch_s": 0.0, "peaks": 0}

    def collect(d, count):
        t0 = time.perf_counter()
        res = d.fetch(reuse=True)
        stats["fetch_s"] += time.perf_counter() - t0
        if count:
            stats["launches"] += d.launch_count()
            stats["humans"] += int(res.info["n_humans"].sum())
            stats["near"] += int(res.info["near_ties"].sum())
            stats["peaks"] += int(res.info["n_peaks"].sum())

    def pipelined(nsteps, count, with_dense):
        depth = n_pipe if with_dense else n_pipe_min
        inflight = []
        for i in range(nsteps):
            j = i % depth
            if len(inflight) == depth:
                collect(inflight.pop(0), count)
            t0 = time.perf_counter()
            decs[j].decode_async(x_dev[i % N_INPUT_SETS], (H, W), dense=denses[j] if with_dense else None,
                                 stream=streams[j])
            stats["enqueue_s"] += time.perf_counter() - t0
            inflight.append(decs[j])
        for d in inflight:
            collect(d, count)
        for st_ in streams[1:]:
            job.stream.wait_stream(st_)

    pipelined(max(warm, n_pipe), False, True)
    stats.update(enqueue_s=0.0, fetch_s=0.0)
    ms_pipe, ms_pipe_mine = timed_region(job, sampler, lambda: pipelined(steps, True, True))
    host = (stats["enqueue_s"] / steps * 1e3, stats["fetch_s"] / steps * 1e3)
    pipelined(max(warm, n_pipe_min), False, False)
    ms_pipe_min, _ = timed_region(job, sampler, lambda: pipelined(steps, False, False))

    frames_job = job.sum_over_ranks(B)                     # frames of the whole job per step
    per_rank = job.all_ranks([ms_pipe_mine / steps, float(st_dense[:, 0].mean()), host[0], host[1], B])
    humans_job = job.sum_over_ranks(stats["humans"])
    for d in decs:
        d.close()
    del denses, dense, x_dev
    torch.cuda.empty_cache()

    peak, peak_src = measured_peak_gbs()
    k1 = float(st_dense[:, 0].mean())
    alg = dense_bytes_per_frame(h, w) * B
    k1_min = float(st_min[:, 0].mean())
    flops = polyphase_flops_per_frame(h, w) * B
    out = {
        "workload": cfg["name"] % B, "frames_per_gpu": B, "frames_per_step_job": int(frames_job),
        "scaling": scaling_of(cfg), "steps": steps,
        "value": frames_job * steps / (ms_pipe * 1e-3), "ms_per_step": ms_pipe / steps,
        "value_one_call_at_a_time": frames_job * steps / (ms_serial * 1e-3),
        "value_minimal_mode_pipelined": frames_job * steps / (ms_pipe_min * 1e-3),
        "value_minimal_mode": frames_job * steps / (ms_serial_min * 1e-3),
        "stage_ms": {"k1_maps": k1, "peak_sort": float(st_dense[:, 1].mean()), "k2_limbs": float(st_dense[:, 2].mean()),
                     "k3_assemble": float(st_dense[:, 3].mean())},
        "stage_ms_minimal": {"k1_maps": k1_min, "k2_limbs": float(st_min[:, 2].mean()),
                             "k3_assemble": float(st_min[:, 3].mean())},
        "roofline": {"bound": "hbm", "kernel": "k1_fused_f4 (upsample+smoother+NMS, dense)",
                     "achieved": alg / (k1 * 1e-3) / 1e9, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                     "frac": alg / (k1 * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_launch": alg, "kernel_ms": k1,
                     "traffic": profile_constant("k1_traffic.json", "dram_bytes_per_launch") if key == "c2" and B == 64 else None},
        "roofline_minimal": {"bound": "fp32", "kernel": "k1_fused_f4 (smoother+NMS only, nothing dense written)",
                             "achieved_tflops": flops / (k1_min * 1e-3) / 1e12, "algorithmic_flops_per_launch": flops,
                             "kernel_ms": k1_min, "note": "polyphase form: 2*(7 + 7/4) flop per output pixel of 19 channels",
                             "pipe_fma_pct": profile_constant("k1_minimal.json", "pipe_fma_pct") if key == "c2" else None},
        "per_rank": [{"rank": r, "ms_per_step": float(v[0]), "k1_ms": float(v[1]), "host_enqueue_ms_per_step": float(v[2]),
                      "host_fetch_wait_ms_per_step": float(v[3]), "frames": int(v[4])} for r, v in enumerate(per_rank)],
        "l2": "%d rotating input batches (%.0f MB each) and %.0f MB of dense outputs per step against the 126 MB L2"
              % (N_INPUT_SETS, B * h * w * 228 / 1e6, B * 76 * H * W * 4 / 1e6),
        "humans_decoded": int(humans_job), "peaks_per_frame": stats["peaks"] / max(1, steps * B),
        "near_tie_pixels": stats["near"], "gpu_launches": stats["launches"],
    }
    return out, sets_host


def measure_e2e(job, key, args, sampler, steps, sets_host):
    """End to end through the host-buffer C-ABI call: pinned host input -> results on the host, every step paying
    its own H2D and D2H inside the timed region (wall clock around synchronised regions, max over ranks)."""
    import tfpose_b200 as tp
    from tfpose_b200 import _lib
    torch, dev = job.torch, job.dev
    cfg = CONFIGS[key]
    B = len(sets_host[0])
    h, w = cfg["h"], cfg["w"]
    H, W = FACTOR * h, FACTOR * w
    xs = [torch.from_numpy(s).pin_memory().numpy() for s in sets_host]
    decs = [tp.PoseDecoder(max_batch=B, max_upsample=(H, W), device=job.local_rank, **cfg["caps"]) for _ in range(2)]
    frames_job = job.sum_over_ranks(B)

    # two contexts: the pinned input of step i+1 crosses PCIe (tfpd_decode_host_async) while step i is decoded and
    # its tables come back (fetch)
    def run(nsteps, objects):
        inflight, last, n_obj = [], None, 0
        for i in range(nsteps):
            d = decs[i % 2]
            if len(inflight) == 2:
                last = inflight.pop(0).fetch(reuse=True)
                if objects:
                    n_obj += sum(len(f) for f in last.all_humans())
            d.decode_host_async(xs[i % N_INPUT_SETS], (H, W))
            inflight.append(d)
        for d in inflight:
            last = d.fetch(reuse=True)
            if objects:
                n_obj += sum(len(f) for f in last.all_humans())
        return last, n_obj

    def wall(fn):
        job.barrier()
        t0 = time.time()
        r = fn()
        torch.cuda.synchronize(dev)
        dt = time.time() - t0
        if sampler is not None:
            sampler.window(t0, t0 + dt)
        ms = job.max_over_ranks(dt * 1e3)
        job.barrier()
        return ms, r

    run(max(args.warmup, 3), False)
    ms, (last, _) = wall(lambda: run(steps, False))
    out = {"e2e": {"value": frames_job * steps / (ms * 1e-3), "unit": "frames/s",
                   "h2d_bytes_per_step": B * h * w * 57 * 4,
                   "d2h_bytes_per_step": B * _lib.INFO_DTYPE.itemsize + B * decs[0].max_humans * _lib.HUMAN_DTYPE.itemsize,
                   "call": "PoseDecoder.decode_host_async + fetch (tfpd_decode_host_async / tfpd_fetch_humans): pinned "
                           "host input -> human tables on host, two contexts alternating; heatMat / pafMat stay on the "
                           "device (inference() also returns them to the host: that variant is e2e_dense_maps)"},
           "e2e_humans_last_step": int(last.info["n_humans"].sum())}
    # the same, plus the list[list[Human]] the reference's inference() returns (built by the _objbuild extension)
    run(max(args.warmup, 3), True)
    ms, (_, n_obj) = wall(lambda: run(steps, True))
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        hs = last.all_humans()
    obj_s = (time.perf_counter() - t0) / reps
    out["e2e_objects"] = {"value": frames_job * steps / (ms * 1e-3), "unit": "frames/s",
                          "humans_built": int(n_obj), "call": "e2e + DecodeResult.all_humans(): host input -> list[list[Human]]",
                          "objects_only_frames_per_s_one_core": B / obj_s,
                          "objects_only_ms_per_step": obj_s * 1e3, "humans_per_step": sum(len(f) for f in hs)}
    decs[1].close()
    if args.e2e_dense:
        ksteps = max(2, min(steps, 5))
        out_pin = tuple(torch.empty((B, H, W, c), dtype=torch.float32).pin_memory().numpy() for c in (19, 38, 19))
        decs[0].decode_host(xs[0], (H, W), out=out_pin)
        ms, _ = wall(lambda: [decs[0].decode_host(xs[i % N_INPUT_SETS], (H, W), out=out_pin) for i in range(ksteps)])
        out["e2e_dense_maps"] = {"value": frames_job * ksteps / (ms * 1e-3), "unit": "frames/s",
                                 "d2h_bytes_per_step": B * 76 * H * W * 4,
                                 "call": "decode_host with the three dense maps copied back into page-locked arrays"}
    decs[0].close()
    return out


def run_ours(args, rank, world, local_rank):
    import tfpose_b200  # noqa: F401  (fails loudly when libtfpose_decode.so is not built: no CPU path)
    from tfpose_b200 import shard
    # before any host buffer exists: pinned inputs on the GPU's own NUMA node (matters with 8 ranks on two sockets)
    numa = shard.bind_host_to_gpu_node(local_rank) if not args.no_numa else {"node": -1, "mempolicy": False, "cpus": 0}
    job = Job(rank, world, local_rank)
    sampler = ClockSampler(list(range(world)), float(os.environ.get("BENCH_SAMPLE_PERIOD", "0.002"))) if rank == 0 else None
    # the timed regions are a few milliseconds long: keep the collector's pauses out of them
    gc.collect()
    gc.disable()
    key = args.config
    main, sets_host = measure_device(job, key, args, sampler, args.steps, 16)
    e2e = measure_e2e(job, key, args, sampler, args.steps, sets_host) if not args.no_e2e else {}
    per_config = {}
    if args.per_config:
        for other in ("c2", "c3", "c4", "c5"):
            if other == key:
                continue
            res, _ = measure_device(job, other, args, sampler, max(5, min(args.steps, 20)), 8)
            keep = ("workload", "frames_per_gpu", "frames_per_step_job", "scaling", "steps", "value", "ms_per_step",
                    "value_one_call_at_a_time", "value_minimal_mode_pipelined", "stage_ms", "per_rank",
                    "humans_decoded", "peaks_per_frame", "l2")
            per_config[other] = {k: res[k] for k in keep}
            per_config[other]["k1_roofline_frac"] = res["roofline"]["frac"]
            per_config[other]["k1_achieved_gbs"] = res["roofline"]["achieved"]
    gc.enable()
    clocks = sampler.result() if sampler is not None else None

    cpu_base = None
    shard.unbind_host(numa)          # the CPU leg below uses every host core
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_base = cpu_arm(sets_host[0][:32], 32, 2, 1)

    if rank == 0:
        cfg = CONFIGS[key]
        line = {
            "metric": "decoded_frames_per_s", "value": main["value"], "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": main["ms_per_step"],
            "higher_is_better": True, "scaling": main["scaling"], "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": main["workload"], "config": key, "batch_per_gpu": main["frames_per_gpu"],
                       "global_batch": main["frames_per_step_job"], "sharding": "by image, no collective",
                       "host_numa": {k: numa[k] for k in ("node", "mempolicy", "cpus")}, "l2": main["l2"],
                       "mode": "asbuilt",
                       "pipeline_minimal_contexts": args.pipeline_minimal,
                       "pipeline": "%d contexts round-robin on %s stream(s): later steps are enqueued before step i's "
                                   "tables are fetched; the HBM-bound map kernel runs on %s so that the latency-bound "
                                   "scoring/assembly kernels of the previous batches overlap it; roofline/stage_ms from "
                                   "a second K-step region run one call at a time with the same settings"
                                   % (args.pipeline, "%d" % args.pipeline if args.pipe_streams > 1 else "1",
                                      ("%d CTAs" % args.k1_ctas) if args.k1_ctas else "all SMs")},
            "roofline": main["roofline"], "roofline_minimal": main["roofline_minimal"],
            "stage_ms": main["stage_ms"], "stage_ms_minimal": main["stage_ms_minimal"],
            "value_one_call_at_a_time": main["value_one_call_at_a_time"],
            "value_minimal_mode": main["value_minimal_mode"],
            "value_minimal_mode_pipelined": main["value_minimal_mode_pipelined"],
            "per_rank": main["per_rank"],
            "gpu_launches": main["gpu_launches"], "clocks": clocks, "humans_decoded": main["humans_decoded"],
            "near_tie_pixels": main["near_tie_pixels"],
        }
        line.update(e2e)
        if "e2e" not in line:
            line["e2e"] = None
        if per_config:
            line["per_config"] = per_config
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        emit(line)
    job.close()


def run_reference(args, rank, world):
    """The reference's CPU decode on the box's host cores, same workload and frames per step as the repo arm
    (rank 0 alone; the other ranks leave at once)."""
    if rank != 0:
        return
    from tfpose_b200 import synth
    cfg = CONFIGS[args.config]
    B = frames_per_rank(cfg, 1 if "total" in cfg else world, args.batch)    # one host: the whole job's step
    kw = {"people": cfg["people"]} if cfg["people"] else {}
    x = synth.synth_batch(min(B, 64), cfg["h"], cfg["w"], config_id=cfg["id"], **kw)
    # which CPU form is faster on this host: one short step each, then the timed run with that one
    probe = cpu_arm(x, B, 1, 1, detail=False, intraop_cap=None)
    form = "procs" if probe["form_worker_per_core"]["value"] >= probe["form_one_process_all_cores"]["value"] else "intraop"
    res = cpu_arm(x, B, args.steps, args.warmup, forms=(form,), detail=True, intraop_cap=None)
    res["probe_frames_per_s"] = {"worker_per_core": probe["form_worker_per_core"]["value"],
                                 "one_process_all_cores": probe["form_one_process_all_cores"]["value"]}
    dt = (res.get("form_worker_per_core") or res.get("form_one_process_all_cores"))["seconds"]
    line = {
        "impl": "reference", "metric": "decoded_frames_per_s", "value": res["value"], "unit": "frames/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": scaling_of(cfg), "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["name"] % frames_per_rank(cfg, world, args.batch), "config": args.config,
                   "batch_per_gpu": frames_per_rank(cfg, world, args.batch), "frames_per_step": B,
                   "sharding": "by image, no collective", "mode": "asbuilt"},
        "cpu_baseline": res,
        "e2e": {"value": res["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "humans_decoded": res["humans_decoded"],
    }
    emit(line)


_REAL_STDOUT = None


def quiet_stdout():
    """stdout carries exactly one JSON line: everything libraries print there (NCCL's version banner comes from C
    code, on file descriptor 1) is sent to stderr; emit() writes the line to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS), help="BASELINE.json config the line is quoted on")
    ap.add_argument("--batch", type=int, default=0, help="frames per GPU of --config (0 = the config's own)")
    ap.add_argument("--no-per-config", dest="per_config", action="store_false", default=True,
                    help="skip the short runs of the other configs (per_config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-numa", action="store_true", help="leave host memory policy and CPU affinity alone")
    ap.add_argument("--pipeline", type=int, default=3, help="contexts in flight in the device-resident timed region")
    ap.add_argument("--pipeline-minimal", type=int, default=4, help="contexts in flight in the minimal-mode pipelined region")
    ap.add_argument("--k1-ctas", type=int, default=128, help="CTAs of the fused map kernel (0 = one per SM)")
    ap.add_argument("--pipe-streams", type=int, default=2, help="1: all contexts on one stream, 2: one stream each")
    ap.add_argument("--no-e2e-dense", dest="e2e_dense", action="store_false", default=True)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
