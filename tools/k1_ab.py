"""A/B timing of the fused map kernel (K1) between two builds of the library on the same box.

usage: python tools/k1_ab.py [libA.so libB.so ...]      (default: tools/ab/libtfpose_decode_base.so and the in-tree build)

Each library runs in its own process (TFPOSE_DECODE_LIB selects it).  Workload = the bench's: 64 frames per call,
4 rotating input batches (inputs + 773 MB of outputs per call exceed the L2), K1 timed alone with CUDA events
(tfpd_last_stage_ms), dense and minimal mode, for the BASELINE configs' shapes.  Minimal mode is timed with the
early-out on and off when the build knows the option.
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import json, os, sys
import numpy as np, torch
sys.path.insert(0, %(root)r)
import tfpose_b200 as tp
from tfpose_b200 import synth
out = {}
for name, h, w, cfg, people, caps in (("c2", 46, 54, 2, None, {}), ("c4", 46, 54, 4, 30, {"max_peaks": 4096, "max_cands": 8192, "max_humans": 128}),
                                      ("c3", 46, 82, 3, None, {}), ("c5", 92, 164, 5, None, {})):
    B, H, W = 64, 4 * h, 4 * w
    kw = {"people": people} if people else {}
    xs = []
    for s in range(4):
        base = synth.synth_batch(8, h, w, config_id=cfg, first_frame=8 * s, **kw)
        xs.append(torch.from_numpy(np.concatenate([base] * 8)).cuda())
    dense = tuple(torch.empty((B, H, W, c), device="cuda") for c in (19, 38, 19))
    dec = tp.PoseDecoder(max_batch=B, max_upsample=(H, W), **caps)
    dec.set_profiling(True)
    rec = {}
    for ctas in (128, 0):
        dec.set_option("k1_ctas", ctas)
        for mode in ("dense", "minimal", "minimal_no_early_out"):
            if mode == "minimal_no_early_out":
                try:
                    dec.set_option("k1_early_out", 0)
                except Exception:
                    continue
            t, humans = [], 0
            for i in range(13):
                res = dec.decode(xs[i %% 4], (H, W), dense=dense if mode == "dense" else None)
                if i >= 3:
                    t.append(dec.stage_ms()[0])
                    humans += int(res.info["n_humans"].sum())
            rec["%%s_ctas%%d" %% (mode, ctas or 148)] = {"k1_us": float(np.mean(t)) * 1e3, "k1_us_min": float(np.min(t)) * 1e3, "humans": humans}
            if mode == "minimal_no_early_out":
                dec.set_option("k1_early_out", 1)
    dec.set_option("k1_ctas", 0)
    for band in (2, 4, 8, 16):
        try:
            dec.set_option("k1_min_band", band)
        except Exception:
            break
        t = []
        for i in range(13):
            dec.decode(xs[i %% 4], (H, W))
            if i >= 3:
                t.append(dec.stage_ms()[0])
        rec["minimal_band%%d_ctas148" %% band] = {"k1_us": float(np.mean(t)) * 1e3, "k1_us_min": float(np.min(t)) * 1e3, "humans": 0}
    rec["dense_gbs_ctas128"] = 4 * (57 * h * w + 76 * H * W) * B / (rec["dense_ctas128"]["k1_us"] * 1e-6) / 1e9
    out[name] = rec
    dec.close()
    del dense, xs
    torch.cuda.empty_cache()
print("RESULT " + json.dumps(out))
'''


def main():
    libs = sys.argv[1:] or [os.path.join(ROOT, "tools", "ab", "libtfpose_decode_base.so"),
                            os.path.join(ROOT, "tf-pose-estimation_b200", "libtfpose_decode.so")]
    results = {}
    for lib in libs:
        env = dict(os.environ, TFPOSE_DECODE_LIB=lib)
        p = subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT}], env=env, capture_output=True, text=True)
        line = [ln for ln in p.stdout.splitlines() if ln.startswith("RESULT ")]
        if not line:
            print("FAILED", lib, p.stderr[-2000:])
            continue
        results[lib] = json.loads(line[0][7:])
    for cfg in ("c2", "c4", "c3", "c5"):
        for key in ("dense_ctas128", "dense_ctas148", "minimal_ctas128", "minimal_ctas148", "minimal_no_early_out_ctas148",
                    "minimal_band2_ctas148", "minimal_band4_ctas148", "minimal_band8_ctas148", "minimal_band16_ctas148"):
            row = []
            for lib in libs:
                r = results.get(lib, {}).get(cfg, {}).get(key)
                row.append("%8.1f us (humans %d)" % (r["k1_us"], r["humans"]) if r else "       n/a")
            print("%s %-30s %s" % (cfg, key, "   |   ".join(row)))
        print("%s %-30s %s" % (cfg, "dense GB/s @128 CTAs", "   |   ".join(
            "%8.0f" % results[lib][cfg]["dense_gbs_ctas128"] for lib in libs if lib in results)))
    print("libs:", libs)
    print("JSON " + json.dumps(results))


if __name__ == "__main__":
    main()
