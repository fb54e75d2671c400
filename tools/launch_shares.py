"""Per-kernel time and share of a decode step from an ncu launch list
(`ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file X.csv <command>`), next to the bench's own
event-timed stage shares when a bench JSON line is given.  ncu's per-launch times are cold-cache and serialised: the
SHARES are what must agree, not the absolute times.
usage: python tools/launch_shares.py launches.csv [bench.json]"""
import collections
import csv
import json
import sys


def load(path):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr, agg = None, collections.OrderedDict()
    for r in rows:
        if r and r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            if d.get("Metric Name") == "gpu__time_duration.sum":
                v = float(d["Metric Value"].replace(",", ""))
                v *= {"nsecond": 1e-3, "ns": 1e-3, "usecond": 1.0, "us": 1.0, "msecond": 1e3, "ms": 1e3}.get(d["Metric Unit"], 1.0)
                agg.setdefault(d["Kernel Name"], []).append(v)
    return agg


def short(name):
    for key, label in (("k1_fused_f4<4, 1>", "K1 maps (dense)"), ("k1_fused_f4<4, 0>", "K1 maps (minimal)"),
                       ("k_limb_score", "K2 limb scoring"), ("k_assemble", "K3 assembly"), ("k_sort_peaks", "peak sort")):
        if key in name:
            return label
    return name[:40]


def main():
    agg = load(sys.argv[1])
    total = sum(sum(v) for v in agg.values())
    print("%-22s %9s %12s %8s" % ("kernel", "launches", "mean us", "share"))
    for k, v in agg.items():
        print("%-22s %9d %12.1f %7.1f%%" % (short(k), len(v), sum(v) / len(v), 100.0 * sum(v) / total))
    if len(sys.argv) > 2:
        st = json.load(open(sys.argv[2]))["stage_ms"]
        tot = st["k1_maps"] + st["k2_limbs"] + st["k3_assemble"]
        print("bench stage_ms (CUDA events, same workload): K1 %.1f us (%.1f%%), K2 %.1f us (%.1f%%), K3 %.1f us (%.1f%%)" % (
            st["k1_maps"] * 1e3, 100 * st["k1_maps"] / tot, st["k2_limbs"] * 1e3, 100 * st["k2_limbs"] / tot,
            st["k3_assemble"] * 1e3, 100 * st["k3_assemble"] / tot))


if __name__ == "__main__":
    main()
