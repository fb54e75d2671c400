"""Shared comparison helpers of the parity tests (CUDA path vs oracle).  The oracle is the checker
here, never the thing under test."""
import numpy as np

MAP_TOL = 1e-5      # north_star: upsampled/blurred maps within 1e-5 relative error


def humans_equal(got, want, H, W):
    """got: list[Human] from the CUDA path; want: oracle/reference tuples [(score, {part: (x, y, s)})].
    Bit-exact: float32 scores, integer pixel coordinates (through x/W, y/H like estimate_paf)."""
    if len(got) != len(want):
        return False, "human count %d != %d" % (len(got), len(want))
    for i, (hg, (score, parts)) in enumerate(zip(got, want)):
        if np.float32(hg.score) != np.float32(score):
            return False, "human %d score %r != %r" % (i, hg.score, score)
        if list(hg.body_parts) != sorted(parts):
            return False, "human %d parts %r != %r" % (i, list(hg.body_parts), sorted(parts))
        for k, (px, py, ps) in parts.items():
            bp = hg.body_parts[k]
            if bp.x != px / W or bp.y != py / H or np.float32(bp.score) != np.float32(ps):
                return False, "human %d part %d (%r,%r,%r) != (%r,%r,%r)" % (i, k, bp.x, bp.y, bp.score, px / W,
                                                                            py / H, ps)
            if bp.uidx != "%d-%d" % (i, k) or bp.part_idx != k:
                return False, "human %d part %d uidx %r" % (i, k, bp.uidx)
    return True, ""


def peaks_equal(pk, R):
    """pk: PoseDecoder.peaks() dict; R: oracle PafResult."""
    return (len(pk["x"]) == len(R.peaks["x"]) and np.array_equal(pk["part"], R.peaks["part"])
            and np.array_equal(pk["x"], R.peaks["x"]) and np.array_equal(pk["y"], R.peaks["y"])
            and np.array_equal(pk["score"], R.peaks["score"]))


def conns_equal(conns, R):
    """conns: PoseDecoder.connections() dict; R: oracle PafResult (limb-major acceptance order)."""
    want = [(l, int(a), int(b), np.float32(s)) for l in range(19) for a, b, s in zip(*R.conns[l])]
    got = list(zip(conns["limb"].tolist(), conns["cid1"].tolist(), conns["cid2"].tolist(),
                   [np.float32(s) for s in conns["score"]]))
    return got == want


def rel_err(a, b, scale):
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max() / max(float(np.abs(scale).max()), 1e-30))


def table_tuples(res, frame):
    """The raw tfpd_human records of a frame as comparable tuples."""
    n = int(res.info["n_humans"][frame])
    out = []
    for r in res.table[frame][:n]:
        out.append((np.float32(r["score"]).tobytes(), r["cid"].tobytes(), r["x"].tobytes(), r["y"].tobytes(),
                    r["part_score"].tobytes()))
    return out


def check_peaks_tensor(oracle, gp, peaks, g, fragile, what):
    """Dense peaks tensor vs oracle: values within MAP_TOL on the common support; a pixel may be a local
    maximum on one side only where the oracle's own NMS decision is a near-tie at the map tolerance
    (flat noise floor, the never-read background channel, x4-replication plateaus) -- and never on a
    pixel that matters downstream unless the frame is flagged fragile."""
    both = (gp != 0) & (peaks != 0)
    assert rel_err(np.where(both, gp, 0), np.where(both, peaks, 0), g) <= MAP_TOL, what
    mism = (gp != 0) != (peaks != 0)
    if mism.any():
        _, margin = oracle.nms(g, want_margin=True)
        assert (np.abs(margin[mism]) <= 2 * MAP_TOL * float(np.abs(g).max())).all(), \
            "%s: peaks-tensor support differs at a pixel that is not a near-tie" % what
        if fragile == 0:
            assert not (mism[..., :18] & (g[..., :18] > 0.05 * (1 - 4e-5))).any(), \
                "%s: support differs above the heat threshold" % what


_TIE_CHAINS = [(1, 2, 3, 4), (1, 5, 6, 7), (1, 8, 9, 10), (1, 11, 12, 13)]   # neck -> shoulder -> elbow -> wrist, neck -> hip -> knee -> ankle


def chained_tie_case(seed, H=160, W=120):
    """Dense (peaks, heat, paf) on which the reference's result DEPENDS on std::sort's order between bit-equal scores
    (pafprocess.cpp:97): two 4-part chains of the skeleton, each laid out on one horizontal line with every part's peaks
    in its own x range and a constant PAF (1, 0) on the three limbs' channel pairs along that line.  Every candidate of
    a limb then has direction (1, 0) and -- the map is tall, so the length penalty of :83 vanishes -- the same score:
    25..49 bit-equal candidates that compete for the same peaks, more than std::sort's 16-element insertion-sort
    (stable) range.  Humans have 4 parts, so they pass the reference's compiled-in part-count filter."""
    from tfpose_b200.common import CocoPairs, CocoPairsNetwork
    rng = np.random.default_rng(seed)
    peaks = np.zeros((H, W, 19), np.float32)
    heat = np.zeros((H, W, 19), np.float32)
    paf = np.zeros((H, W, 38), np.float32)
    y = 8
    for chain in [_TIE_CHAINS[i] for i in rng.permutation(4)[:2]]:
        for k, part in enumerate(chain):
            n = int(rng.integers(5, 8))
            xs = np.sort(rng.choice(np.arange(2 + 30 * k, 28 + 30 * k), n, replace=False))
            for x in xs:
                v = np.float32(rng.integers(2, 5) / 4.0)
                peaks[y, x, part] = v
                heat[y, x, part] = v
        for a, b in zip(chain[:-1], chain[1:]):
            limb = [i for i, p in enumerate(CocoPairs) if tuple(p) == (a, b)][0]
            paf[y, :, CocoPairsNetwork[limb][0]] = 1.0
        y += 40
    return peaks, heat, paf
