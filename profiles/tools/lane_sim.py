"""CPU simulation of the column blocks of the tiled band kernels (build_items_range in csrc/hc_gpu.cu) over the ORACLE's per-pair
column counts: how many lane-columns of the warp steps are used, and where the rest goes (lanes without a walked pair, lanes
stopped at an indel mismatch, overlaps shorter than the longest of the block), with and without the grouping of the column
reads by indel signature. Test / measurement infrastructure (it runs the oracle); nothing here is on the product path.
  python profiles/tools/lane_sim.py C2            # 20 groups of 24 rows over the genome, window 288
  python profiles/tools/lane_sim.py C4 200000     # 6 groups of 16 rows, window 2912"""
import collections
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench                                     # noqa: E402
from oracle import pyoracle                      # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "C2"
n_frag = int(sys.argv[2]) if len(sys.argv) > 2 else None
b, cfg = bench.generate(name, n_frag, workers=8)
kw = bench.calc_kwargs(b.max_position(), "defaults + no_prob0")
cov_first, cov_last, single = np.asarray(b.cov_first), np.asarray(b.cov_last), np.asarray(b.single_end) != 0
pmax = np.maximum.accumulate(cov_last)
orc = pyoracle.Oracle(b, **kw)
groups, rows_per, W = (20, 24, 288) if name == "C2" else (6, 16, 2912)


def signature(r):
    s = []
    for seg in range(2):
        k = 2 * r + seg
        n = int(b.cig_len[k])
        c = b.cig[int(b.cig_off[k]):int(b.cig_off[k]) + n].tobytes()
        if n == 0 or (b"I" not in c and b"D" not in c):
            continue
        pos = int(b.seg_start[k])
        for ch in c:
            if ch in b"SH":
                continue
            if ch == ord("M"):
                pos += 1
            elif ch == ord("D"):
                s.append((pos, "D")); pos += 1
            else:
                s.append((pos, "I"))
    return hash(tuple(s)) & 0x7fffffff if s else 0


cache = {}
for by_sig in (0, 1):
    loss = collections.Counter(); used = 0; longest = 0; steps = 0; gaps = 0; walked_n = 0
    for g in range(groups):
        r0 = int(b.n * (0.06 + 0.88 * g / groups))
        rows = np.arange(r0, r0 + rows_per)
        jlo = int(np.searchsorted(pmax, cov_first[rows[0]], side="left")) // 32 * 32
        ncol = rows[0] - jlo
        if g not in cache:
            out = orc.pairs(np.repeat(rows, ncol), np.tile(np.arange(jlo, rows[0]), len(rows)))
            cache[g] = (out, np.array([signature(r) for r in range(jlo, rows[0])]))
        out, sig = cache[g]
        st, cc, gp = (out[f].reshape(len(rows), -1) for f in ("status", "cc", "gap"))
        walked = st == 2
        gaps += int(((gp != 0) & walked).sum()); walked_n += int(walked.sum())
        for c0 in range(0, ncol, W):
            js = np.arange(c0, min(c0 + W, ncol))
            cnt = collections.Counter(sig[js])
            grp = np.array([0 if s == 0 else (s if cnt[s] >= 8 else 0xffffffff) for s in sig[js]]) if by_sig else np.zeros(len(js), np.int64)
            js = js[np.lexsort((js, cov_last[js + jlo], grp, ~single[js + jlo]))]
            for p in range(0, len(js), 32):
                blk = js[p:p + 32]
                reach = cov_first[rows] <= cov_last[blk + jlo].max()
                w = walked[:, blk][reach]
                sub = np.where(w, cc[:, blk][reach], 0)
                m = sub.max(1)
                stopped = (gp[:, blk][reach] != 0) & w
                used += int(sub.sum()); longest += int(m.sum()); steps += int(reach.sum())
                loss["no walked pair"] += int(((~w) * m[:, None]).sum() + (32 - len(blk)) * m.sum())
                loss["stopped at an indel mismatch"] += int((stopped * (m[:, None] - sub)).sum())
                loss["overlap shorter than the block's longest"] += int(((w & ~stopped) * (m[:, None] - sub)).sum())
    print("%s signature grouping %s: used %.3f of the lane-columns; %s; longest lane per step %.1f columns; %.3f of the walked pairs stop at a gap" % (
        name, "on" if by_sig else "off", used / (32 * longest), {k: round(v / (32 * longest), 4) for k, v in loss.items()}, longest / steps, gaps / walked_n))
