import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from haploclique_b200 import edge
from oracle import pyoracle
from tests import common
b = common.load_batch("hiv50_records.npz")
kw = common.quasi_kwargs({}, b.max_position())
pi, pj = common.quasi_pair_sample(b.n)
calc = edge.QuasispeciesEdgeCalculator(**kw); calc.upload(b)
dev = calc.score_pairs(pi, pj); orc = pyoracle.QuasiOracle(b, **kw).pairs(pi, pj)
bad = np.flatnonzero((dev["q"] == 0) != (orc["q"] == 0))
print("mismatching q==0 pattern:", bad.size, "of", pi.size)
for k in bad[:6]:
    i, j = int(pi[k]), int(pj[k])
    print(i, j, "dev", dev[k], "orc", orc[k])
    for r in (i, j):
        kk = 2 * r
        cig = bytes(b.cig[int(b.cig_off[kk]):int(b.cig_off[kk]) + int(b.cig_len[kk])]).decode()
        import itertools
        rle = "".join("%d%s" % (len(list(g)), c) for c, g in itertools.groupby(cig))
        print("   read", r, "start", int(b.seg_start[kk]), "end", int(b.seg_end[kk]), "seq_len", int(b.seq_len[kk]), rle)
