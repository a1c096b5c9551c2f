import sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from somatic_sniper_b200 import api, synth
import prefilter_np
n = 300000
cols = synth.make_columns(n, seed=41)
sc = api.Scorer(api.default_params(), device=0, n_slots=1, max_columns=n, max_entries=int(cols["n_entries"]) + 64)
_, dump = sc.score(cols, dump=True)
sc.close()
dropped = dump[:, 29] == 1
want = prefilter_np.decide(cols)
x = np.nonzero(dropped ^ want)[0]
print("kernel-only drops", (dropped & ~want).sum(), "numpy-only drops", (~dropped & want).sum())
off = cols["off"].astype(np.int64)
ref16 = api.nt16_lut()[cols["ref_char"]]
for c in x[:12]:
    out = [int(c), int(c) & 31, bool(dropped[c])]
    for s in (0, 1):
        a, b = off[2 * c + s], off[2 * c + s + 1]
        bs, bq, mq = cols["bs"][a:b], cols["bq"][a:b], cols["mq"][a:b]
        nr = (bs & 15) != ref16[c]
        out.append(dict(n=int(b - a), m=int(nr.sum()), S=int(bq[nr].sum()), n1=int((~nr & ((bs >> 4) == 1)).sum()), start=int(a) & 3))
    print(out)
