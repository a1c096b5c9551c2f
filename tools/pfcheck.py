import sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from somatic_sniper_b200 import api, synth
import prefilter_np
for name, kw in (("wgs", dict(ncol=300000, seed=41)), ("edge", dict(ncol=100000, seed=43, edge=0.5, p_som=5e-3)), ("shallow", dict(ncol=100000, depth_t=12, depth_n=6, seed=44, edge=0.1)), ("wgs100x", dict(ncol=100000, depth_t=110, depth_n=100, seed=46, binned=False)),
                 ("q0", dict(ncol=100000, seed=50, q0=True)),
                 ("ragged", dict(ncol=100000, depth_t=150, depth_n=20, seed=49, edge=0.0, min_depth=0))):
    n = kw.pop("ncol"); q0 = kw.pop("q0", False); cols = synth.make_columns(n, **kw)
    if q0:      # N read-bases and a sprinkle of others carry quality 0, as in many real BAMs
        rng = np.random.default_rng(3); z = ((cols["bs"] & 15) == 15) | (rng.random(len(cols["bq"])) < 2e-3); cols["bq"] = np.where(z, 0, cols["bq"]).astype(np.uint8)
    sc = api.Scorer(api.default_params(), device=0, n_slots=1, max_columns=n, max_entries=int(cols["n_entries"]) + 64)
    _, dump = sc.score(cols, dump=True)
    dropped = dump[:, 29] == 1
    tg, ng = prefilter_np.consensus_genotypes(dump)
    ref16 = api.nt16_lut()[cols["ref_char"]]
    bad = dropped & ((tg != ref16) | (ng != ref16))
    full, _ = sc.score(cols, flags=api.SUBMIT_NO_PREFILTER)
    fast, _ = sc.score(cols)
    want = prefilter_np.decide(cols)
    print(name, "dropped %.4f numpy %.4f bad %d identical %s sites %d survivors %d xor %d" % (dropped.mean(), want.mean(), bad.sum(), fast.tobytes() == full.tobytes(), len(full), sc.last_survivors, (dropped ^ want).sum()), flush=True)
    sc.close()
