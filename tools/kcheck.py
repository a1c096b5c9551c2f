#!/usr/bin/env python
"""Quick kernel iteration loop on the GPU box: parity of a few batches against the oracle (test infrastructure), then the
resident-batch rate of the bench workload.  python tools/kcheck.py [--no-parity]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from somatic_sniper_b200 import api, synth

def parity():
    from oracle.loader import Oracle
    ok = True
    cases = [("wgs", dict(ncol=60000, seed=3)), ("edge", dict(ncol=40000, seed=4, edge=0.5, p_som=5e-3)),
             ("shallow", dict(ncol=30000, depth_t=12, depth_n=6, seed=5, edge=0.2)), ("mid", dict(ncol=8000, depth_t=100, depth_n=70, seed=6, binned=False)),
             ("deep", dict(ncol=600, depth_t=1500, depth_n=900, seed=7))]
    for joint in (0, 1):
        orc = Oracle(use_joint=joint, min_somatic_qual=0 if joint else 15)
        prm = api.default_params(use_joint_priors=joint, min_somatic_qual=0 if joint else 15)
        for name, kw in cases:
            kw = dict(kw); n = kw.pop("ncol")
            cols = synth.make_columns(n, **kw)
            sc = api.Scorer(prm, device=0, n_slots=1, max_columns=n, max_entries=int(cols["n_entries"]) + 64)
            sites, dump = sc.score(cols, dump=True)
            sc.close()
            want = orc.score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"], want_glf=True)
            g = np.array_equal(dump[:, :28], want["glf"])
            s = np.array_equal(sites["column"], want["emit_col"])
            idx = [i for i in range(53) if i != 27]
            r = s and np.array_equal(api.sites_to_records(sites, cols["ref_char"])[:, idx], want["recs"][:len(sites)][:, idx])
            print("parity %-8s joint=%d glf=%s sites=%s recs=%s (%d sites)" % (name, joint, g, s, r, len(sites)), flush=True)
            if not g:
                bad = np.nonzero((dump[:, :28] != want["glf"]).any(1))[0]
                print("   first bad columns", bad[:5], "of", len(bad))
            ok &= g and s and r
    return ok

def rate(ncol=1_000_000, steps=10, nslots=2):
    cols = synth.make_columns(ncol, seed=1234)
    sc = api.Scorer(api.default_params(), device=0, n_slots=nslots, max_columns=ncol, max_entries=int(cols["n_entries"]) + 64)
    bs = []
    for _ in range(nslots):
        b = sc.acquire(); b.fill(cols); sc.upload(b); bs.append(b)
    ref = None
    for name, flags, use in (("full model, every column", api.SUBMIT_NO_PREFILTER, bs[:1]), ("pre-filter + survivors, 1 stream", 0, bs[:1]),
                             ("pre-filter + survivors, %d streams" % nslots, 0, bs)):
        for i in range(4): sc.score_resident(use[i % len(use)], flags)
        sc.sync(); [sc.fetch(b) for b in use]; sc.kernel_time(reset=True)
        sc.timer_start()
        for i in range(steps): sc.score_resident(use[i % len(use)], flags)
        ms = sc.timer_stop()
        sites = sc.fetch(use[0])
        pf, scms, nb = sc.kernel_time_split(reset=True)
        if ref is None: ref = sites
        same = sites.tobytes() == ref.tobytes()
        print("rate %-38s %.4g columns/s  (%.3f ms per 1e6-column step; pre-filter %.3f ms, scoring %.3f ms; %d sites, %d survivors, identical=%s)"
              % (name, ncol * steps / (ms * 1e-3), ms / steps, pf / max(nb, 1), scms / max(nb, 1), len(sites), use[0].survivors(), same), flush=True)
    sc.close()

if __name__ == "__main__":
    ok = True
    if "--no-parity" not in sys.argv:
        ok = parity()
    rate(nslots=int(os.environ.get("KCHECK_SLOTS", "2")))
    sys.exit(0 if ok else 1)
