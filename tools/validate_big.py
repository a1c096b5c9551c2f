#!/usr/bin/env python
"""Size-independent check at BASELINE configs[2] scale (50 Mb, -J joint genotyping): every 1 Mb shard is scored twice through
the C ABI - product path (pre-filter + survivors) and every-column path - and the site records must agree byte for byte.
    python tools/validate_big.py [--mb 50] [--out gpurun_out/validate_big.json]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from somatic_sniper_b200 import api, synth

ap = argparse.ArgumentParser()
ap.add_argument("--mb", type=int, default=50)
ap.add_argument("--out", default=None)
a = ap.parse_args()
ncol = 1_000_000
prm = api.default_params(use_joint_priors=1, somatic_mutation_rate=0.01)
sc = None
tot = dict(columns=0, read_bases=0, sites=0, survivors=0, mismatching_shards=0)
t0 = time.time()
for shard in range(a.mb):
    cols = synth.make_columns(ncol, seed=5000 + shard, binned=(shard % 2 == 0))
    if sc is None:
        sc = api.Scorer(prm, device=0, n_slots=1, max_columns=ncol, max_entries=int(1.02 * cols["n_entries"]) + 64)
    fast, _ = sc.score(cols)
    surv = sc.last_survivors
    full, _ = sc.score(cols, flags=api.SUBMIT_NO_PREFILTER)
    same = fast.tobytes() == full.tobytes()
    tot["columns"] += ncol; tot["read_bases"] += int(cols["n_entries"]); tot["sites"] += len(full); tot["survivors"] += surv
    tot["mismatching_shards"] += 0 if same else 1
    print("shard %d: %d sites, %d survivors, identical=%s" % (shard, len(full), surv, same), flush=True)
sc.close()
tot["identical"] = tot["mismatching_shards"] == 0
tot["seconds"] = time.time() - t0
tot["what"] = "%d shards of 1e6 synthetic columns (60x/30x, binned and unbinned qualities alternating), -J joint priors; product path vs SSN_SUBMIT_NO_PREFILTER" % a.mb
print(json.dumps(tot))
if a.out:
    json.dump(tot, open(a.out, "w"), indent=1)
sys.exit(0 if tot["identical"] else 1)
