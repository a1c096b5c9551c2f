#!/usr/bin/env python
"""Small workloads for compute-sanitizer (memcheck / racecheck / synccheck) covering every kernel: the smoke test (fast
tiles, pre-filter, survivors, device pileup), a panel-depth batch (bin-counting deep path and, with SSN_NO_HIST=1, the
sorting path with its global key scratch), a ragged batch whose tiles need several staging passes in the pre-filter, and
base qualities outside [1,63] (generic tiles).  Sizes are kept tiny: the tools slow kernels down 10-100x.
    compute-sanitizer --tool memcheck python tools/sanitize_gpu.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as g
from somatic_sniper_b200 import api, synth


def run(cols, flags_list=(0, api.SUBMIT_NO_PREFILTER, api.SUBMIT_COLUMN_DUMP), slots=2):
    sc = api.Scorer(api.default_params(min_somatic_qual=0), device=0, n_slots=slots, max_columns=cols["ncol"], max_entries=int(cols["n_entries"]) + 64)
    out = []
    for f in flags_list:
        b = sc.acquire(); b.fill(cols); sc.submit(b, f)
        s, _ = sc.results_next()
        out.append(s)
    sc.close()
    assert out[0].tobytes() == out[1].tobytes()
    return len(out[0])


if __name__ == "__main__":
    g.smoke()
    print("deep", run(synth.make_columns(96, depth_t=1400, depth_n=900, seed=3, binned=False, p_som=5e-2)))
    print("ragged", run(synth.make_columns(2048, depth_t=150, depth_n=20, seed=4, min_depth=0)))
    print("edge", run(synth.make_columns(2048, seed=5, edge=0.5, p_som=1e-2)))
    print("mid", run(synth.make_columns(512, depth_t=200, depth_n=180, seed=6)))
    print("sanitize workloads ok")
