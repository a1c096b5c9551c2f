#!/usr/bin/env python
"""Kernel experiments on one GPU (env knobs are read once per process, so run one configuration per invocation):
    python tools/kexp.py wgs      # configs[1] batch, one batch at a time: pre-filter / survivors kernel durations
    python tools/kexp.py panel    # 16384 columns at 1500x/1500x: product path and every-column path
    python tools/kexp.py full     # configs[1] batch, every column through the full model"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from somatic_sniper_b200 import api, synth


def timed(sc, b, flags, reps):
    for _ in range(3):
        sc.score_resident(b, flags)
    sc.sync()
    sites = sc.fetch(b)
    sc.kernel_time(reset=True)
    sc.timer_start()
    for _ in range(reps):
        sc.score_resident(b, flags)
    ms = sc.timer_stop() / reps
    pf, score, n = sc.kernel_time_split(reset=True)
    return dict(ms_per_step=ms, prefilter_ms=pf / max(n, 1), score_ms=score / max(n, 1), sites=len(sites), survivors=b.survivors()), sites


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "wgs"
    knobs = {k: v for k, v in os.environ.items() if k.startswith("SSN_")}
    if mode == "panel":
        import numpy as np
        unit = synth.make_columns(16384, depth_t=1500, depth_n=1500, seed=1734)
        reps, ue = int(os.environ.get("KEXP_REPS", 4)), int(unit["n_entries"])
        cols = dict(off=np.concatenate([unit["off"][:-1].astype(np.int64) + r * ue for r in range(reps)] + [np.array([reps * ue], np.int64)]).astype(np.uint32),
                    bs=np.tile(unit["bs"], reps), bq=np.tile(unit["bq"], reps), mq=np.tile(unit["mq"], reps),
                    ref_char=np.tile(unit["ref_char"], reps), ncol=reps * unit["ncol"], n_entries=reps * ue)
    else:
        cols = synth.make_columns(1_000_000, seed=1234)
    sc = api.Scorer(api.default_params(), device=0, n_slots=1, max_columns=cols["ncol"], max_entries=int(cols["n_entries"]) + 64)
    b = sc.acquire(); b.fill(cols); sc.upload(b)
    out = dict(mode=mode, knobs=knobs, columns=cols["ncol"], read_bases=int(cols["n_entries"]), algorithmic_bytes=synth.algorithmic_bytes(cols))
    if mode in ("wgs", "panel"):
        out["product"], s1 = timed(sc, b, 0, 20 if mode == "wgs" else 5)
    if mode in ("full", "panel"):
        out["every_column"], s2 = timed(sc, b, api.SUBMIT_NO_PREFILTER, 5)
    if mode == "panel":
        assert s1.tobytes() == s2.tobytes()
    for k in ("product", "every_column"):
        if k in out:
            out[k]["hbm_GBs"] = out["algorithmic_bytes"] / (out[k]["ms_per_step"] * 1e-3) / 1e9
    sc.close()
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
