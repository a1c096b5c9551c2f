"""Device pileup at bench scale: one 16-window read batch (60x/30x, 150 bp reads) checked against the host packer,
then timed: the pileup kernels alone and the pipelined end-to-end rate (pinned reads -> sites) over 8 slots.
Usage (GPU box): python tools/pileup_check.py [--windows 16] [--steps 20] [--no-parity]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from somatic_sniper_b200 import api, hostapi, reads as R  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--windows", type=int, default=16)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--slots", type=int, default=8)
    ap.add_argument("--no-parity", action="store_true")
    a = ap.parse_args()
    length = a.windows * 8192
    ref = R.make_reference(1, length + 1024)
    ref_t = ref.copy()
    ref_t[::5003] = np.where(ref_t[::5003] == 1, 2, 1)
    t0 = time.time()
    t = R.make_reads(2, ref_t, 60, 150, 0, length, low_mapq_frac=0.02)
    n = R.make_reads(3, ref, 30, 150, 0, length, low_mapq_frac=0.02)
    rw = R.build_read_windows(t, n, ref, first_window=0, n_windows=a.windows)
    print("built %d windows: %d records, %.1f MB of records in %.1f s" % (a.windows, rw["rec_off"].size, rw["recs"].size / 1e6, time.time() - t0), file=sys.stderr)
    sc = api.Scorer(n_slots=a.slots, max_columns=length, max_entries=rw["entries_bound"])
    sc.reads_reserve(rw["recs"].size + 64, rw["rec_off"].size + 1, a.windows)
    # every slot's pinned read arrays get the same batch; the data stays there across acquire/release
    held = [sc.acquire() for _ in range(a.slots)]
    for b in held:
        args = sc.fill_reads(b, rw)
    for b in held:
        sc.release(b)
    out = {}
    b = sc.acquire()
    sc.submit_reads(b, *args)
    sites, _ = sc.results_next()
    pos, nent = sc.column_positions(b)
    out.update(columns=int(pos.size), entries=nent, sites=int(sites.size), rec_bytes=int(rw["recs"].size),
               h2d_bytes=int(rw["recs"].size + 4 * rw["rec_off"].size + 24 * a.windows + 8192 * a.windows),
               packed_bytes=int(3 * nent + 9 * pos.size))
    if not a.no_parity:
        want = hostapi.pack_read_windows(rw)
        got = sc.download_columns(b)
        ok = all(np.array_equal(got[k], want[k]) for k in ("off", "bs", "bq", "mq", "ref_nt16", "pos"))
        out["identical_to_host_packer"] = bool(ok)
        t1 = time.time()
        hostapi.pack_read_windows(rw)
        out["host_pack_1thread_s"] = time.time() - t1
    sc.release(b)
    # one batch at a time: the pileup kernels' own duration comes from the events around them
    sc.pileup_time(reset=True)
    sc.kernel_time_split(reset=True)
    for _ in range(4):
        b = sc.acquire()
        sc.submit_reads(b, *args)
        sc.results_next()
        sc.release(b)
    ms, nb = sc.pileup_time(reset=True)
    pf, score, kb = sc.kernel_time_split(reset=True)
    out.update(pileup_kernels_ms=ms / nb, prefilter_ms=pf / max(kb, 1), score_ms=score / max(kb, 1))
    # pipeline: prime every slot, then per step drain all / submit all
    live = []
    for _ in range(a.slots):
        b = sc.acquire()
        sc.submit_reads(b, *args)
        live.append(b)
    for phase in ("warm", "timed"):
        steps = 3 if phase == "warm" else a.steps
        sc.timer_start()
        w0 = time.time()
        for _ in range(steps):
            for k in range(a.slots):
                sc.results_next()
                sc.release(live[k])
                live[k] = sc.acquire()
                sc.submit_reads(live[k], *args)
        dev_ms = sc.timer_stop()
        wall = time.time() - w0
    for _ in range(a.slots):
        sc.results_next()
    cols = out["columns"] * a.slots * a.steps
    out.update(e2e_columns_per_s=cols / (dev_ms / 1e3), e2e_device_ms_per_step=dev_ms / a.steps, e2e_wall_ms_per_step=1e3 * wall / a.steps,
               h2d_gb_per_s=out["h2d_bytes"] * a.slots * a.steps / (dev_ms / 1e3) / 1e9)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
