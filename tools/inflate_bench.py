#!/usr/bin/env python
"""Raw throughput of the GPU inflater on a synthetic BAM: the whole file's blocks resident in one lane, kernel + copies
timed together and the kernel's share estimated from a second run with the copies already done (events are inside the
library, so this measures what a reader sees: H2D + kernel + D2H per batch).
    python tools/inflate_bench.py [--len 16000000] [--level 1] [--batch 256] [--threads 1,8,32]"""
import argparse, json, os, struct, sys, tempfile, threading, time, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from somatic_sniper_b200 import api, cli_tools as ct


def blocks_of(path, limit):
    raw = open(path, "rb").read(limit)
    at, pl, sz = 0, [], []
    while at + 18 < len(raw):
        xlen = struct.unpack_from("<H", raw, at + 10)[0]
        bsize = struct.unpack_from("<H", raw, at + 16)[0] + 1
        if at + bsize > len(raw):
            break
        n = struct.unpack_from("<I", raw, at + bsize - 4)[0]
        if n:
            pl.append(raw[at + 12 + xlen:at + bsize - 8]); sz.append(n)
        at += bsize
    return pl, sz


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--len", type=int, default=8_000_000)
    ap.add_argument("--level", type=int, default=1)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--threads", default="1,4,16,48")
    ap.add_argument("--seconds", type=float, default=2.0)
    a = ap.parse_args()
    d = tempfile.mkdtemp(prefix="ssn_infl_")
    pair = ct.make_pair_fast(d, [("chr1", a.len)], 60, 30, seed=1, level=a.level)
    pl, sz = blocks_of(pair["tumor"], 600 << 20)
    res = dict(blocks=len(pl), compressed=sum(map(len, pl)), inflated=sum(sz), level=a.level, batch=a.batch, runs=[])
    t0 = time.perf_counter()
    for p in pl[:2000]:
        zlib.decompress(p, -15)
    res["zlib_python_MBs_1core"] = sum(sz[:2000]) / (time.perf_counter() - t0) / 1e6
    for nthreads in [int(x) for x in a.threads.split(",")]:
        inf = api.Inflater(n_lanes=nthreads, in_cap=a.batch * 66000, out_cap=a.batch * 65536, max_blocks=a.batch)
        done = [0] * nthreads
        stop = time.perf_counter() + a.seconds + 0.5
        def work(t):
            lane, hin, hout = inf.acquire()
            k = (t * 997) % max(1, len(pl) - a.batch)
            blk, x, y = [], 0, 0
            for j in range(k, k + a.batch):
                hin[x:x + len(pl[j])] = np.frombuffer(pl[j], np.uint8)
                blk.append((x, len(pl[j]), y, sz[j])); x += len(pl[j]); y += sz[j]
            blk = np.array(blk, api.INFLATE_BLOCK_DTYPE)
            r = inf.run(lane, blk, x, y)          # warm-up (lane buffers, module load)
            assert not r.any()
            assert hout[:sz[k]].tobytes() == zlib.decompress(pl[k], -15)
            t_start = time.perf_counter()
            while time.perf_counter() < stop:
                inf.run(lane, blk, x, y)
                done[t] += y
            done[t] = (done[t], time.perf_counter() - t_start)
            inf.release(lane)
        th = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
        [t.start() for t in th]; [t.join() for t in th]
        total = sum(b / s for b, s in done)
        res["runs"].append(dict(threads=nthreads, GBs=total / 1e9, per_thread_MBs=total / nthreads / 1e6))
        inf.close()
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
