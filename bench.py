#!/usr/bin/env python
"""bench.py — pileup columns/s scored (BASELINE.json metric) on N B200s of one node.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (libssn.so, C ABI)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU code on the host cores

A step = one pass of the hot path over one batch of synthetic columns.  Workload at every N:
BASELINE.json configs[1] — a synthetic 1 Mb region, normal 30x / tumor 60x, default options
(-Q 15, independent priors): 1,000,000 columns per GPU (weak scaling: region shards are independent,
no collective on the data path; torch.distributed only carries the barrier and the max-over-ranks).

value      columns/s, all ranks, batches already resident in HBM, timed with CUDA events; the product path:
           the exact hom-reference pre-filter kernel over every column, then the full model on the survivors;
           8 resident batches in flight on their slot streams (the survivors' kernel of one batch overlaps the
           next batch's pre-filter); single_stream = the same with one batch at a time
full_model the same with SSN_SUBMIT_NO_PREFILTER: every column through the full model (identical sites)
e2e        the same through ssn_batch_submit/ssn_results_next with pinned HOST batches, 8 batches in flight
           (steady-state pipeline): every timed step drains 8 results and submits 8 batches, i.e. H2D of every
           column + kernels + D2H of the site ring inside the timed region
roofline   HBM: algorithmic bytes (3 B per read-base + 9 B per column) per launch /
           the dominant kernel's mean launch duration (events around each launch on its own stream); the dominant
           kernel is the pre-filter (it reads every byte of the batch); step_frac = the same over both kernels
cpu_baseline  the unmodified reference (oracle/_ref, 1 thread) on a bounded sample, rank 0, N=1 only
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pileup_columns_per_sec_scored"
UNIT = "columns/s"
COLUMNS = 1_000_000
CHUNKS = 8                      # e2e: the step's columns travel as 8 pinned batches on 8 streams
STREAMS = 8                     # value: resident batches in flight (slot streams), the way the packer pipeline runs (same as CHUNKS)
WORKLOAD = "synthetic 1 Mb region, normal 30x / tumor 60x, -Q 15 independent priors (BASELINE configs[1])"


REPEATS = 5                     # the K-step timed region of `value` is repeated; the median is reported, the spread beside it


def workload_config(ncol, depth_t, depth_n, alg_bytes):
    """`config` is the workload and nothing else, identical in both arms (the driver compares them)."""
    return {"workload": WORKLOAD, "columns_per_gpu_per_step": int(ncol), "depth_tumor": depth_t, "depth_normal": depth_n,
            "l2": "inputs (%.0f MB per step) exceed the 126 MB L2; no flush" % (alg_bytes / 1e6)}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--columns", type=int, default=COLUMNS)
    ap.add_argument("--depth-t", type=float, default=60.0)
    ap.add_argument("--depth-n", type=float, default=30.0)
    ap.add_argument("--cpu-sample", type=int, default=1_000_000, help="columns of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-e2e-reads", action="store_true")
    ap.add_argument("--no-e2e-bam", action="store_true")
    ap.add_argument("--no-panel", action="store_true")
    ap.add_argument("--bam-contigs", type=int, default=4)
    ap.add_argument("--bam-len", type=int, default=8_000_000)
    ap.add_argument("--backend", default="nccl", help=argparse.SUPPRESS)
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured)"
    except Exception:
        return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


def ncu_traffic(key="dram_bytes_per_launch"):
    """dram bytes per launch of a kernel from the committed ncu capture (profiles/traffic.json), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(key)
    except Exception:
        return None


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.rows = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) >= 7:
                self.rows.append(f)
        self.proc = None

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for f in self.rows:
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def slice_columns(cols, c0, c1):
    off = cols["off"].astype(np.int64)
    e0, e1 = int(off[2 * c0]), int(off[2 * c1])
    return dict(off=(off[2 * c0:2 * c1 + 1] - e0).astype(np.uint32), bs=cols["bs"][e0:e1], bq=cols["bq"][e0:e1],
                mq=cols["mq"][e0:e1], ref_char=cols["ref_char"][c0:c1], ncol=c1 - c0, n_entries=e1 - e0)


def dist_setup(n, backend="nccl"):
    """torch.distributed is plumbing only: barrier + max/sum over ranks.  Region shards need no data-path collective."""
    if n <= 1:
        return None, 0, 0
    import torch
    import torch.distributed as dist
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    if backend == "nccl":
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:                                   # gloo: the CPU tests of the N>1 bookkeeping
        dist.init_process_group(backend)
    return dist, rank, local


def bind_near_gpu(local):
    """N > 1: keep this rank (and so the first touch of its pinned batches) on the CPUs of the NUMA node its GPU hangs off.
    Round 1's scaling run lost half of e2e at N = 8 to the host side of eight concurrent pinned H2D streams; where the box
    has several nodes this keeps each stream's source memory local.  A no-op on a one-node box or when sysfs says nothing."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open("/sys/bus/pci/devices/%s/local_cpulist" % bdf) as f:
            txt = f.read().strip()
        cpus = set()
        for part in txt.split(","):
            if part:
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        have = os.sched_getaffinity(0)
        want = cpus & have
        if want and want != have:
            os.sched_setaffinity(0, want)
            return "GPU %s: bound to its node's CPUs %s (%d of %d)" % (bdf, txt, len(want), len(have))
        return "GPU %s: local CPUs %s = every CPU of this process, nothing to bind" % (bdf, txt or "?")
    except Exception as e:              # no sysfs entry, no such property: run unbound
        return "unbound (%s)" % type(e).__name__


def _dist_device(dist):
    return "cuda" if dist.get_backend() == "nccl" else "cpu"


def dist_barrier(dist):
    if dist is not None:
        import torch
        dist.barrier()
        if _dist_device(dist) == "cuda":
            torch.cuda.synchronize()


def dist_max(dist, x):
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=_dist_device(dist))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def dist_sum(dist, x):
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=_dist_device(dist))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def shard_seed(rank):
    """Rank r scores its own region shard: an independent 1 Mb region drawn with seed 1234 + r (weak scaling)."""
    return 1234 + rank


def whole_job_rate(dist, columns_this_rank, steps, ms_this_rank):
    """columns/s of the whole job: units of all ranks / the slowest rank's device time."""
    ms_max = dist_max(dist, ms_this_rank)
    total = dist_sum(dist, float(columns_this_rank))
    return total * steps / (ms_max * 1e-3), ms_max, total


# ---------------------------------------------------------------------------------------------------
# CPU arms (the only places bench.py executes anything under oracle/)

_W = {}


def _cpu_worker_init(kind):
    from oracle.loader import Oracle, Reference
    _W["impl"] = Reference() if kind == "reference" else Oracle()


def _cpu_worker_run(shard):
    if isinstance(shard, int):          # forked workers inherit the shards; only the index travels
        shard = _W["shards"][shard]
    r = _W["impl"].score_batch(shard["off"], shard["bs"], shard["bq"], shard["mq"], shard["ref_char"], max_recs=4096)
    return int(r["n_emit"])


def cpu_kind():
    from oracle.loader import Reference
    return "reference" if Reference.available() else "port"


def cpu_baseline_single(cols, sample_cols):
    """The reference's per-column path (glf_somatic over fabricated pileups), 1 thread, bounded sample."""
    kind = cpu_kind()
    _cpu_worker_init(kind)
    shard = slice_columns(cols, 0, min(sample_cols, cols["ncol"]))
    t = time.perf_counter()
    _cpu_worker_run(shard)
    dt = time.perf_counter() - t
    return {"value": shard["ncol"] / dt, "unit": UNIT, "cores": 1, "kind": kind,
            "sample": "first %d columns of the step's batch, %.1f s, oracle/%s" % (
                shard["ncol"], dt, "_ref/libref_harness.so (unmodified reference)" if kind == "reference" else "ssn_oracle.c")}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    from somatic_sniper_b200 import synth
    kind = cpu_kind()
    cores = len(os.sched_getaffinity(0))
    sample_cols = min(args.columns, args.cpu_sample)
    cols = synth.make_columns(sample_cols, depth_t=args.depth_t, depth_n=args.depth_n, seed=1234)
    bounds = np.linspace(0, sample_cols, cores + 1).astype(int)
    shards = [slice_columns(cols, int(bounds[i]), int(bounds[i + 1])) for i in range(cores)]
    ctx = mp.get_context("fork")
    _W["shards"] = shards
    idx = list(range(cores))
    with ctx.Pool(cores, initializer=_cpu_worker_init, initargs=(kind,)) as pool:
        for _ in range(args.warmup):
            pool.map(_cpu_worker_run, idx, chunksize=1)
        t = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_cpu_worker_run, idx, chunksize=1)
        dt = time.perf_counter() - t
    value = sample_cols * args.steps / dt
    sample = "%d columns per step region-split over %d processes, oracle/%s" % (
        sample_cols, cores, "_ref (unmodified reference glf_somatic)" if kind == "reference" else "ssn_oracle.c port")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64/f32+int", "data": "synthetic",
            "config": workload_config(sample_cols, args.depth_t, args.depth_n, synth.algorithmic_bytes(cols)),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=_JSON_OUT, flush=True)


# ---------------------------------------------------------------------------------------------------

def bam_level(args):
    """bam-somaticsniper (this repo's CLI, SSN_GPUS = N) on a synthetic 30x/60x BAM pair: columns/s from the program's own
    statistics (streaming = wall minus set-up).  The pair is generated once per box under /tmp and reused by later runs."""
    import tempfile
    from somatic_sniper_b200 import cli_tools as ct
    if not (os.path.exists(ct.CLI) and os.path.exists(ct.SIMBAM)):
        return {"unavailable": "bam-somaticsniper / ssn-simbam are not built"}
    contigs = [("chr%d" % (i + 1), args.bam_len) for i in range(args.bam_contigs)]
    d = os.path.join(tempfile.gettempdir(), "ssn_bench_bam_%dx%d_%g_%g" % (args.bam_contigs, args.bam_len, args.depth_t, args.depth_n))
    t0 = time.perf_counter()
    marker = os.path.join(d, "pair.json")
    if os.path.exists(marker):
        with open(marker) as f:
            pair = json.load(f)
    else:
        pair = ct.make_pair_fast(d, contigs, args.depth_t, args.depth_n, seed=1234, level=1)
        with open(marker, "w") as f:
            json.dump(pair, f)
    t_gen = time.perf_counter() - t0
    out = os.path.join(d, "out.%d" % args.gpus)
    runs = []
    for _ in range(3):
        r = ct.run_cli(ct.CLI, pair["fasta"], pair["tumor"], pair["normal"], out, [], env={"SSN_GPUS": args.gpus}, stats=os.path.join(d, "stats.json"))
        if r["rc"] != 0:
            return {"unavailable": "bam-somaticsniper failed: %s" % r["stderr"][-300:]}
        runs.append((r["wall_s"], r["stats"]))
    wall, st = min(runs, key=lambda x: x[0])
    bam_bytes = os.path.getsize(pair["tumor"]) + os.path.getsize(pair["normal"])
    return {"value": st["columns_per_s"], "unit": UNIT, "columns": st["columns"], "read_bases": st["read_bases"], "sites": st["sites"],
            "wall_s": wall, "setup_s": st["setup_s"], "stream_s": st["wall_s"] - st["setup_s"], "columns_per_s_wall": st["columns"] / wall,
            "all_walls_s": [w for w, _ in runs], "bam_bytes": bam_bytes, "shards": st["shards"], "threads": st["threads"],
            "host_cores": os.cpu_count(), "gpus": args.gpus, "pack_wait_s": st["pack_wait_s"], "gpu_wait_s": st["gpu_wait_s"],
            "gpu_inflate": st.get("gpu_inflate"), "gpu_inflate_blocks": st.get("gpu_inflate_blocks"), "gpu_inflate_refused": st.get("gpu_inflate_refused"),
            "generate_s": t_gen,
            "workload": "%d contigs x %d bp, tumor %gx / normal %gx, 150 bp reads, BGZF level 1 (tools/simbam_fast.c); classic output, "
                        "default options; file cache warm" % (args.bam_contigs, args.bam_len, args.depth_t, args.depth_n)}


def run_b200(args):
    from somatic_sniper_b200 import api, synth
    dist, rank, local = dist_setup(args.gpus, args.backend)
    binding = bind_near_gpu(local) if args.gpus > 1 and args.backend == "nccl" else "single rank, unbound"
    ncol = args.columns
    cols = synth.make_columns(ncol, depth_t=args.depth_t, depth_n=args.depth_n, seed=shard_seed(rank))
    nent = int(cols["n_entries"])
    alg_bytes = synth.algorithmic_bytes(cols)
    bounds = np.linspace(0, ncol, CHUNKS + 1).astype(int)
    chunks = [slice_columns(cols, int(bounds[i]), int(bounds[i + 1])) for i in range(CHUNKS)]

    # ---- value: resident batch, kernel only --------------------------------------------------------
    sc = api.Scorer(api.default_params(), device=local, n_slots=STREAMS, max_columns=ncol, max_entries=nent + 64)
    resident = []
    for _ in range(STREAMS):
        b = sc.acquire()
        b.fill(cols)
        sc.upload(b)
        resident.append(b)
    b = resident[0]
    filtered = sc.prefilter_enabled()
    # every column through the full model (the pre-filter switched off): same sites, reported beside the product path
    for _ in range(max(args.warmup, 3)):
        sc.score_resident(b, api.SUBMIT_NO_PREFILTER)
    sc.sync()
    sites_full = sc.fetch(b)
    sc.kernel_time(reset=True)
    dist_barrier(dist)
    sc.timer_start()
    for _ in range(args.steps):
        sc.score_resident(b, api.SUBMIT_NO_PREFILTER)
    ms_full = sc.timer_stop()
    dist_barrier(dist)
    _, full_kms, full_kn = sc.kernel_time_split(reset=True)
    for _ in range(max(args.warmup, 3)):
        sc.score_resident(b)
    sc.sync()
    sites_ref = sc.fetch(b)
    survivors = b.survivors()
    assert sites_ref.tobytes() == sites_full.tobytes(), "pre-filtered path and full-model path disagree"
    sc.kernel_time(reset=True)
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)                    # nvidia-smi needs a moment before its first sample
    # one batch at a time: the kernels run back to back, so their durations are their own (roofline)
    dist_barrier(dist)
    sc.sync()
    sc.timer_start()
    for _ in range(args.steps):
        sc.score_resident(b)
    ms_single = sc.timer_stop()
    dist_barrier(dist)
    pf_ms, score_ms, kn = sc.kernel_time_split(reset=True)
    kms = pf_ms + score_ms
    n_sites = len(sites_ref)
    # STREAMS batches in flight
    for i in range(2 * STREAMS):
        sc.score_resident(resident[i % STREAMS])
    sc.sync()
    for r in resident:
        assert sc.fetch(r).tobytes() == sites_ref.tobytes()
    sc.kernel_time(reset=True)
    rep_ms = []
    for _ in range(REPEATS):            # each repeat: exactly K steps between barrier + sync on both sides
        l0 = sc.kernel_launches()
        dist_barrier(dist)
        sc.sync()
        sc.timer_start()
        for i in range(args.steps):
            sc.score_resident(resident[i % STREAMS])
        rep_ms.append(dist_max(dist, sc.timer_stop()))
        dist_barrier(dist)
        launches = sc.kernel_launches() - l0
    ms = float(np.median(rep_ms))
    ov_pf_ms, ov_score_ms, ov_kn = sc.kernel_time_split(reset=True)
    for r in resident:
        sc.release(r)
    sc.close()
    value, ms_max, total_cols = whole_job_rate(dist, ncol, args.steps, ms)
    single_value, single_ms_max, _ = whole_job_rate(dist, ncol, args.steps, ms_single)
    peak, peak_src = measured_peak()
    k_avg_ms = kms / max(kn, 1)
    full_value, full_ms_max, _ = whole_job_rate(dist, ncol, args.steps, ms_full)
    full_k_ms = full_kms / max(full_kn, 1)
    if filtered:
        dom_name, dom_ms, dom_traffic = "ssn_prefilter_kernel", pf_ms / max(kn, 1), ncu_traffic("prefilter_dram_bytes_per_launch")
    else:
        dom_name, dom_ms, dom_traffic = "ssn_score_kernel<false>", k_avg_ms, ncu_traffic()
    achieved = alg_bytes / (dom_ms * 1e-3) / 1e9

    # ---- e2e: pinned host batches through submit/results_next ----------------------------------------
    e2e = None
    if not args.no_e2e:
        cmax = max(c["ncol"] for c in chunks)
        emax = max(int(c["n_entries"]) for c in chunks)
        sc = api.Scorer(api.default_params(), device=local, n_slots=CHUNKS, max_columns=cmax, max_entries=emax + 64)
        slots = []
        for c in chunks:                       # the packer's job: columns land in pinned memory once
            bb = sc.acquire()
            bb.fill(c)
            slots.append((bb, c["ncol"], int(c["n_entries"])))

        # steady-state pipeline, the way the packer drives it: CHUNKS batches stay in flight; a step drains the CHUNKS
        # oldest results and submits the same CHUNKS pinned batches again (H2D + kernels + D2H of every one of them)
        for bb, nc, ne in slots:
            sc.submit(bb)

        def one_step():
            got = 0
            for i in range(CHUNKS):
                s, _h = sc.results_next()
                got += len(s)
                bb, nc, ne = slots[i]
                nb = sc.acquire()                  # the slot just drained: its pinned contents travel again
                assert nb.h.value == bb.h.value
                nb.set_size(nc, ne)
                sc.submit(nb)
                slots[i] = (nb, nc, ne)
            return got

        for _ in range(max(args.warmup, 3)):
            got = one_step()
        assert got == n_sites, "e2e path found %d sites, resident path %d" % (got, n_sites)
        dist_barrier(dist)
        sc.sync()
        sc.timer_start()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            one_step()
        ems = sc.timer_stop()
        wall_ms = (time.perf_counter() - t0) * 1e3
        dist_barrier(dist)
        for _ in range(CHUNKS):                # drain what is still in flight
            sc.results_next()
        sc.close()
        dev_ms, host_ms = ems, wall_ms
        ems = max(ems, wall_ms)                # host work (submit, sort of sites) counts too
        ems_max = dist_max(dist, ems)
        h2d_step = 3 * nent + 4 * (2 * ncol + CHUNKS) + ncol
        my_gbs = h2d_step * args.steps / (ems * 1e-3) / 1e9
        e2e = {"h2d_GBs_per_rank": {"this_box_sum": dist_sum(dist, my_gbs), "slowest_rank": -dist_max(dist, -my_gbs), "fastest_rank": dist_max(dist, my_gbs)},
               "value": total_cols * args.steps / (ems_max * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": 3 * nent + 4 * (2 * ncol + CHUNKS) + ncol,
               "d2h_bytes_per_step": 8 * CHUNKS + 216 * n_sites, "ms_per_step": ems_max / args.steps,
               "batches_per_step": CHUNKS, "device_ms_per_step": dev_ms / args.steps, "host_wall_ms_per_step": host_ms / args.steps}

    # ---- e2e_reads: the filtered reads themselves travel; the device builds the columns, then scores them ----
    e2e_reads = None
    if not args.no_e2e and not args.no_e2e_reads:
        from somatic_sniper_b200 import reads as rd
        n_win = max(1, int(round(ncol / CHUNKS / api.READ_WINDOW)))
        length = n_win * api.READ_WINDOW
        ref = rd.make_reference(shard_seed(rank) + 11, length + 1024)
        ref_t = ref.copy()
        ref_t[::5003] = np.where(ref_t[::5003] == 1, 2, 1)          # a tumor-only variant every 5 kb
        rw = rd.build_read_windows(rd.make_reads(shard_seed(rank) + 12, ref_t, args.depth_t, 150, 0, length, low_mapq_frac=0.02),
                                   rd.make_reads(shard_seed(rank) + 13, ref, args.depth_n, 150, 0, length, low_mapq_frac=0.02),
                                   ref, first_window=0, n_windows=n_win)
        sc = api.Scorer(api.default_params(), device=local, n_slots=CHUNKS, max_columns=length, max_entries=rw["entries_bound"])
        sc.reads_reserve(rw["recs"].size + 64, rw["rec_off"].size + 1, n_win)
        held = [sc.acquire() for _ in range(CHUNKS)]
        for bb in held:                        # the BAM reader's job: records land in pinned memory once
            rargs = sc.fill_reads(bb, rw)
        for bb in held:
            sc.release(bb)
        live = []
        for _ in range(CHUNKS):
            bb = sc.acquire()
            sc.submit_reads(bb, *rargs)
            live.append(bb)

        def reads_step():
            got = 0
            for i in range(CHUNKS):
                s, _h = sc.results_next()
                got += len(s)
                sc.release(live[i])
                live[i] = sc.acquire()
                sc.submit_reads(live[i], *rargs)
            return got

        for _ in range(max(args.warmup, 3)):
            r_sites = reads_step()
        dist_barrier(dist)
        sc.sync()
        sc.pileup_time(reset=True)
        sc.timer_start()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            reads_step()
        rms = sc.timer_stop()
        rwall = (time.perf_counter() - t0) * 1e3
        dist_barrier(dist)
        for i in range(CHUNKS):
            sc.results_next()
        r_pos, r_nent = sc.column_positions(live[0])
        p_ms, p_n = sc.pileup_time()
        sc.close()
        r_cols = int(r_pos.size) * CHUNKS
        rms_max = dist_max(dist, max(rms, rwall))
        r_total = dist_sum(dist, r_cols)
        h2d = CHUNKS * int(rw["recs"].size + 4 * rw["rec_off"].size + 24 * n_win + api.READ_WINDOW * n_win)
        r_gbs = h2d * args.steps / (max(rms, rwall) * 1e-3) / 1e9
        e2e_reads = {"h2d_GBs_per_rank": {"this_box_sum": dist_sum(dist, r_gbs), "slowest_rank": -dist_max(dist, -r_gbs), "fastest_rank": dist_max(dist, r_gbs)},
                     "value": r_total * args.steps / (rms_max * 1e-3), "unit": UNIT, "ms_per_step": rms_max / args.steps,
                     "columns_per_gpu_per_step": r_cols, "read_bases_per_step": int(r_nent) * CHUNKS, "sites_per_step": int(r_sites),
                     "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": CHUNKS * 32 + 4 * r_cols + 216 * int(r_sites) + 32 * CHUNKS,
                     "batches_per_step": CHUNKS, "windows_per_batch": n_win,
                     "device_ms_per_step": rms / args.steps, "host_wall_ms_per_step": rwall / args.steps,
                     "pileup_kernels_ms_per_batch": p_ms / max(p_n, 1),
                     "input": "filtered reads (ssn_read_rec + cigar + 4-bit bases + qualities, %.2f bytes per read-base) filed under "
                              "8192-position windows; ssn_batch_submit_reads builds the host packer's column batch on the device "
                              "(tests/test_pileup.py: byte-identical) and scores it" % (h2d / max(1, int(r_nent) * CHUNKS))}

    # ---- panel: BASELINE configs[3] depth (1500x / 1500x), resident, the product path and every-column path ----
    panel = None
    if not args.no_panel:
        # 16384 distinct columns (148 MB), laid out four times over: 65536 columns fill every resident warp of the grid
        unit = synth.make_columns(16384, depth_t=1500, depth_n=1500, seed=shard_seed(rank) + 500)
        reps, ue = 4, int(unit["n_entries"])
        pcols = dict(off=np.concatenate([unit["off"][:-1].astype(np.int64) + r * ue for r in range(reps)] + [np.array([reps * ue], np.int64)]).astype(np.uint32),
                     bs=np.tile(unit["bs"], reps), bq=np.tile(unit["bq"], reps), mq=np.tile(unit["mq"], reps),
                     ref_char=np.tile(unit["ref_char"], reps), ncol=reps * unit["ncol"], n_entries=reps * ue)
        p_alg = synth.algorithmic_bytes(pcols)
        sc = api.Scorer(api.default_params(), device=local, n_slots=1, max_columns=pcols["ncol"], max_entries=int(pcols["n_entries"]) + 64)
        pb = sc.acquire()
        pb.fill(pcols)
        sc.upload(pb)
        res = {}
        for name, flags in (("product", 0), ("every_column", api.SUBMIT_NO_PREFILTER)):
            for _ in range(3):
                sc.score_resident(pb, flags)
            sc.sync()
            sites_p = sc.fetch(pb)
            sc.kernel_time(reset=True)
            dist_barrier(dist)
            sc.timer_start()
            for _ in range(max(3, args.steps // 2)):
                sc.score_resident(pb, flags)
            p_ms = sc.timer_stop() / max(3, args.steps // 2)
            dist_barrier(dist)
            ppf, psc, pkn = sc.kernel_time_split(reset=True)
            res[name] = (p_ms, ppf / max(pkn, 1), psc / max(pkn, 1), sites_p)
        assert res["product"][3].tobytes() == res["every_column"][3].tobytes()
        sc.close()
        pm = dist_max(dist, res["product"][0])
        ptot = dist_sum(dist, float(pcols["ncol"]))
        panel = {"value": ptot / (pm * 1e-3), "unit": UNIT, "ms_per_step": pm,
                 "workload": "synthetic targeted panel, 65536 columns (16384 distinct, four times over) at 1500x / 1500x (BASELINE configs[3] depth), %d read-bases, resident" % int(pcols["n_entries"]),
                 "sites_per_step": int(len(res["product"][3])),
                 "kernels_ms": {"ssn_prefilter_kernel": res["product"][1], "ssn_score_kernel<false>": res["product"][2]},
                 "every_column_ms_per_step": res["every_column"][0],
                 "roofline": {"bound": "hbm", "achieved": p_alg / (pm * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                              "frac": p_alg / (pm * 1e-3) / 1e9 / peak,
                              "traffic": (int(ncu_traffic("panel_dram_bytes_per_launch") * (p_alg / 295195500.0)) if ncu_traffic("panel_dram_bytes_per_launch") else None),
                              "traffic_note": "ncu capture of half this batch (32768 columns, 295.2 MB algorithmic, 329.1 MB DRAM), scaled by size",
                              "kernel": "ssn_score_kernel<false> (deep columns)",
                              "algorithmic_bytes_per_launch": p_alg, "ceiling_columns_per_s": peak * 1e9 / (p_alg / pcols["ncol"])},
                 "l2": "inputs (%.0f MB per step) exceed the 126 MB L2; no flush" % (p_alg / 1e6)}

    # ---- e2e_bam: the whole program on BAM files (region shards -> inflate -> reads -> device pileup -> kernels -> writer) ----
    e2e_bam = None
    if not args.no_e2e_bam:
        dist_barrier(dist)
        if rank == 0:
            e2e_bam = bam_level(args)
        dist_barrier(dist)

    sampler.stop()
    if rank != 0:
        return
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64/f32+int", "data": "synthetic",
            "config": workload_config(ncol, args.depth_t, args.depth_n, alg_bytes),
            "run": {"read_bases_per_step": nent, "sites_per_step": n_sites,
                    "prefilter": "exact hom-reference pre-filter on; %d of %d columns go on to the full model" % (survivors, ncol)
                                 if filtered else "off",
                    "resident_batches_in_flight": STREAMS, "parallelism": "region shards, %d GPU(s), no collective" % args.gpus,
                    "cpu_binding": binding, "timed_region_repeats": REPEATS, "ms_per_step_each_repeat": [m / args.steps for m in rep_ms],
                    "value_spread": (max(rep_ms) - min(rep_ms)) / ms},
            "clocks": sampler.summary(),
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": dom_traffic, "peak_source": peak_src, "kernel": dom_name,
                         "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms_per_launch": dom_ms,
                         "step_kernels_ms": {"ssn_prefilter_kernel": pf_ms / max(kn, 1), "ssn_score_kernel<false> (survivors)": score_ms / max(kn, 1)},
                         "step_frac": alg_bytes / (k_avg_ms * 1e-3) / 1e9 / peak,
                         "measured_in": "single_stream pass (kernels back to back); with %d batches in flight the kernels share the SMs: "
                                        "pre-filter %.4f ms, survivors %.4f ms per launch" % (STREAMS, ov_pf_ms / max(ov_kn, 1), ov_score_ms / max(ov_kn, 1)),
                         "whole_step_frac": alg_bytes / (ms_max / args.steps * 1e-3) / 1e9 / peak},
            "single_stream": {"value": single_value, "unit": UNIT, "ms_per_step": single_ms_max / args.steps},
            "full_model": {"value": full_value, "unit": UNIT, "ms_per_step": full_ms_max / args.steps,
                           "note": "SSN_SUBMIT_NO_PREFILTER: every column through ssn_score_kernel<false>; identical sites",
                           "roofline": {"bound": "hbm", "achieved": alg_bytes / (full_k_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                        "frac": alg_bytes / (full_k_ms * 1e-3) / 1e9 / peak, "traffic": ncu_traffic(),
                                        "kernel": "ssn_score_kernel<false>", "kernel_ms_per_launch": full_k_ms}},
            }
    # The headline e2e is the route the shipped program takes by default since round 2: the filtered reads travel and the
    # device builds the columns (ssn_batch_submit_reads).  The packed-column route stays beside it as e2e_columns.
    if e2e_reads is not None:
        line["e2e"] = dict(e2e_reads, route="reads (ssn_batch_submit_reads): the CLI's default; e2e_columns = packed columns (ssn_batch_submit)")
        line["e2e_reads"] = e2e_reads
        if e2e is not None:
            line["e2e_columns"] = e2e
    elif e2e is not None:
        line["e2e"] = dict(e2e, route="packed columns (ssn_batch_submit)")
    if panel is not None:
        line["panel"] = panel
    if e2e_bam is not None:
        line["e2e_bam"] = e2e_bam
    if args.gpus == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_single(cols, args.cpu_sample)
    print(json.dumps(line), file=_JSON_OUT, flush=True)
    if dist is not None:
        dist.destroy_process_group()


_JSON_OUT = sys.stdout


def main():
    # stdout carries exactly one JSON line: libraries that write to fd 1 (NCCL prints its version there) go to stderr
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
