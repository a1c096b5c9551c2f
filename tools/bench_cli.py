#!/usr/bin/env python
"""BAM-level end-to-end timing (BASELINE.json configs[1]-style): this repo's bam-somaticsniper (BGZF inflate -> packer ->
H2D -> kernel -> writers) against the unmodified reference binary (oracle/_ref, single process) on the same synthetic
tumor/normal pair, with a byte comparison of the two output files.
    python tools/bench_cli.py --len 1000000 [--depth-t 60 --depth-n 30] [--args "-J -F vcf"] [--out gpurun_out/cli_bench.json]"""
import argparse, json, os, shlex, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from somatic_sniper_b200 import cli_tools as ct


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--len", type=int, default=1_000_000)
    ap.add_argument("--depth-t", type=float, default=60.0)
    ap.add_argument("--depth-n", type=float, default=30.0)
    ap.add_argument("--targets", type=int, default=0, help="panel mode: this many 5 kb targets instead of one contig")
    ap.add_argument("--args", default="")
    ap.add_argument("--repeat", type=int, default=3)
    ap.add_argument("--no-reference", action="store_true")
    ap.add_argument("--threads", default=None)
    ap.add_argument("--device-pileup", action="store_true", help="SSN_DEVICE_PILEUP=1: ship the reads, build the columns on the GPU")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    w = tempfile.mkdtemp(prefix="ssn_bench_")
    t0 = time.perf_counter()
    contigs = [("chr1", a.len)] if not a.targets else [("t%03d" % i, 5000) for i in range(a.targets)]
    pair = ct.make_pair(w, contigs, a.depth_t, a.depth_n, seed=1234, n_runs=0, special=True)
    t_sim = time.perf_counter() - t0
    opts = shlex.split(a.args)
    env = {"SSN_THREADS": a.threads} if a.threads else {}
    if a.device_pileup:
        env["SSN_DEVICE_PILEUP"] = 1
    runs = []
    for i in range(a.repeat):
        mine = os.path.join(w, "mine.out")
        r = ct.run_cli(ct.CLI, pair["fasta"], pair["tumor"], pair["normal"], mine, opts, env=env, stats=os.path.join(w, "stats.json"))
        assert r["rc"] == 0, r["stderr"]
        runs.append(dict(wall_s=r["wall_s"], stats=r["stats"]))
    best = min(runs, key=lambda x: x["wall_s"])
    st = best["stats"]
    res = dict(workload="synthetic %d bp, tumor %gx / normal %gx, 2x150 reads, args '%s'" % (pair["columns"], a.depth_t, a.depth_n, a.args),
               reads=[pair["reads_tumor"], pair["reads_normal"]], bam_bytes=[os.path.getsize(pair["tumor"]), os.path.getsize(pair["normal"])],
               simulate_s=t_sim, b200=dict(wall_s=best["wall_s"], setup_s=st["setup_s"], stream_s=st["wall_s"] - st["setup_s"], columns=st["columns"],
                                           columns_per_s_stream=st["columns_per_s"], columns_per_s_wall=st["columns"] / best["wall_s"],
                                           all_walls=[x["wall_s"] for x in runs], stats=st))
    if not a.no_reference and ct.reference_available():
        ref = os.path.join(w, "ref.out")
        r = ct.run_cli(ct.REF_CLI, pair["fasta"], pair["tumor"], pair["normal"], ref, opts)
        same = ct.comparable(ref) == ct.comparable(os.path.join(w, "mine.out"))
        res["reference"] = dict(wall_s=r["wall_s"], columns_per_s_wall=st["columns"] / r["wall_s"], processes=1, rc=r["rc"])
        res["identical_output"] = bool(same)
        res["output_lines"] = sum(1 for _ in open(ref, "rb"))
        res["speedup_wall"] = r["wall_s"] / best["wall_s"]
    print(json.dumps(res), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
