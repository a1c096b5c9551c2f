#!/usr/bin/env python
"""Runs this repo's bam-somaticsniper (GPU) and the unmodified reference binary (oracle/_ref, test infrastructure) on the
same synthetic tumor/normal BAM pair for a set of option strings and compares the output files byte for byte.
    python tools/cli_parity.py [--len 200000] [--out gpurun_out/cli_parity.json]"""
import argparse, json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from somatic_sniper_b200 import cli_tools as ct

OPTION_SETS = [[], ["-F", "vcf"], ["-F", "bed"], ["-J", "-F", "vcf"], ["-s", "0.001", "-Q", "0", "-F", "vcf"],
               ["-q", "1", "-Q", "0", "-L", "-G"], ["-p", "-Q", "0"], ["-T", "0.7", "-N", "3", "-r", "0.01", "-Q", "5"], ["-q", "30", "-F", "vcf", "-n", "nrm", "-t", "tum"]]

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--len", type=int, default=200000)
    ap.add_argument("--out", default=None)
    ap.add_argument("--devices", default=None)
    a = ap.parse_args()
    w = tempfile.mkdtemp(prefix="ssn_cli_")
    pair = ct.make_pair(w, [("chr1", a.len), ("chr2", a.len // 4), ("chrU", 300), ("chr3", a.len // 2)], seed=11, n_runs=2,
                        variant_kw=dict(het_every=1000, hom_every=5000, som_every=2000, loh_every=10000, gor_every=10000))
    res = []
    ok = True
    for i, opts in enumerate(OPTION_SETS):
        mine, ref = os.path.join(w, "mine_%d.out" % i), os.path.join(w, "ref_%d.out" % i)
        env = {"SSN_DEVICES": a.devices} if a.devices else None
        r1 = ct.run_cli(ct.CLI, pair["fasta"], pair["tumor"], pair["normal"], mine, opts, env=env, stats=os.path.join(w, "stats.json"))
        r2 = ct.run_cli(ct.REF_CLI, pair["fasta"], pair["tumor"], pair["normal"], ref, opts)
        same = r1["rc"] == 0 and r2["rc"] == 0 and ct.comparable(mine) == ct.comparable(ref)
        ok &= same
        nlines = sum(1 for _ in open(ref, "rb"))
        res.append(dict(opts=opts, identical=same, rc=[r1["rc"], r2["rc"]], lines=nlines, wall_s=[r1["wall_s"], r2["wall_s"]], stats=r1["stats"]))
        print(("OK  " if same else "DIFF"), " ".join(opts) or "(default)", "lines", nlines, "wall %.2f / %.2f s" % (r1["wall_s"], r2["wall_s"]), flush=True)
        if not same:
            print(r1["stderr"][-2000:])
    if a.out:
        with open(a.out, "w") as f:
            json.dump(dict(workload="synthetic %d bp over 4 contigs, 60x/30x" % (a.len + a.len // 4 + 300 + a.len // 2), results=res), f, indent=1)
    sys.exit(0 if ok else 1)

if __name__ == "__main__":
    main()
