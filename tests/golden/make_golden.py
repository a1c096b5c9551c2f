#!/usr/bin/env python
"""Generates the committed golden fixtures from the UNMODIFIED reference.

Run in the build container only (needs /root/reference and oracle/_ref/, see oracle/build_ref.sh):

    python tests/golden/make_golden.py

Outputs (all small, committed):
    tests/golden/columns_in.npz        seeded SoA column batch (synth.make_columns + KATs + deep/ragged columns)
    tests/golden/columns_<cfg>.npz     what the reference's glf_somatic / sniper_maqcns_glfgen produced for it
    tests/golden/tables.json           SHA-256 of every model table's bytes per parameter set + spot values
    tests/golden/tables_small.npz      the small integer tables verbatim (prior, jointprior, qadd, fk, q_r)
    tests/golden/integration.npz       the pileup columns of integration-test/data at the three expected.vcf
                                       sites (built from `samtools view` text of the reference's own BAMs)
    tests/golden/integration/          the reference's tiny end-to-end fixtures (BAM pair, FASTA, expected.vcf)
                                       plus classic/bed/-J outputs produced by the reference binary here
Nothing here is imported by the product.
"""
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import loader  # noqa: E402
from somatic_sniper_b200 import synth  # noqa: E402
from conftest import CONFIGS  # noqa: E402

REF_DATA = "/root/reference/integration-test/data"


def golden_columns():
    """One mixed batch: KATs, random 30x/60x with and without edge cases, ragged depths, a few deep columns."""
    parts = [synth.kat_columns(),
             synth.make_columns(700, seed=101, p_som=1e-2, p_het=1e-2, p_loh=5e-3, p_gor=5e-3),
             synth.make_columns(500, seed=102, binned=False, edge=1.0, p_som=1e-2, p_het=1e-2, p_loh=5e-3, p_gor=5e-3),
             synth.make_columns(300, seed=103, depth_t=8, depth_n=5, edge=0.5, p_som=2e-2, p_het=2e-2),
             synth.make_columns(24, seed=104, depth_t=1500, depth_n=1100, binned=False, p_som=5e-2, p_het=5e-2),
             synth.make_columns(40, seed=105, depth_t=130, depth_n=126, edge=0.3, p_som=5e-2)]
    off = [np.zeros(1, np.int64)]
    base = 0
    for p in parts:
        o = p["off"].astype(np.int64)
        off.append(o[1:] + base)
        base += int(o[-1])
    cat = lambda k: np.concatenate([p[k] for p in parts])
    return dict(off=np.concatenate(off).astype(np.uint32), bs=cat("bs"), bq=cat("bq"), mq=cat("mq"),
                ref_char=cat("ref_char"), ncol=sum(p["ncol"] for p in parts), n_entries=base)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def table_sets():
    return {"default": dict(), "TNr": dict(theta=0.7, n_hap=3, het_rate=0.01),
            "joint_s001": dict(use_joint=1, somatic_rate=0.001), "nopriors": dict(use_priors=0),
            "N4": dict(n_hap=4, theta=0.9)}


def run_child(cfg_kind, name, out_path):
    """The reference keeps global state; every parameter set runs in its own process."""
    subprocess.run([sys.executable, os.path.abspath(__file__), "--child", cfg_kind, name, out_path], check=True)


def child(cfg_kind, name, out_path):
    if cfg_kind == "columns":
        cols = dict(np.load(os.path.join(HERE, "columns_in.npz")))
        ref = loader.Reference(**CONFIGS[name])
        r = ref.score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"], want_glf=True)
        np.savez_compressed(out_path, score=r["score"], emit_col=r["emit_col"], recs=r["recs"], glf=r["glf"])
    elif cfg_kind == "tables":
        ref = loader.Reference(**table_sets()[name])
        t = ref.tables()
        np.savez_compressed(out_path, **t)
    elif cfg_kind == "integration":
        cols = dict(np.load(os.path.join(HERE, "integration.npz")))
        ref = loader.Reference(**CONFIGS[name])
        r = ref.score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"], want_glf=True)
        np.savez_compressed(out_path, score=r["score"], emit_col=r["emit_col"], recs=r["recs"], glf=r["glf"])


def sam_pileup(bam, fasta_seq):
    """Pileup of a BAM whose reads are plain M-only alignments (the reference's integration data), from SAM text."""
    txt = subprocess.run([loader.REF_SAMTOOLS, "view", bam], check=True, capture_output=True, text=True).stdout
    cols = {}
    for line in txt.splitlines():
        f = line.split("\t")
        flag, pos, mapq, cigar, seq, qual = int(f[1]), int(f[3]) - 1, int(f[4]), f[5], f[9], f[10]
        assert cigar == "%dM" % len(seq), cigar
        if flag & (4 | 256 | 512 | 1024):
            continue
        for i, (b, q) in enumerate(zip(seq, qual)):
            cols.setdefault(pos + i, []).append((b.upper() if b.upper() in "ACGTN=" else "N", ord(q) - 33, mapq, 1 if flag & 16 else 0))
    return cols


def integration():
    dst = os.path.join(HERE, "integration")
    os.makedirs(dst, exist_ok=True)
    for f in ("t-small.bam", "n-small.bam", "small.fa", "expected.vcf"):
        shutil.copyfile(os.path.join(REF_DATA, f), os.path.join(dst, f))
    seq = "".join(l.strip() for l in open(os.path.join(dst, "small.fa")) if not l.startswith(">"))
    t = sam_pileup(os.path.join(dst, "t-small.bam"), seq)
    n = sam_pileup(os.path.join(dst, "n-small.bam"), seq)
    # NB the reference drops nothing here: one contig only (QUIRK B1 needs a second contig)
    positions = sorted(set(t) & set(n))
    batch = synth.columns_from_lists([(seq[p] if p < len(seq) else "N", t[p], n[p]) for p in positions])
    np.savez_compressed(os.path.join(HERE, "integration.npz"), pos=np.array(positions, np.int32),
                        **{k: batch[k] for k in ("off", "bs", "bq", "mq", "ref_char")})
    # other output formats / options from the reference binary itself
    with tempfile.TemporaryDirectory() as tmp:
        fa = os.path.join(tmp, "small.fa")
        shutil.copyfile(os.path.join(dst, "small.fa"), fa)
        runs = {"classic": ["-F", "classic"], "bed": ["-F", "bed"], "vcf": ["-F", "vcf"],
                "vcf_J_Q0": ["-F", "vcf", "-J", "-Q", "0"], "classic_J_Q0": ["-F", "classic", "-J", "-Q", "0"],
                "classic_q1_Q0_LG": ["-F", "classic", "-q", "1", "-Q", "0", "-L", "-G"],
                "vcf_p_Q0_names": ["-F", "vcf", "-p", "-Q", "0", "-n", "NRM", "-t", "TMR"],
                "classic_T07_N3_r01": ["-F", "classic", "-T", "0.7", "-N", "3", "-r", "0.01", "-Q", "0"],
                "vcf_s0001": ["-F", "vcf", "-s", "0.001", "-Q", "0"]}
        for name, opts in runs.items():
            out = os.path.join(tmp, name)
            subprocess.run([loader.REF_CLI] + opts + ["-f", fa, os.path.join(dst, "t-small.bam"),
                            os.path.join(dst, "n-small.bam"), out], check=True, capture_output=True)
            lines = [l for l in open(out) if not (l.startswith("##fileDate") or l.startswith("##reference="))]
            with open(os.path.join(dst, "ref_" + name + ".out"), "w") as fh:
                fh.writelines(lines)
        with open(os.path.join(dst, "ref_runs.json"), "w") as fh:
            json.dump(runs, fh, indent=1)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--child":
        child(*sys.argv[2:5])
        return
    assert loader.Reference.available(), "build oracle/_ref first (oracle/build_ref.sh)"
    cols = golden_columns()
    np.savez_compressed(os.path.join(HERE, "columns_in.npz"), **{k: cols[k] for k in ("off", "bs", "bq", "mq", "ref_char")})
    for cfg in CONFIGS:
        run_child("columns", cfg, os.path.join(HERE, "columns_%s.npz" % cfg))
    integration()
    for cfg in ("default", "joint_s", "nopriors_q0"):
        run_child("integration", cfg, os.path.join(HERE, "integration_%s.npz" % cfg))
    meta, small = {}, {}
    with tempfile.TemporaryDirectory() as tmp:
        for name, kw in table_sets().items():
            p = os.path.join(tmp, name + ".npz")
            run_child("tables", name, p)
            t = dict(np.load(p))
            meta[name] = dict(params=kw, q_r=float(t["q_r"]), coef_sha256=sha(t["coef"]), lhet_sha256=sha(t["lhet"]),
                              fk_sha256=sha(t["fk"]), coef_nonzero=int(np.count_nonzero(t["coef"])),
                              coef_20_30_3=float(t["coef"][20 << 16 | 30 << 8 | 3]),
                              lhet_10_10=float(t["lhet"][10 << 8 | 10]), lhet_30_0=float(t["lhet"][30 << 8]))
            for k in ("prior", "jointprior", "qadd", "fk"):
                small["%s__%s" % (name, k)] = t[k]
    with open(os.path.join(HERE, "tables.json"), "w") as fh:
        json.dump(meta, fh, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "tables_small.npz"), **small)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
