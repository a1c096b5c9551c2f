#!/usr/bin/env python
"""Post-filters on bam-somaticsniper output, with the options and the line-by-line behaviour of the reference's Perl scripts
(SURVEY.md §8(f) row 4 — downstream text filters, not part of the scored path):

    python -m somatic_sniper_b200.postfilter snpfilter      --snp-file F [--indel-file I] [options]   src/scripts/snpfilter.pl
    python -m somatic_sniper_b200.postfilter highconfidence --snp-file F [options]                    src/scripts/highconfidence.pl

Both read classic or VCF output (auto-detected from a `##fileformat=VCF` line, snpfilter.pl:118-121), pass comment lines
through and split the sites into the kept file and an optional low-quality file.  Defaults as in the scripts."""
import argparse
import sys


def _vcf_keys(fields, sample_col):
    keys = fields[8].split(":")
    vals = fields[sample_col].split(":")
    return {k: (vals[i] if i < len(vals) else None) for i, k in enumerate(keys)}


_IUB = {"A": "AA", "C": "CC", "G": "GG", "T": "TT", "M": "AC", "K": "GT", "Y": "CT", "R": "AG", "W": "AT", "S": "CG",
        "D": "AGT", "B": "CGT", "H": "ACT", "V": "ACG", "N": "ACGT"}


def _is_loh(tumor, normal):
    """snpfilter.pl:is_loh — the tumor genotype's alleles are a proper subset of the normal's (classic output, IUB codes)"""
    t, n = _IUB.get(tumor), _IUB.get(normal)
    if t is None or n is None or tumor == normal:
        return False
    ts, ns = set(t), set(n)
    return ts < ns or (ts <= ns and len(ts) < len(ns))


def snpfilter(a):
    indel = set()
    if a.indel_file:
        import re
        pat = re.compile(r"^(\S+)\s+(\S+)\s+(\S+)\s+(\S+)\s+\S+\s+(\S+)\s+")
        for line in open(a.indel_file):
            m = pat.match(line)
            if not m:
                continue
            chrom, pos, ident, seq, score = m.groups()
            if ident != "*" or seq == "*/*" or float(score) < a.min_indel_score:
                continue
            indel.add((chrom, int(pos)))
    out = open(a.out_file or a.snp_file + ".SNPfilter", "w")
    lq = open(a.lq_output, "w") if a.lq_output else None
    vcf = False
    win, last_chr = [], ""

    def flush_all():
        for s in win:
            if s["pass"]:
                out.write(s["line"])
            elif lq:
                lq.write(s["line"])

    for raw in open(a.snp_file):
        if raw.startswith("##fileformat=VCF"):
            vcf = True
        if raw.startswith("#"):
            out.write(raw)
            continue
        snp = raw.rstrip("\n")
        f = snp.split("\t")
        tumor_gt = somatic_status = normal_var = None
        if vcf:
            chrom, pos, ref, var = f[0], int(f[1]), f[3], f[4]
            k = _vcf_keys(f, 10)                              # samples: NORMAL, TUMOR -> the tumor is the second one
            cns_qual, snp_qual, map_qual, depth = float(k["GQ"]), float(k["VAQ"]), float(k["MQ"]), float(k["DP"])
            tumor_gt, somatic_status = k["GT"], k["SS"]
        else:
            chrom, pos, ref, var, normal_var = f[0], int(f[1]), f[2], f[3], f[4]
            cns_qual, snp_qual, map_qual, depth = float(f[6]), float(f[7]), float(f[8]), float(f[12])
        if any((chrom, p) in indel for p in range(pos - a.indel_win_size, pos + a.indel_win_size + 1)):
            if lq:
                lq.write(snp + "\n")
            continue
        ok = map_qual >= a.min_mapping_quality and a.min_read_depth <= depth <= a.max_read_depth
        if not (cns_qual >= a.min_cns_qual or snp_qual >= a.min_snp_qual):
            ok = False
        if a.tumor_variant_only and ((tumor_gt is not None and tumor_gt == "0/0") or (tumor_gt is None and var == ref)):
            ok = False
        if not a.include_loh and ((somatic_status is not None and somatic_status == "3") or
                                  (somatic_status is None and _is_loh(var, normal_var))):
            ok = False
        if not ok:
            if lq:
                lq.write(snp + "\n")
            continue
        if chrom != last_chr:
            flush_all()
            win, last_chr = [], chrom
        win.append(dict(line=snp + "\n", pos=pos, **{"pass": True}))
        if len(win) - 1 == a.max_snp_per_win:
            if win[-1]["pos"] - win[0]["pos"] < a.snp_win_size:
                for s in win:
                    s["pass"] = False
            if win[0]["pass"]:
                out.write(win[0]["line"])
            elif lq:
                lq.write(win[0]["line"])
            win.pop(0)
    flush_all()
    out.close()
    if lq:
        lq.close()


def highconfidence(a):
    out = open(a.out_file or a.snp_file + ".hc", "w")
    lq = open(a.lq_output, "w") if a.lq_output else None
    vcf = False
    for raw in open(a.snp_file):
        if raw.startswith("##fileformat=VCF"):
            vcf = True
        if raw.startswith("#"):
            out.write(raw)
            continue
        snp = raw.rstrip("\n")
        f = snp.split("\t")
        if vcf:
            ref, alts = f[3], f[4].split(",")
            k = _vcf_keys(f, 10)
            alleles = [ref] + alts
            used = sorted({alleles[int(i)] for i in k["GT"].split("/")})
            amq = dict(zip(used, k["AMQ"].split(",")))
            amq.pop(ref, None)
            mapqs, score = [float(x) for x in amq.values()], float(k["SSC"])
        else:
            mapqs, score = [float(x) for x in f[18].split(",")], float(f[5])
        ok = any(m >= a.min_mapping_quality for m in mapqs) and score >= a.min_somatic_score
        if ok:
            out.write(snp + "\n")
        elif lq:
            lq.write(snp + "\n")
    out.close()
    if lq:
        lq.close()


def main(argv=None):
    ap = argparse.ArgumentParser(prog="postfilter")
    sub = ap.add_subparsers(dest="cmd", required=True)
    s = sub.add_parser("snpfilter")
    s.add_argument("--snp-file", required=True); s.add_argument("--lq-output"); s.add_argument("--out-file"); s.add_argument("--indel-file")
    s.add_argument("--min-mapping-quality", type=int, default=40); s.add_argument("--min-cns-qual", type=int, default=20)
    s.add_argument("--min-read-depth", type=int, default=3); s.add_argument("--max-read-depth", type=int, default=100_000_000)
    s.add_argument("--snp-win-size", type=int, default=10); s.add_argument("--max-snp-per-win", type=int, default=2)
    s.add_argument("--min-snp-qual", type=int, default=20); s.add_argument("--indel-win-size", type=int, default=10)
    s.add_argument("--min-indel-score", type=int, default=50)
    s.add_argument("--tumor-variant-only", action="store_true"); s.add_argument("--include-loh", action="store_true")
    h = sub.add_parser("highconfidence")
    h.add_argument("--snp-file", required=True); h.add_argument("--lq-output"); h.add_argument("--out-file")
    h.add_argument("--min-mapping-quality", type=int, default=40); h.add_argument("--min-somatic-score", type=int, default=40)
    a = ap.parse_args(argv)
    (snpfilter if a.cmd == "snpfilter" else highconfidence)(a)


if __name__ == "__main__":
    main()
