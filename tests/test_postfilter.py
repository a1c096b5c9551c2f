"""somatic_sniper_b200/postfilter.py against the reference's Perl post-filters (src/scripts/snpfilter.pl,
highconfidence.pl) on classic and VCF output of the unmodified reference binary: same kept file,
same low-quality file, for the default options and a few others.  Golden outputs of the Perl scripts are committed
(tests/golden/postfilter/, made by this file run as a script in the container that holds /root/reference)."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "postfilter")
SCRIPTS = "/root/reference/src/scripts"
CASES = {
    "snpfilter_default": ("snpfilter", []),
    "snpfilter_loose": ("snpfilter", ["--min-mapping-quality", "20", "--min-cns-qual", "5", "--snp-win-size", "50", "--max-snp-per-win", "1", "--include-loh"]),
    "snpfilter_tumor_only": ("snpfilter", ["--tumor-variant-only", "--min-read-depth", "20", "--max-read-depth", "70"]),
    "snpfilter_indels": ("snpfilter", ["--indel-file", "INDELS", "--indel-win-size", "200"]),
    "highconfidence_default": ("highconfidence", []),
    "highconfidence_loose": ("highconfidence", ["--min-mapping-quality", "30", "--min-somatic-score", "20"]),
}


def _run_mine(cmd, snp, args, out, lq):
    from somatic_sniper_b200 import postfilter
    postfilter.main([cmd, "--snp-file", snp, "--out-file", out, "--lq-output", lq] + args)


@pytest.mark.parametrize("fmt", ["classic", "vcf"])
@pytest.mark.parametrize("case", sorted(CASES))
def test_matches_the_perl_scripts_golden_output(case, fmt, tmp_path):
    cmd, args = CASES[case]
    snp = os.path.join(GOLD, "input.%s" % fmt)
    args = [os.path.join(GOLD, "indels.pileup") if a == "INDELS" else a for a in args]
    out, lq = str(tmp_path / "out"), str(tmp_path / "lq")
    _run_mine(cmd, snp, args, out, lq)
    assert open(out).read() == open(os.path.join(GOLD, "%s.%s.out" % (case, fmt))).read()
    assert open(lq).read() == open(os.path.join(GOLD, "%s.%s.lq" % (case, fmt))).read()
    assert sum(1 for l in open(out) if not l.startswith("#")) + sum(1 for _ in open(lq)) == sum(1 for l in open(snp) if not l.startswith("#"))


def test_live_perl_scripts_when_present(tmp_path):
    if not (os.path.exists(SCRIPTS) and shutil.which("perl")):
        pytest.skip("the reference's scripts (or perl) are not in this environment")
    for fmt in ("classic", "vcf"):
        snp = os.path.join(GOLD, "input.%s" % fmt)
        for case, (cmd, args) in CASES.items():
            args = [os.path.join(GOLD, "indels.pileup") if a == "INDELS" else a for a in args]
            o1, l1, o2, l2 = (str(tmp_path / x) for x in ("o1", "l1", "o2", "l2"))
            subprocess.run(["perl", os.path.join(SCRIPTS, cmd + ".pl"), "--snp-file", snp, "--out-file", o1, "--lq-output", l1] + args, check=True)
            _run_mine(cmd, snp, args, o2, l2)
            assert open(o1).read() == open(o2).read() and open(l1).read() == open(l2).read(), (case, fmt)


if __name__ == "__main__":
    # regenerate the goldens: reference binary on a simulated pair -> input.classic / input.vcf, Perl scripts -> *.out / *.lq
    sys.path.insert(0, ROOT)
    import tempfile
    from somatic_sniper_b200 import cli_tools as ct
    os.makedirs(GOLD, exist_ok=True)
    d = tempfile.mkdtemp()
    pair = ct.make_pair_fast(d, [("chr1", 400000), ("chr2", 150000)], 60, 30, seed=99)
    for fmt in ("classic", "vcf"):
        ct.run_cli(ct.REF_CLI, pair["fasta"], pair["tumor"], pair["normal"], os.path.join(GOLD, "input.%s" % fmt), ["-F", fmt, "-Q", "10"])
    with open(os.path.join(GOLD, "input.vcf")) as f:
        lines = [l for l in f if not (l.startswith("##fileDate") or l.startswith("##reference="))]
    open(os.path.join(GOLD, "input.vcf"), "w").writelines(lines)
    sites = [l.split("\t") for l in open(os.path.join(GOLD, "input.classic"))]
    with open(os.path.join(GOLD, "indels.pileup"), "w") as f:       # samtools pileup -i style indel lines near every 7th site
        for i, s in enumerate(sites[::7]):
            f.write("%s\t%d\t*\t%s\t60\t%d\t60\t30\t+A\t*\t5\t25\t0\t0\t0\n" % (s[0], int(s[1]) + (i % 5) * 60 - 100, "*/+A" if i % 3 else "*/*", 40 + 7 * (i % 4)))
    for fmt in ("classic", "vcf"):
        snp = os.path.join(GOLD, "input.%s" % fmt)
        for case, (cmd, args) in CASES.items():
            args = [os.path.join(GOLD, "indels.pileup") if a == "INDELS" else a for a in args]
            subprocess.run(["perl", os.path.join(SCRIPTS, cmd + ".pl"), "--snp-file", snp, "--out-file", os.path.join(GOLD, "%s.%s.out" % (case, fmt)),
                            "--lq-output", os.path.join(GOLD, "%s.%s.lq" % (case, fmt))] + args, check=True)
    print("goldens written to", GOLD, sum(os.path.getsize(os.path.join(GOLD, f)) for f in os.listdir(GOLD)), "bytes")
