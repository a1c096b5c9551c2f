"""Helpers around the two command line tools: this repo's bam-somaticsniper (GPU) and — test infrastructure only — the
unmodified reference binary built under oracle/_ref.  Used by tests/ and tools/bench_cli.py; nothing here scores."""
import json
import os
import subprocess
import time

from . import simbam

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
CLI = os.path.join(_HERE, "bin", "bam-somaticsniper")
REF_CLI = os.path.join(ROOT, "oracle", "_ref", "bam-somaticsniper")


def reference_available():
    return os.path.exists(REF_CLI)


def make_pair(workdir, contigs, depth_t=60.0, depth_n=30.0, seed=7, n_runs=1, level=6, variant_kw=None, special=True):
    """Writes ref.fa, tumor.bam, normal.bam under workdir; returns their paths."""
    os.makedirs(workdir, exist_ok=True)
    ref = simbam.make_reference(contigs, seed=seed, n_runs=n_runs)
    fa = os.path.join(workdir, "ref.fa")
    simbam.write_fasta(fa, ref)
    if os.path.exists(fa + ".fai"):
        os.remove(fa + ".fai")
    var = simbam.make_variants(ref, seed=seed + 1, **(variant_kw or {}))
    t, n = os.path.join(workdir, "tumor.bam"), os.path.join(workdir, "normal.bam")
    nt = simbam.simulate_bam(t, ref, var, "tumor", depth_t, seed=seed + 2, level=level, special=special)
    nn = simbam.simulate_bam(n, ref, var, "normal", depth_n, seed=seed + 3, level=level, special=special)
    return dict(fasta=fa, tumor=t, normal=n, reads_tumor=nt, reads_normal=nn, columns=sum(l for _, l in contigs))


SIMBAM = os.path.join(_HERE, "bin", "ssn-simbam")


def make_pair_fast(workdir, contigs, depth_t=60.0, depth_n=30.0, seed=7, level=1, threads=None):
    """The same kind of pair from the multi-threaded C generator (tools/simbam_fast.c): for inputs a Python loop cannot
    produce in reasonable time.  Returns the paths and read counts."""
    os.makedirs(workdir, exist_ok=True)
    threads = threads or max(1, (os.cpu_count() or 2) - 1)
    p = subprocess.run([SIMBAM, workdir, str(seed), str(depth_t), str(depth_n), str(level), str(threads)] +
                       ["%s:%d" % c for c in contigs], stdout=subprocess.PIPE, check=True)
    r = json.loads(p.stdout.decode())
    return dict(fasta=os.path.join(workdir, "ref.fa"), tumor=os.path.join(workdir, "tumor.bam"), normal=os.path.join(workdir, "normal.bam"),
                reads_tumor=r["reads_tumor"], reads_normal=r["reads_normal"], columns=sum(l for _, l in contigs))


def run_cli(exe, fasta, tumor, normal, out, args=(), env=None, stats=None, timeout=3600):
    e = dict(os.environ)
    if env:
        e.update({k: str(v) for k, v in env.items()})
    if stats:
        e["SSN_STATS"] = stats
    t = time.perf_counter()
    p = subprocess.run([exe, "-f", fasta] + list(args) + [tumor, normal, out], env=e, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, timeout=timeout)
    dt = time.perf_counter() - t
    st = None
    if stats and os.path.exists(stats):
        with open(stats) as f:
            st = json.loads(f.read())
    return dict(rc=p.returncode, stderr=p.stderr.decode(errors="replace"), stdout=p.stdout.decode(errors="replace"), wall_s=dt, stats=st)


def comparable(path):
    """Output bytes minus the two VCF header lines that carry the date and the FASTA path (integration-test/test_sniper.py:8-19)."""
    with open(path, "rb") as f:
        return b"".join(l for l in f if not (l.startswith(b"##fileDate") or l.startswith(b"##reference=")))
