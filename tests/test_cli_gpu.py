"""GPU tests of the whole drop-in: this repo's bam-somaticsniper (packer -> C ABI -> CUDA kernel -> writers) must write
the bytes the unmodified reference binary writes — against the committed golden files of the reference's own integration
data (BASELINE.json configs[0]) and against the reference binary run live on synthetic BAM pairs."""
import json
import os
import shutil

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INTEG = os.path.join(ROOT, "tests", "golden", "integration")
RUNS = json.load(open(os.path.join(INTEG, "ref_runs.json")))


@pytest.fixture(scope="module")
def ct(built):
    from somatic_sniper_b200 import cli_tools
    return cli_tools


@pytest.mark.parametrize("run", sorted(RUNS))
def test_integration_data_matches_reference_files(ct, run, tmp_path):
    fa = str(tmp_path / "small.fa")
    shutil.copy(os.path.join(INTEG, "small.fa"), fa)
    out = str(tmp_path / "out")
    r = ct.run_cli(ct.CLI, fa, os.path.join(INTEG, "t-small.bam"), os.path.join(INTEG, "n-small.bam"), out, RUNS[run])
    assert r["rc"] == 0, r["stderr"]
    assert ct.comparable(out) == ct.comparable(os.path.join(INTEG, "ref_%s.out" % run))
    assert "Preparing to snipe some somatics" in r["stderr"] and "Tumor bam is " in r["stderr"]


def test_expected_vcf_of_the_reference_repo(ct, tmp_path):
    """integration-test/test_sniper.py:8-19 — the reference's own end-to-end test."""
    fa = str(tmp_path / "small.fa")
    shutil.copy(os.path.join(INTEG, "small.fa"), fa)
    out = str(tmp_path / "out.vcf")
    r = ct.run_cli(ct.CLI, fa, os.path.join(INTEG, "t-small.bam"), os.path.join(INTEG, "n-small.bam"), out, ["-F", "vcf"])
    assert r["rc"] == 0, r["stderr"]
    assert ct.comparable(out) == ct.comparable(os.path.join(INTEG, "expected.vcf"))


@pytest.fixture(scope="module")
def synthetic_pair(ct, tmp_path_factory):
    d = str(tmp_path_factory.mktemp("pair"))
    return ct.make_pair(d, [("chr1", 120000), ("chr2", 30000), ("chrU", 300), ("chr3", 50000)], seed=21, n_runs=2,
                        variant_kw=dict(het_every=1000, hom_every=5000, som_every=2000, loh_every=10000, gor_every=10000))


OPTION_SETS = [[], ["-F", "vcf"], ["-F", "bed"], ["-J", "-F", "vcf"], ["-s", "0.001", "-Q", "0", "-F", "vcf"],
               ["-q", "1", "-Q", "0", "-L", "-G"], ["-p", "-Q", "0"], ["-T", "0.7", "-N", "3", "-r", "0.01", "-Q", "5"]]


@pytest.mark.parametrize("opts", OPTION_SETS, ids=lambda o: "_".join(o) or "default")
def test_synthetic_pair_matches_live_reference(ct, synthetic_pair, opts, tmp_path):
    if not ct.reference_available():
        pytest.skip("oracle/_ref is not built")
    p = synthetic_pair
    mine, ref = str(tmp_path / "mine"), str(tmp_path / "ref")
    r1 = ct.run_cli(ct.CLI, p["fasta"], p["tumor"], p["normal"], mine, opts, stats=str(tmp_path / "stats.json"))
    r2 = ct.run_cli(ct.REF_CLI, p["fasta"], p["tumor"], p["normal"], ref, opts)
    assert r1["rc"] == 0 and r2["rc"] == 0, r1["stderr"]
    assert ct.comparable(mine) == ct.comparable(ref)
    assert r1["stats"]["kernel_launches"] >= 1 and r1["stats"]["columns"] > 150000
    assert sum(1 for _ in open(ref)) > 20


@pytest.mark.parametrize("run", sorted(RUNS))
def test_device_pileup_on_integration_data(ct, run, tmp_path):
    """SSN_DEVICE_PILEUP=1: the reads travel and the GPU builds the columns (csrc/ssn_pileup.cuh); same bytes out."""
    fa = str(tmp_path / "small.fa")
    shutil.copy(os.path.join(INTEG, "small.fa"), fa)
    out = str(tmp_path / "out")
    r = ct.run_cli(ct.CLI, fa, os.path.join(INTEG, "t-small.bam"), os.path.join(INTEG, "n-small.bam"), out, RUNS[run],
                   env={"SSN_DEVICE_PILEUP": 1}, stats=str(tmp_path / "stats.json"))
    assert r["rc"] == 0, r["stderr"]
    assert ct.comparable(out) == ct.comparable(os.path.join(INTEG, "ref_%s.out" % run))
    assert r["stats"]["device_pileup"] == 1


@pytest.mark.parametrize("opts", [[], ["-J", "-F", "vcf"], ["-q", "1", "-Q", "0", "-L", "-G"]], ids=lambda o: "_".join(o) or "default")
def test_device_pileup_matches_host_packer_and_reference(ct, synthetic_pair, opts, tmp_path):
    p = synthetic_pair
    host, dev, ref = str(tmp_path / "host"), str(tmp_path / "dev"), str(tmp_path / "ref")
    r0 = ct.run_cli(ct.CLI, p["fasta"], p["tumor"], p["normal"], host, opts, stats=str(tmp_path / "s0.json"))
    r1 = ct.run_cli(ct.CLI, p["fasta"], p["tumor"], p["normal"], dev, opts, env={"SSN_DEVICE_PILEUP": 1, "SSN_DEVICES": "0,0"},
                    stats=str(tmp_path / "s1.json"))
    assert r0["rc"] == 0 and r1["rc"] == 0, r1["stderr"]
    assert ct.comparable(dev) == ct.comparable(host)
    assert r1["stats"]["device_pileup"] == 1 and r1["stats"]["columns"] == r0["stats"]["columns"]
    assert r1["stats"]["read_bases"] == r0["stats"]["read_bases"] and r1["stats"]["sites"] == r0["stats"]["sites"] > 20
    if ct.reference_available():
        ct.run_cli(ct.REF_CLI, p["fasta"], p["tumor"], p["normal"], ref, opts)
        assert ct.comparable(dev) == ct.comparable(ref)


def test_two_scoring_contexts_and_small_slots(ct, synthetic_pair, tmp_path):
    """Region batches alternate between scoring contexts (the multi-GPU path, here two contexts on one device) and the
    ordered merge still reproduces the reference's file."""
    if not ct.reference_available():
        pytest.skip("oracle/_ref is not built")
    p = synthetic_pair
    mine, ref = str(tmp_path / "mine"), str(tmp_path / "ref")
    r1 = ct.run_cli(ct.CLI, p["fasta"], p["tumor"], p["normal"], mine, ["-F", "vcf", "-Q", "0"],
                    env={"SSN_DEVICES": "0,0", "SSN_SLOTS": 1, "SSN_THREADS": 3}, stats=str(tmp_path / "stats.json"))
    r2 = ct.run_cli(ct.REF_CLI, p["fasta"], p["tumor"], p["normal"], ref, ["-F", "vcf", "-Q", "0"])
    assert r1["rc"] == 0, r1["stderr"]
    assert ct.comparable(mine) == ct.comparable(ref)
    assert r1["stats"]["gpus"] == 2


def test_stdin_tumor_and_bad_format(ct, synthetic_pair, tmp_path):
    import subprocess
    p = synthetic_pair
    out = str(tmp_path / "o")
    with open(p["tumor"], "rb") as f:
        r = subprocess.run([ct.CLI, "-f", p["fasta"], "-F", "bed", "-", p["normal"], out], stdin=f, stderr=subprocess.PIPE)
    assert r.returncode == 0
    ref = str(tmp_path / "r")
    if ct.reference_available():
        ct.run_cli(ct.REF_CLI, p["fasta"], p["tumor"], p["normal"], ref, ["-F", "bed"])
        assert ct.comparable(out) == ct.comparable(ref)
    r = subprocess.run([ct.CLI, "-f", p["fasta"], "-F", "gff", p["tumor"], p["normal"], out], stderr=subprocess.PIPE)
    assert r.returncode == 1 and b"unknown output format: 'gff'. Abort!" in r.stderr
