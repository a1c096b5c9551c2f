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
    r0 = ct.run_cli(ct.CLI, p["fasta"], p["tumor"], p["normal"], host, opts, env={"SSN_DEVICE_PILEUP": 0}, stats=str(tmp_path / "s0.json"))
    r1 = ct.run_cli(ct.CLI, p["fasta"], p["tumor"], p["normal"], dev, opts, env={"SSN_DEVICE_PILEUP": 1, "SSN_DEVICES": "0,0"},
                    stats=str(tmp_path / "s1.json"))
    assert r0["rc"] == 0 and r1["rc"] == 0, r1["stderr"]
    assert ct.comparable(dev) == ct.comparable(host)
    assert r0["stats"]["device_pileup"] == 0 and r1["stats"]["device_pileup"] == 1 and r1["stats"]["columns"] == r0["stats"]["columns"]
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
                    env={"SSN_DEVICES": "0,0", "SSN_SLOTS": 1, "SSN_THREADS": 3, "SSN_DEVICE_PILEUP": 0, "SSN_SHARD_WINDOWS": 3}, stats=str(tmp_path / "stats.json"))
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


# ---- region shards through the whole program ---------------------------------------------------

@pytest.mark.parametrize("env,opts", [
    ({"SSN_SHARD_WINDOWS": 2, "SSN_SHARD_LOOKBACK": 1, "SSN_GPU_INFLATE": 1}, []),
    ({"SSN_SHARD_WINDOWS": 2, "SSN_SHARD_LOOKBACK": 1, "SSN_DEVICES": "0,0", "SSN_GPU_INFLATE": 0}, ["-F", "vcf", "-Q", "0"]),
    ({"SSN_SHARD_WINDOWS": 1, "SSN_SHARD_LOOKBACK": 1, "SSN_DEVICES": "0,0,0", "SSN_THREADS": 3, "SSN_SLOTS": 2, "SSN_GPU_INFLATE": 1,
      "SSN_INFLATE_BLOCKS": 8, "SSN_INFLATE_LANES": 2}, ["-J", "-F", "vcf"]),
    ({"SSN_SHARD_WINDOWS": 3, "SSN_SHARD_LOOKBACK": 2, "SSN_DEVICE_PILEUP": 0, "SSN_DEVICES": "0,0", "SSN_GPU_INFLATE": 0}, ["-q", "1", "-Q", "0", "-L", "-G"]),
    ({"SSN_SHARD_WINDOWS": 2, "SSN_SHARD_LOOKBACK": 1, "SSN_GPU_INFLATE": 1, "SSN_INFLATE_DIRECT": 1, "SSN_INFLATE_WAVES": 2}, ["-F", "bed"]),
], ids=["1ctx_gpu_inflate", "2ctx_vcf_host_inflate", "3ctx_joint_gpu_inflate_small_lanes", "2ctx_host_columns", "1ctx_gpu_inflate_direct"])
def test_region_shards_match_the_single_process_reference(ct, synthetic_pair, env, opts, tmp_path):
    """The 4-contig pair cut into >= 8 intervals, scored on 1-3 contexts (one feeder thread each, shards taken in genome
    order, output merged in order): byte-identical to the unmodified reference binary's single front-to-back walk."""
    if not ct.reference_available():
        pytest.skip("oracle/_ref is not built")
    p = synthetic_pair
    mine, ref = str(tmp_path / "mine"), str(tmp_path / "ref")
    r1 = ct.run_cli(ct.CLI, p["fasta"], p["tumor"], p["normal"], mine, opts, env=env, stats=str(tmp_path / "stats.json"))
    r2 = ct.run_cli(ct.REF_CLI, p["fasta"], p["tumor"], p["normal"], ref, opts)
    assert r1["rc"] == 0 and r2["rc"] == 0, r1["stderr"]
    assert ct.comparable(mine) == ct.comparable(ref)
    st = r1["stats"]
    assert st["shards"] >= 8 and st["passes"] == 1 and st["gpus"] == len(str(env.get("SSN_DEVICES", "0")).split(","))
    assert st["gpu_inflate"] == env["SSN_GPU_INFLATE"] and (st["gpu_inflate_blocks"] > 100) == bool(env["SSN_GPU_INFLATE"])
    assert sum(1 for _ in open(ref)) > 20


def test_keep_rule_on_a_shard_edge_through_the_cli(ct, tmp_path):
    """QUIRK B1 (sniper_pileup.c:216) with the target change exactly on an interval edge: the reference binary drops
    chr2's first read (its end 10 is not beyond chr1's last passing read start 10), and so must the sharded run."""
    import numpy as np
    from somatic_sniper_b200 import simbam as sim
    from test_packer import _simple_reads, _write_pair
    contigs = [("chr1", 8192), ("chr2", 20000)]
    ref = [("chr1", np.frombuffer(b"ACGT" * 2048, np.uint8)), ("chr2", np.frombuffer(b"TGCA" * 5000, np.uint8))]
    outs = {}
    for last_start in (9, 10):
        d = tmp_path / str(last_start)
        d.mkdir()
        def mk(alt):
            return (_simple_reads(0, [0, 5, last_start]) + _simple_reads(0, [40], flag=0x400)
                    + _simple_reads(1, [0], length=10, base=alt) + _simple_reads(1, [0, 0, 4], length=10, base=alt)
                    + _simple_reads(1, [8190, 8200, 16380], length=30))
        t, n, fa = _write_pair(sim, str(d), contigs, ref, mk("C"), mk("T"))
        mine, want = str(d / "mine"), str(d / "ref")
        r1 = ct.run_cli(ct.CLI, fa, t, n, mine, ["-Q", "0"], env={"SSN_SHARD_WINDOWS": 1, "SSN_SHARD_LOOKBACK": 1}, stats=str(d / "s.json"))
        assert r1["rc"] == 0, r1["stderr"]
        assert r1["stats"]["shards"] == 4
        if ct.reference_available():
            ct.run_cli(ct.REF_CLI, fa, t, n, want, ["-Q", "0"])
            assert ct.comparable(mine) == ct.comparable(want)
        outs[last_start] = open(mine).read()
    # the dropped read changes the tumor depth at chr2:1 (column 8 of the classic format is not depth; compare whole lines)
    assert outs[9] != outs[10] and "chr2\t1\t" in outs[9]


def test_long_read_across_an_interval_edge_falls_back_to_one_walk(ct, tmp_path):
    """A read with a 20 kb N skip begins further in front of an interval edge than the look-back and reaches behind it: the
    sharded pass notices, the program walks the pair again as one interval, and the output equals the reference's."""
    import numpy as np
    from somatic_sniper_b200 import simbam as sim
    from test_packer import _simple_reads, _write_pair
    contigs = [("chr1", 60000)]
    ref = [("chr1", np.frombuffer(b"ACGT" * 15000, np.uint8))]
    def mk(alt):
        long_read = [(0, 100, 60, 0, "10M20000N10M", alt * 20, [30] * 20)]
        return (_simple_reads(0, [50, 90]) + long_read + _simple_reads(0, [200, 8300], length=30)
                + _simple_reads(0, [20100, 20105, 20108], length=30, base=alt) + _simple_reads(0, [30000], length=30))
    t, n, fa = _write_pair(sim, str(tmp_path), contigs, ref, mk("G"), mk("A"))
    mine, want = str(tmp_path / "mine"), str(tmp_path / "ref")
    r1 = ct.run_cli(ct.CLI, fa, t, n, mine, ["-Q", "0", "-F", "vcf"], env={"SSN_SHARD_WINDOWS": 1, "SSN_SHARD_LOOKBACK": 1}, stats=str(tmp_path / "s.json"))
    assert r1["rc"] == 0, r1["stderr"]
    assert r1["stats"]["passes"] == 2 and "one interval" in r1["stderr"]
    if ct.reference_available():
        ct.run_cli(ct.REF_CLI, fa, t, n, want, ["-Q", "0", "-F", "vcf"])
        assert ct.comparable(mine) == ct.comparable(want)
        assert sum(1 for l in open(want) if not l.startswith("#")) >= 1


def test_window_too_deep_for_a_read_batch_takes_the_column_route(ct, tmp_path):
    """Windows whose reads exceed a read batch (~4600x over 8192 positions; a slot holds 32 Mi read-bases) are packed into
    columns by the host, and the reads route resumes behind them (the contig's last, partial window fits)."""
    if not ct.reference_available():
        pytest.skip("oracle/_ref is not built")
    pair = ct.make_pair_fast(str(tmp_path), [("chr1", 20000)], 3000, 1600, seed=5)
    mine, ref = str(tmp_path / "mine"), str(tmp_path / "ref")
    r1 = ct.run_cli(ct.CLI, pair["fasta"], pair["tumor"], pair["normal"], mine, ["-Q", "30"], stats=str(tmp_path / "s.json"))
    r2 = ct.run_cli(ct.REF_CLI, pair["fasta"], pair["tumor"], pair["normal"], ref, ["-Q", "30"])
    assert r1["rc"] == 0 and r2["rc"] == 0, r1["stderr"]
    assert ct.comparable(mine) == ct.comparable(ref)
    st = r1["stats"]
    assert st["column_route_batches"] >= 2 and st["batches"] > st["column_route_batches"] and st["device_pileup"] == 1


# ---- BASELINE.json configs 2-4 against the live reference binary ---------------------------------------------------

def test_config2_one_megabase_classic(ct, tmp_path):
    """configs[1]: synthetic 1 Mb region, normal 30x / tumor 60x, -q 1 -Q 15, classic output."""
    if not ct.reference_available():
        pytest.skip("oracle/_ref is not built")
    pair = ct.make_pair_fast(str(tmp_path), [("chr1", 1000000)], 60, 30, seed=1234, level=6)
    mine, ref = str(tmp_path / "mine"), str(tmp_path / "ref")
    opts = ["-q", "1", "-Q", "15", "-F", "classic"]
    r1 = ct.run_cli(ct.CLI, pair["fasta"], pair["tumor"], pair["normal"], mine, opts, stats=str(tmp_path / "s.json"))
    r2 = ct.run_cli(ct.REF_CLI, pair["fasta"], pair["tumor"], pair["normal"], ref, opts)
    assert r1["rc"] == 0 and r2["rc"] == 0, r1["stderr"]
    assert ct.comparable(mine) == ct.comparable(ref)
    assert r1["stats"]["columns"] > 990000 and sum(1 for _ in open(ref)) > 100


def test_config3_joint_vcf_on_several_megabases(ct, tmp_path):
    """configs[2] at 6 Mb (the 50 Mb run lives in profiles/): -J -s 0.01 -T 0.85 -F vcf, two contigs, region shards on."""
    if not ct.reference_available():
        pytest.skip("oracle/_ref is not built")
    pair = ct.make_pair_fast(str(tmp_path), [("chr1", 4000000), ("chr2", 2000000)], 60, 30, seed=77)
    mine, ref = str(tmp_path / "mine"), str(tmp_path / "ref")
    opts = ["-J", "-s", "0.01", "-T", "0.85", "-F", "vcf"]
    r1 = ct.run_cli(ct.CLI, pair["fasta"], pair["tumor"], pair["normal"], mine, opts, env={"SSN_DEVICES": "0,0"}, stats=str(tmp_path / "s.json"))
    r2 = ct.run_cli(ct.REF_CLI, pair["fasta"], pair["tumor"], pair["normal"], ref, opts)
    assert r1["rc"] == 0 and r2["rc"] == 0, r1["stderr"]
    assert ct.comparable(mine) == ct.comparable(ref)
    assert r1["stats"]["shards"] >= 4 and r1["stats"]["columns"] > 5900000
    assert sum(1 for l in open(ref) if not l.startswith("#")) > 300


def test_config4_targeted_panel_depth(ct, tmp_path):
    """configs[3]: targeted panel at 2000x / 1000x (per-column sort and dependency model under depth): 8 targets x 5 kb."""
    if not ct.reference_available():
        pytest.skip("oracle/_ref is not built")
    pair = ct.make_pair_fast(str(tmp_path), [("t%02d" % i, 5000) for i in range(8)], 2000, 1000, seed=9)
    mine, ref = str(tmp_path / "mine"), str(tmp_path / "ref")
    for opts in ([], ["-J", "-F", "vcf", "-Q", "40"]):
        r1 = ct.run_cli(ct.CLI, pair["fasta"], pair["tumor"], pair["normal"], mine, opts, stats=str(tmp_path / "s.json"))
        r2 = ct.run_cli(ct.REF_CLI, pair["fasta"], pair["tumor"], pair["normal"], ref, opts)
        assert r1["rc"] == 0 and r2["rc"] == 0, r1["stderr"]
        assert ct.comparable(mine) == ct.comparable(ref)
    assert r1["stats"]["read_bases"] > 8 * 4000 * 2900


def test_region_shards_on_every_physical_gpu(ct, tmp_path):
    """With more than one GPU in the box: one feeder thread and one scoring context per physical device, shards taken in
    genome order, output merged in order — byte-identical to the run on one GPU and to the reference binary."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU in this box (the multi-context path runs as SSN_DEVICES=0,0 above)")
    pair = ct.make_pair_fast(str(tmp_path), [("chr1", 1500000), ("chr2", 700000), ("chr3", 300000)], 60, 30, seed=21)
    outs = {}
    for g in sorted({1, 2, n}):
        o = str(tmp_path / ("out%d" % g))
        r = ct.run_cli(ct.CLI, pair["fasta"], pair["tumor"], pair["normal"], o, ["-F", "vcf"], env={"SSN_GPUS": g, "SSN_SHARD_WINDOWS": 8},
                       stats=str(tmp_path / "s.json"))
        assert r["rc"] == 0, r["stderr"]
        assert r["stats"]["gpus"] == g and r["stats"]["shards"] >= 30
        outs[g] = ct.comparable(o)
    assert all(v == outs[1] for v in outs.values())
    if ct.reference_available():
        ref = str(tmp_path / "ref")
        ct.run_cli(ct.REF_CLI, pair["fasta"], pair["tumor"], pair["normal"], ref, ["-F", "vcf"])
        assert ct.comparable(ref) == outs[1]


# ---- INTEGRATION.md's binding, compiled against the unmodified reference (tools/shim/ssn_shim.c) -------------------

SHIM = os.path.join(ROOT, "oracle", "_ref", "bam-somaticsniper-shim")


def test_reference_walker_through_the_shim_reproduces_expected_vcf(ct, tmp_path):
    """Reference main() + reference walker (bam_sspileup_file) -> shim callback -> libssn.so -> reference formatter:
    the reference's own integration test (integration-test/test_sniper.py:8-19) passes with the GPU behind main.c:146."""
    if not os.path.exists(SHIM):
        pytest.skip("oracle/_ref/bam-somaticsniper-shim is not built")
    fa = str(tmp_path / "small.fa")
    shutil.copy(os.path.join(INTEG, "small.fa"), fa)
    out = str(tmp_path / "out.vcf")
    r = ct.run_cli(SHIM, fa, os.path.join(INTEG, "t-small.bam"), os.path.join(INTEG, "n-small.bam"), out, ["-F", "vcf"])
    assert r["rc"] == 0, r["stderr"]
    assert ct.comparable(out) == ct.comparable(os.path.join(INTEG, "expected.vcf"))
    for run in sorted(RUNS):
        r = ct.run_cli(SHIM, fa, os.path.join(INTEG, "t-small.bam"), os.path.join(INTEG, "n-small.bam"), out, RUNS[run])
        assert r["rc"] == 0, r["stderr"]
        assert ct.comparable(out) == ct.comparable(os.path.join(INTEG, "ref_%s.out" % run)), run


@pytest.mark.parametrize("opts", [[], ["-J", "-F", "vcf"], ["-p", "-Q", "0", "-F", "bed"], ["-q", "1", "-Q", "0", "-L", "-G", "-T", "0.7"]],
                         ids=lambda o: "_".join(o) or "default")
def test_shim_matches_reference_on_the_synthetic_pair(ct, synthetic_pair, opts, tmp_path):
    if not os.path.exists(SHIM) or not ct.reference_available():
        pytest.skip("oracle/_ref is not built")
    p = synthetic_pair
    mine, ref = str(tmp_path / "shim"), str(tmp_path / "ref")
    r1 = ct.run_cli(SHIM, p["fasta"], p["tumor"], p["normal"], mine, opts)
    r2 = ct.run_cli(ct.REF_CLI, p["fasta"], p["tumor"], p["normal"], ref, opts)
    assert r1["rc"] == 0 and r2["rc"] == 0, r1["stderr"]
    assert ct.comparable(mine) == ct.comparable(ref)
    assert sum(1 for _ in open(ref)) > 20
