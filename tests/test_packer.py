"""CPU tests of the host side above the C ABI: the BAM -> SoA column packer (libssn_host.so) against the reference's own
dual-BAM walker (bam_sspileup_file, sniper_pileup.c:226-266, driven through oracle/_ref) and against committed goldens;
the line writers against the reference binary's golden output files; the command line's argument handling."""
import ctypes as C
import json
import os
import re
import struct
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
INTEG = os.path.join(GOLD, "integration")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_harness.so")


@pytest.fixture(scope="module")
def host(built):
    from somatic_sniper_b200 import hostapi
    hostapi.load_library()
    return hostapi


@pytest.fixture(scope="module")
def sim(built):
    from somatic_sniper_b200 import simbam
    return simbam


def _reference_walk(tumor, normal, mapq):
    """Columns the unmodified reference walker hands to its callback: (cols[n,6], entries[m,3])."""
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref is not built")
    lib = C.CDLL(REF_SO)
    lib.ref_pileup_dump.restype = C.c_long
    cols, ent, nent = C.POINTER(C.c_int32)(), C.POINTER(C.c_uint8)(), C.c_long()
    nc = lib.ref_pileup_dump(os.fsencode(tumor), os.fsencode(normal), int(mapq), C.byref(cols), C.byref(ent), C.byref(nent))
    assert nc >= 0
    rc = np.ctypeslib.as_array(cols, shape=(max(nc, 1), 6))[:nc].copy() if nc else np.zeros((0, 6), np.int32)
    re_ = np.ctypeslib.as_array(ent, shape=(max(nent.value, 1), 3))[:nent.value].copy() if nent.value else np.zeros((0, 3), np.uint8)
    return rc, re_


def _expected_columns(rc, re_, ref_lookup, nt16):
    """Filters the reference walker's columns to those glf_somatic can emit for (somatic_sniper.c:127,156) and returns
    a list of (tid, pos, ref_char, sorted tumor entries, sorted normal entries)."""
    out = []
    e = 0
    for tid, pos, _n1, _n2, nt, nn in rc:
        t, n = re_[e:e + nt], re_[e + nt:e + nt + nn]
        e += nt + nn
        rb = ref_lookup(tid, pos)
        if nt > 0 and nn > 0 and nt16[rb] != 15:
            out.append((int(tid), int(pos), int(rb), sorted(map(tuple, t.tolist())), sorted(map(tuple, n.tolist()))))
    return out


def _packed_columns(batches):
    out = []
    for b in batches:
        off = b["off"].astype(np.int64)
        assert off[0] == 0 and off[-1] == b["n_entries"]
        for c in range(b["ncol"]):
            t = np.stack([b[k][off[2 * c]:off[2 * c + 1]] for k in ("bs", "bq", "mq")], 1)
            n = np.stack([b[k][off[2 * c + 1]:off[2 * c + 2]] for k in ("bs", "bq", "mq")], 1)
            out.append((int(b["tid"][c]), int(b["pos"][c]), int(b["ref_char"][c]), sorted(map(tuple, t.tolist())), sorted(map(tuple, n.tolist()))))
    return out


def _check_pair(host, sim, tumor, normal, fasta, ref, mapq=0, **kw):
    from somatic_sniper_b200 import api
    nt16 = api.nt16_lut()
    seqs = [s for _, s in ref]
    rc, re_ = _reference_walk(tumor, normal, mapq)
    want = _expected_columns(rc, re_, lambda t, p: seqs[t][p] if t < len(seqs) and p < len(seqs[t]) else ord("N"), nt16)
    pk = host.Packer(tumor, normal, fasta, mapq, **kw)
    batches = list(pk)
    st = pk.stats()
    pk.close()
    got = _packed_columns(batches)
    assert len(got) == len(want)
    assert [g[:3] for g in got] == [w[:3] for w in want], "column sets differ"
    assert got == want, "column contents differ"
    for b in batches:
        assert np.array_equal(b["ref_nt16"], nt16[b["ref_char"]])
        assert b["ncol"] <= kw.get("max_columns", 1 << 16) and b["n_entries"] <= kw.get("max_entries", 1 << 23)
    assert st.columns == len(got)
    return got


@pytest.mark.parametrize("mapq,threads,caps", [(0, 3, (1 << 16, 1 << 23)), (20, 0, (5000, 200000)), (1, 2, (700, 1 << 23))])
def test_packer_matches_reference_walker(host, sim, tmp_path, mapq, threads, caps):
    """Simulated multi-contig pair with indels, clips, N skips, duplicates, QC-fail/secondary reads, mapQ 0, N bases,
    an empty contig and N runs in the reference; batch capacities small enough to split windows."""
    from somatic_sniper_b200 import cli_tools
    pair = cli_tools.make_pair(str(tmp_path), [("chr1", 30000), ("chr2", 12000), ("chrEmpty", 100), ("chr3", 20000)], 40, 20,
                               seed=5 + mapq, n_runs=2, variant_kw=dict(het_every=500, hom_every=2000, som_every=1000, loh_every=5000, gor_every=5000))
    ref = sim.make_reference([("chr1", 30000), ("chr2", 12000), ("chrEmpty", 100), ("chr3", 20000)], seed=5 + mapq, n_runs=2)
    got = _check_pair(host, sim, pair["tumor"], pair["normal"], pair["fasta"], ref, mapq, threads=threads, max_columns=caps[0], max_entries=caps[1])
    assert len(got) > 50000


def test_packer_integration_bams_match_golden(host):
    """The reference's own integration BAMs: columns equal the golden dump of the unmodified reference's walk."""
    import shutil, tempfile
    from somatic_sniper_b200 import api
    d = tempfile.mkdtemp()
    fa = os.path.join(d, "small.fa")
    shutil.copy(os.path.join(INTEG, "small.fa"), fa)
    pk = host.Packer(os.path.join(INTEG, "t-small.bam"), os.path.join(INTEG, "n-small.bam"), fa, 0, threads=1)
    assert pk.target_names() == ["17"]
    got = _packed_columns(list(pk))
    pk.close()
    z = np.load(os.path.join(GOLD, "integration.npz"))
    off = z["off"].astype(np.int64)
    nt16 = api.nt16_lut()
    want = []
    for c in range(len(z["pos"])):
        t = np.stack([z[k][off[2 * c]:off[2 * c + 1]] for k in ("bs", "bq", "mq")], 1)
        n = np.stack([z[k][off[2 * c + 1]:off[2 * c + 2]] for k in ("bs", "bq", "mq")], 1)
        if len(t) and len(n) and nt16[z["ref_char"][c]] != 15:
            want.append((0, int(z["pos"][c]), int(z["ref_char"][c]), sorted(map(tuple, t.tolist())), sorted(map(tuple, n.tolist()))))
    assert got == want and len(got) == 20


# ---- crafted inputs ----------------------------------------------------------------------------------------------

def _write_pair(sim, d, contigs, ref, reads_t, reads_n, eof=True):
    """reads: list of (tid, pos, mapq, flag, cigar string like '5M2D3M', bases, quals)"""
    def enc(reads):
        out = []
        for tid, pos, mq, flag, cig, bases, quals in reads:
            cigar = [(m[1], int(m[0])) for m in re.findall(r"(\d+)([MIDNSHP=X])", cig)]
            codes = sim._NT16[np.frombuffer(bases.encode(), np.uint8)]
            out.append(sim.bam_record(tid, pos, mq, flag, cigar, codes, np.asarray(quals, np.uint8)))
        return out
    fa = os.path.join(d, "ref.fa")
    sim.write_fasta(fa, ref)
    t, n = os.path.join(d, "t.bam"), os.path.join(d, "n.bam")
    sim.write_bam(t, contigs, enc(reads_t), eof_marker=eof)
    sim.write_bam(n, contigs, enc(reads_n), eof_marker=eof)
    return t, n, fa


def _simple_reads(tid, starts, length=20, mq=60, flag=0, base="A", q=30):
    return [(tid, s, mq, flag, "%dM" % length, base * length, [q] * length) for s in starts]


@pytest.mark.parametrize("last_start,kept", [(9, True), (10, False), (30, False)])
def test_first_read_of_later_target_is_dropped(host, sim, tmp_path, last_start, kept):
    """QUIRK B1 (sniper_pileup.c:216): the first passing read of a later target survives only if its end lies beyond
    the start of the previous target's last passing read.  Here that read covers chr2:0-9 (end 10)."""
    contigs = [("chr1", 200), ("chr2", 200)]
    ref = [("chr1", np.frombuffer(b"ACGT" * 50, np.uint8)), ("chr2", np.frombuffer(b"TGCA" * 50, np.uint8))]
    mk = lambda: _simple_reads(0, [0, 5, last_start]) + _simple_reads(1, [0], length=10) + _simple_reads(1, [0, 0, 4], length=10)
    t, n, fa = _write_pair(sim, str(tmp_path), contigs, ref, mk(), mk())
    got = _check_pair(host, sim, t, n, fa, ref)
    depth_chr2_0 = [len(g[3]) for g in got if g[0] == 1 and g[1] == 0]
    assert depth_chr2_0 == [3 if kept else 2]


def test_cigar_operators_and_filters(host, sim, tmp_path):
    """D positions give no entry, N skips give nothing, I/S consume only the query, =/X (unknown to samtools-0.1.6) advance
    nothing (QUIRK B6); unmapped/secondary/QC-fail/duplicate reads and reads below -q never enter; supplementary ones do (B11)."""
    contigs = [("c", 400)]
    ref = [("c", np.frombuffer((b"ACGTTGCA" * 50), np.uint8))]
    L = "ACGTACGTACGTACGTACGT"
    reads = [
        (0, 10, 60, 0, "5M3D15M", L, list(range(20, 40))),
        (0, 12, 50, 16, "4S6M2I8M", L, list(range(10, 30))),
        (0, 14, 40, 0, "6M50N14M", L, [33] * 20),
        (0, 15, 30, 0x800, "20M", L, [35] * 20),
        (0, 16, 60, 4, "20M", L, [35] * 20),
        (0, 16, 60, 0x100, "20M", L, [35] * 20),
        (0, 17, 60, 0x200, "20M", L, [35] * 20),
        (0, 17, 60, 0x400, "20M", L, [35] * 20),
        (0, 18, 3, 0, "20M", L, [35] * 20),
        (0, 20, 60, 0, "5M4=6X5M", L, [36] * 20),
        (0, 21, 60, 0, "2H10M3P10M2H", L, [37] * 20),
        (0, 22, 60, 0, "20S", L, [38] * 20),                   # no reference length at all
        (0, 60, 255, 16, "20M", "NNNN=RYKMACGTACGTACG", [0, 64, 93, 255] * 5),
        (0, 390, 60, 0, "20M", L, [30] * 20),                  # runs past the end of the reference sequence
    ]
    t, n, fa = _write_pair(sim, str(tmp_path), contigs, ref, reads, reads[::1])
    for mapq in (0, 4, 41):
        _check_pair(host, sim, t, n, fa, ref, mapq, threads=1)


def test_disjoint_and_empty_inputs(host, sim, tmp_path):
    contigs = [("a", 300), ("b", 300), ("c", 300)]
    ref = [(nm, np.frombuffer(b"ACGT" * 75, np.uint8)) for nm, _ in contigs]
    # tumor covers a and c, normal covers b and c: only c can produce columns
    t, n, fa = _write_pair(sim, str(tmp_path), contigs, ref, _simple_reads(0, [5, 50]) + _simple_reads(2, [100, 110]),
                           _simple_reads(1, [7]) + _simple_reads(2, [105, 250]))
    got = _check_pair(host, sim, t, n, fa, ref)
    assert {g[0] for g in got} == {2} and len(got) == 20
    d2 = tmp_path / "empty"
    d2.mkdir()
    t, n, fa = _write_pair(sim, str(d2), contigs, ref, [], _simple_reads(0, [5]))
    assert _check_pair(host, sim, t, n, fa, ref) == []
    d3 = tmp_path / "noeof"
    d3.mkdir()
    t, n, fa = _write_pair(sim, str(d3), contigs, ref, _simple_reads(0, [5, 6]), _simple_reads(0, [5]), eof=False)
    assert len(_check_pair(host, sim, t, n, fa, ref)) == 20


def test_truncated_bam_stops_like_the_reference(host, sim, tmp_path):
    """A file cut in the middle of a BGZF block: both walkers stop at the last complete block."""
    contigs = [("a", 100000)]
    ref = sim.make_reference(contigs, seed=3)
    var = sim.make_variants(ref, seed=4)
    fa = str(tmp_path / "ref.fa")
    sim.write_fasta(fa, ref)
    t, n = str(tmp_path / "t.bam"), str(tmp_path / "n.bam")
    sim.simulate_bam(t, ref, var, "tumor", 12, seed=1)
    sim.simulate_bam(n, ref, var, "normal", 8, seed=2)
    size = os.path.getsize(t)
    with open(t, "r+b") as f:
        f.truncate(size * 2 // 3)
    got = _check_pair(host, sim, t, n, fa, ref, threads=2)
    assert 1000 < len(got) < 100000


def test_reference_lookup_quirks(host, sim, tmp_path):
    """Unknown contig -> every base 'N' -> nothing emitted (B12); lower-case and IUPAC reference characters are kept raw."""
    contigs = [("known", 200), ("missing", 200)]
    seq = np.frombuffer((b"acgtRYacgtNNACGTnnAC" * 10), np.uint8)
    t, n, fa = _write_pair(sim, str(tmp_path), contigs, [("known", seq)], _simple_reads(0, [0, 30]) + _simple_reads(1, [0]) * 3,
                           _simple_reads(0, [0, 25]) + _simple_reads(1, [0]) * 3)
    got = _check_pair(host, sim, t, n, fa, [("known", seq), ("missing", np.zeros(0, np.uint8))])
    assert {g[0] for g in got} == {0}
    assert {chr(g[2]) for g in got} == set("acgtRYACGT")


def test_unsorted_input_aborts(host, sim, tmp_path):
    contigs = [("a", 300), ("b", 300)]
    ref = [(nm, np.frombuffer(b"ACGT" * 75, np.uint8)) for nm, _ in contigs]
    t, n, fa = _write_pair(sim, str(tmp_path), contigs, ref, _simple_reads(1, [5]) + _simple_reads(0, [50]), _simple_reads(0, [7]))
    code = ("import sys; sys.path.insert(0, %r); from somatic_sniper_b200 import hostapi; "
            "print(len(list(hostapi.Packer(%r, %r, %r, 0, threads=0))))" % (ROOT, t, n, fa))
    p = subprocess.run([sys.executable, "-c", code], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert p.returncode == -6 and b"the input is not sorted. Abort!" in p.stderr       # abort(), sniper_pileup.c:211-214


def test_host_header_symbols_are_exported(host):
    hdr = open(os.path.join(ROOT, "include", "ssn_host.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ssn_[a-z0-9_]+)\s*\(", hdr))
    lib = host.load_library()
    assert declared == set(host.HOST_SYMBOLS)
    assert not [s for s in declared if not hasattr(lib, s)]


# ---- writers ----------------------------------------------------------------------------------------------------------

def _opts_to_oracle(opts):
    kw, i = {}, 0
    while i < len(opts):
        o = opts[i]
        if o in ("-F", "-n", "-t"):
            i += 1
        elif o == "-J":
            kw["use_joint"] = 1
        elif o == "-s":
            kw["use_joint"] = 1; kw["somatic_rate"] = float(opts[i + 1]); i += 1
        elif o == "-Q":
            kw["min_somatic_qual"] = int(opts[i + 1]); i += 1
        elif o == "-q":
            kw["mapq"] = int(opts[i + 1]); i += 1
        elif o == "-p":
            kw["use_priors"] = 0
        elif o == "-L":
            kw["include_loh"] = 0
        elif o == "-G":
            kw["include_gor"] = 0
        elif o == "-T":
            kw["theta"] = float(opts[i + 1]); i += 1
        elif o == "-N":
            kw["n_hap"] = int(opts[i + 1]); i += 1
        elif o == "-r":
            kw["het_rate"] = float(opts[i + 1]); i += 1
        i += 1
    return kw


def _records_to_sites(api, recs, cols):
    s = np.zeros(len(recs), api.SITE_DTYPE)
    s["column"] = cols
    s["ref_base4"] = recs[:, 1]
    for base, key in ((2, "tumor"), (28, "normal")):
        for i, f in enumerate(["genotype", "joint_genotype", "joint_consensus_quality", "consensus_quality", "variant_allele_quality",
                               "somatic_score", "variant_status"]):
            s[key][f] = recs[:, base + i]
        s[key]["mean_baseQ"] = recs[:, base + 7:base + 11]
        s[key]["mean_mapQ"] = recs[:, base + 11:base + 15]
        s[key]["base_occ"] = recs[:, base + 15:base + 19]
        s[key]["dp4"] = recs[:, base + 19:base + 23]
        s[key]["total_depth"] = recs[:, base + 23]
        s[key]["total_mean_mapQ"] = recs[:, base + 24]
    return s


RUNS = json.load(open(os.path.join(INTEG, "ref_runs.json")))


@pytest.mark.parametrize("run", sorted(RUNS))
def test_writers_reproduce_reference_output_files(host, run, tmp_path):
    """packer (integration BAMs) -> oracle records (the checker standing in for the GPU) -> ssn_writer_* must give the
    bytes the unmodified reference binary wrote for the same options (tests/golden/integration/ref_*.out)."""
    import shutil
    from oracle.loader import Oracle
    from somatic_sniper_b200 import api, cli_tools
    opts = RUNS[run]
    kw = _opts_to_oracle(opts)
    mapq = kw.pop("mapq", 0)
    fa = str(tmp_path / "small.fa")
    shutil.copy(os.path.join(INTEG, "small.fa"), fa)
    pk = host.Packer(os.path.join(INTEG, "t-small.bam"), os.path.join(INTEG, "n-small.bam"), fa, mapq, threads=0)
    names = pk.target_names()
    batches = list(pk)
    pk.close()
    fmt = opts[opts.index("-F") + 1] if "-F" in opts else "classic"
    nid = opts[opts.index("-n") + 1] if "-n" in opts else "NORMAL"
    tid_ = opts[opts.index("-t") + 1] if "-t" in opts else "TUMOR"
    lib = host.load_library()
    libc = C.CDLL(None)
    libc.fopen.restype = C.c_void_p
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.fclose.argtypes = [C.c_void_p]
    lib.ssn_writer_create.restype = C.c_void_p
    lib.ssn_writer_create.argtypes = [C.c_char_p, C.c_void_p]
    lib.ssn_writer_header.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p]
    lib.ssn_writer_site.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.c_int, C.c_void_p]
    lib.ssn_writer_destroy.argtypes = [C.c_void_p]
    out = str(tmp_path / "out.txt")
    fh = libc.fopen(out.encode(), b"w")
    w = lib.ssn_writer_create(fmt.encode(), fh)
    assert w
    lib.ssn_writer_header(w, fa.encode(), nid.encode(), tid_.encode())
    orc = Oracle(**kw)
    for b in batches:
        r = orc.score_batch(b["off"], b["bs"], b["bq"], b["mq"], b["ref_char"])
        sites = _records_to_sites(api, r["recs"][:r["n_emit"]], r["emit_col"][:r["n_emit"]])
        for s in sites:
            c = int(s["column"])
            lib.ssn_writer_site(w, names[b["tid"][c]].encode(), int(b["pos"][c]), int(b["ref_char"][c]), s.tobytes())
    lib.ssn_writer_destroy(w)
    libc.fclose(fh)
    want = os.path.join(INTEG, "ref_%s.out" % run)
    assert cli_tools.comparable(out) == cli_tools.comparable(want)
    assert lib.ssn_writer_create(b"nonsense", None) is None


# ---- command line (argument handling happens before any GPU work) ----------------------------------------------------

def _cli(*args):
    from somatic_sniper_b200 import cli_tools
    return subprocess.run([cli_tools.CLI] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE)


def test_cli_argument_errors(built):
    p = _cli()
    assert p.returncode == 1 and b"[options] -f <ref.fasta> <tumor.bam> <normal.bam> <snp_output_file>" in p.stderr
    assert b"-Q INT    filtering somatic snv output with somatic quality less than  INT [15]" in p.stderr
    p = _cli("-Q", "40", "-s", "0.5")
    assert p.returncode == 1 and b"INT [40]" in p.stderr and b"(implies -J) [0.500000]" in p.stderr
    p = _cli("-z")
    assert p.returncode == 1 and b"Unrecognizd option '-?'." in p.stderr
    p = _cli("-I", "x", "a", "b", "c")
    assert p.returncode == 1 and b"Unrecognizd option '-I'." in p.stderr        # accepted by getopt, no case (main.c:80,97)
    p = _cli("a.bam", "b.bam", "out")
    assert p.returncode == 1 and b"You MUST specify a reference sequence. It isn't optional." in p.stderr
    p = _cli("-v")
    assert p.returncode == 0 and p.stdout.startswith(b"Somatic Sniper version (")


def test_cli_has_no_cpu_fallback(built, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import shutil
    fa = str(tmp_path / "small.fa")
    shutil.copy(os.path.join(INTEG, "small.fa"), fa)
    p = _cli("-f", fa, os.path.join(INTEG, "t-small.bam"), os.path.join(INTEG, "n-small.bam"), str(tmp_path / "o"))
    assert p.returncode == 1 and b"there is no CPU path" in p.stderr


@pytest.mark.parametrize("mapq", [0, 20])
def test_read_window_route_equals_the_column_route(host, sim, tmp_path, mapq):
    """ssn_packer_next_reads (what the CLI ships under SSN_DEVICE_PILEUP=1) + the packer's count/fill over those whole
    windows must give the very columns ssn_packer_next gives: same positions, reference characters and entries in the
    same order — on the simulated multi-contig pair (indels, clips, N skips, filtered reads, quirk B1)."""
    from somatic_sniper_b200 import cli_tools
    contigs = [("chr1", 30000), ("chr2", 12000), ("chrEmpty", 100), ("chr3", 20000)]
    pair = cli_tools.make_pair(str(tmp_path), contigs, 30, 20, seed=11 + mapq, n_runs=2,
                               variant_kw=dict(het_every=500, hom_every=2000, som_every=1000, loh_every=5000, gor_every=5000))
    pk = host.Packer(pair["tumor"], pair["normal"], pair["fasta"], mapq, threads=2, max_columns=1 << 16, max_entries=1 << 23)
    want = list(pk)
    pk.close()
    key = lambda b: (b["tid"].astype(np.int64) << 32) | b["pos"]
    w_key = np.concatenate([key(b) for b in want])
    w_refc = np.concatenate([b["ref_char"] for b in want])
    w_cnt = np.concatenate([np.diff(b["off"].astype(np.int64)) for b in want])
    w_ent = np.concatenate([np.stack([b["bs"], b["bq"], b["mq"]], 1) for b in want])
    pk = host.Packer(pair["tumor"], pair["normal"], pair["fasta"], mapq, threads=2, max_columns=1 << 16, max_entries=1 << 23)
    g_key, g_refc, g_cnt, g_ent, n_batches = [], [], [], [], 0
    for rb in pk.read_batches(max_windows=3):
        cols = host.pack_read_windows(rb, max_entries=1 << 23)
        w, o = cols["pos"] >> 13, cols["pos"] & 8191
        g_key.append((rb["win"]["tid"][w].astype(np.int64) << 32) | (rb["win"]["lo"][w].astype(np.int64) + o))
        g_refc.append(rb["ref_char"][w, o])
        g_cnt.append(np.diff(cols["off"].astype(np.int64)))
        g_ent.append(np.stack([cols["bs"], cols["bq"], cols["mq"]], 1))
        n_batches += 1
    pk.close()
    assert n_batches >= 3 and len(w_key) > 40000
    assert np.array_equal(np.concatenate(g_key), w_key)
    assert np.array_equal(np.concatenate(g_refc), w_refc)
    assert np.array_equal(np.concatenate(g_cnt), w_cnt)
    assert np.array_equal(np.concatenate(g_ent), w_ent)
