"""CPU tests (no GPU): pin the oracle port (oracle/ssn_oracle.c) to the reference.

Three anchors, in the order the task asks for them:
  (a) the reference's own golden: integration-test/data/expected.vcf:19-21 (3 sites, every FORMAT field),
  (b) the known-answer columns of SURVEY.md App. C.3 (dumped from the reference functions),
  (c) outputs of the UNMODIFIED reference on seeded batches, committed as tests/golden/*.npz by
      tests/golden/make_golden.py — and, when oracle/_ref/ is present, the live reference itself.
"""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import CONFIGS, oracle_kwargs

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def mods(built):
    from oracle import loader
    from somatic_sniper_b200 import synth
    return loader, synth


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _gold_cols():
    c = dict(np.load(os.path.join(GOLD, "columns_in.npz")))
    c["ncol"] = len(c["ref_char"])
    c["n_entries"] = len(c["bs"])
    return c


TABLE_SETS = json.load(open(os.path.join(GOLD, "tables.json")))


@pytest.mark.parametrize("name", sorted(TABLE_SETS))
def test_oracle_tables_match_reference_golden(mods, name):
    loader, _ = mods
    meta = TABLE_SETS[name]
    t = loader.Oracle(**meta["params"]).tables()
    assert _sha(t["coef"]) == meta["coef_sha256"]
    assert _sha(t["lhet"]) == meta["lhet_sha256"]
    assert _sha(t["fk"]) == meta["fk_sha256"]
    assert float(t["q_r"]) == meta["q_r"]
    assert int(np.count_nonzero(t["coef"])) == meta["coef_nonzero"]
    small = np.load(os.path.join(GOLD, "tables_small.npz"))
    for k in ("prior", "jointprior", "qadd", "fk"):
        assert np.array_equal(t[k], small["%s__%s" % (name, k)]), k


def test_table_known_answers(mods):
    """SURVEY.md App. C.3 / A.8 values."""
    loader, _ = mods
    t = loader.Oracle().tables()
    assert int(float(t["q_r"]) + 0.5) == 27 and abs(float(t["q_r"]) - 26.9856968) < 1e-6
    assert int(np.count_nonzero(t["coef"])) == 2072385
    assert t["coef"][20 << 16 | 30 << 8 | 3] == -33.353304137588516
    assert t["lhet"][10 << 8 | 10] == -1.7361522965964564
    assert t["lhet"][30 << 8] == -20.794415416798358 and t["lhet"][0] == 0
    assert t["fk"][1] == 0.85450002302601935 and t["fk"][255] == 0.029999999329447746
    assert list(t["prior"][1]) == [0, 30, 30, 30, 33, 60, 60, 33, 60, 33]
    assert list(t["qadd"][500:525]) == [-12, -11, -10, -10, -9, -8, -7, -6, -5, -5, -4, -4, -3, -3, -2, -2, -1, -1,
                                        -1, -1, -1, -1, 0, 0, 0]
    assert t["qadd"][0] == -512 and t["qadd"][999] == 0 and not t["qadd"][1000:].any()
    tj = loader.Oracle(use_joint=1, somatic_rate=0.01).tables()
    assert list(tj["jointprior"][1][0]) == [0, 20, 20, 20, 40, 40, 40, 40, 40, 40]
    assert list(tj["jointprior"][1][5]) == [100, 80, 80, 100, 80, 60, 80, 100, 80, 100]
    assert list(tj["jointprior"][1][9]) == [73, 73, 73, 53, 73, 73, 53, 73, 53, 33]


def test_known_answer_columns(mods):
    loader, synth = mods
    cols = synth.kat_columns()
    r = loader.Oracle().score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"], want_glf=True)
    g = r["glf"]
    assert list(g[0, :14]) == [30, 0, 54, 54, 161, 167, 167, 230, 230, 230, 14, 60, 10, 0x313c0333]
    assert list(g[1, :14]) == [238, 255, 0, 255, 255, 255, 255, 255, 255, 255, 28, 60, 58, 0x513cd311]
    assert list(g[1, 14:]) == [0, 90, 90, 90, 255, 255, 255, 255, 255, 255, 0, 60, 30, 0x133c7500]
    assert list(g[2, :14]) == [0, 155, 255, 255, 255, 255, 255, 255, 255, 255, 255, 52, 1458, 0x1334b649]
    assert list(g[2, 14:]) == [0, 255, 255, 255, 255, 255, 255, 255, 255, 255, 0, 60, 1502, 0x123cff00]
    assert list(g[3, :14]) == [2, 8, 8, 0, 40, 40, 29, 40, 29, 13, 8, 32, 7, 0x18200b0e]
    assert list(g[3, 14:]) == [0, 6, 6, 1, 74, 74, 69, 74, 69, 66, 4, 60, 3, 0x193c1c05]
    assert list(r["emit_col"]) == [1] and r["score"][1] == 120 and r["score"][2] == 255
    rec = r["recs"][0]
    #        Tgt Tjgt Tjcq Tcq  Tvaq SSC  SS
    assert list(rec[2:9]) == [5, 0, 255, 211, 211, 120, 2]
    assert list(rec[28:35])[0] == 1 and rec[28 + 3] == 117 and rec[28 + 4] == 0
    rj = loader.Oracle(use_joint=1, somatic_rate=0.01).score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"],
                                                                   cols["ref_char"])
    recj = rj["recs"][0]
    assert (recj[7], recj[3], recj[28 + 1], recj[4]) == (100, 5, 1, 218)


@pytest.mark.parametrize("cfg", list(CONFIGS))
def test_oracle_matches_reference_golden_columns(mods, cfg):
    """The committed outputs of the unmodified reference on the seeded golden batch."""
    loader, _ = mods
    cols = _gold_cols()
    want = np.load(os.path.join(GOLD, "columns_%s.npz" % cfg))
    got = loader.Oracle(**oracle_kwargs(cfg)).score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"],
                                                          cols["ref_char"], want_glf=True)
    assert np.array_equal(got["glf"], want["glf"])
    assert np.array_equal(got["score"], want["score"])
    assert np.array_equal(got["emit_col"], want["emit_col"])
    assert np.array_equal(got["recs"][:, :53], want["recs"][:, :53])
    assert len(want["emit_col"]) > 0 or cfg == "joint"


# ---- (a) the reference's own expected.vcf -------------------------------------------------------------

def _parse_vcf_sample(s):
    f = s.split(":")
    ints = lambda x: [int(v) for v in x.split(",")]
    return dict(GT=f[0], IGT=f[1], DP=int(f[2]), DP4=ints(f[3]), BCOUNT=ints(f[4]), GQ=int(f[5]), JGQ=f[6], VAQ=int(f[7]),
                BQ=ints(f[8]), MQ=int(f[9]), AMQ=ints(f[10]), SS=int(f[11]), SSC=f[12])


def _gt_string(ref4, g):
    """output_vcf_gt, output_vcf.c:46-79, for the genotypes the three expected sites hold."""
    alts = [b for b in (1, 2, 4, 8) if (g & b) and b != ref4]
    alleles = ([0] if g & ref4 else []) + [1 + i for i in range(len(alts))]
    if len(alleles) == 1:
        alleles = alleles * 2
    return "/".join(str(a) for a in alleles)


@pytest.mark.parametrize("impl", ["oracle", "golden_reference"])
def test_expected_vcf_sites(mods, impl):
    """integration-test/data/expected.vcf:19-21 — the only hot-path results the reference's own tests pin."""
    loader, _ = mods
    cols = dict(np.load(os.path.join(GOLD, "integration.npz")))
    pos = cols.pop("pos")
    if impl == "oracle":
        r = loader.Oracle().score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"])
    else:
        r = np.load(os.path.join(GOLD, "integration_default.npz"))
    lines = [l.rstrip("\n").split("\t") for l in open(os.path.join(GOLD, "integration", "expected.vcf")) if not l.startswith("#")]
    assert [int(l[1]) for l in lines] == [int(pos[c]) + 1 for c in r["emit_col"]] == [63, 73, 260]
    for l, rec in zip(lines, r["recs"]):
        assert chr(rec[0]) == l[3]
        ref4 = int(rec[1])
        for s, base in ((_parse_vcf_sample(l[9]), 28), (_parse_vcf_sample(l[10]), 2)):
            gt, cq, vaq, ssc, ss = (int(rec[base + i]) for i in (0, 3, 4, 5, 6))
            assert _gt_string(ref4, gt) == s["GT"] == s["IGT"]
            assert int(rec[base + 23]) == s["DP"]
            assert [int(x) for x in rec[base + 19:base + 23]] == s["DP4"]
            assert [int(x) for x in rec[base + 15:base + 19]] == s["BCOUNT"]
            assert cq == s["GQ"] and vaq == s["VAQ"] and ss == s["SS"]
            assert int(rec[base + 24]) == s["MQ"]
            sel = [i for i in range(4) if gt & (1 << i)]
            assert [int(rec[base + 7 + i]) for i in sel] == s["BQ"]
            assert [int(rec[base + 11 + i]) for i in sel] == s["AMQ"]
            assert (s["SSC"] == ".") == (ssc == -1) and (ssc == -1 or ssc == int(s["SSC"]))


@pytest.mark.parametrize("cfg", ["default", "joint_s", "nopriors_q0"])
def test_integration_columns_match_reference(mods, cfg):
    loader, _ = mods
    cols = dict(np.load(os.path.join(GOLD, "integration.npz")))
    want = np.load(os.path.join(GOLD, "integration_%s.npz" % cfg))
    got = loader.Oracle(**oracle_kwargs(cfg)).score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"],
                                                          want_glf=True)
    for k in ("glf", "score", "emit_col"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(got["recs"][:, :53], want["recs"][:, :53])


# ---- (c) live reference, when oracle/_ref was built (this container; the prebuilt .so also travels) -------

def _ref_or_skip(loader):
    if not loader.Reference.available():
        pytest.skip("oracle/_ref/libref_harness.so not built")


@pytest.mark.parametrize("cfg", ["default", "joint_s"])
def test_oracle_matches_live_reference(mods, cfg):
    loader, synth = mods
    _ref_or_skip(loader)
    import multiprocessing as mp
    cols = synth.make_columns(3000, seed=77, edge=0.7, binned=(cfg == "default"), p_som=1e-2, p_het=1e-2, p_loh=5e-3, p_gor=5e-3)
    got = loader.Oracle(**oracle_kwargs(cfg)).score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"],
                                                          want_glf=True)
    with mp.get_context("fork").Pool(1) as pool:      # the reference keeps global state: own process
        want = pool.apply(_live_reference, (oracle_kwargs(cfg), cols))
    for k in ("glf", "score", "emit_col"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(got["recs"][:, :53], want["recs"][:, :53])


def _live_reference(kw, cols):
    from oracle import loader
    return loader.Reference(**kw).score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"], want_glf=True)


def test_allele_predicates_truth_table(mods):
    """The reference's unit truth tables (test/lib/sniper/TestAlleleUtil.cpp) via the live reference predicates
    against the oracle's gate, exhaustively over ref x tumor x normal genotype masks."""
    loader, _ = mods
    _ref_or_skip(loader)
    import ctypes as C
    L = C.CDLL(loader.REF_SO)
    pop = [bin(i).count("1") for i in range(16)]
    assert [L.ref_count_alleles(i) for i in range(16)] == pop
    for r in (1, 2, 4, 8):
        for t in range(1, 16):
            for n in range(1, 16):
                loh = int(n != t and (t & n) == t)
                gor = int((not (n != r and (r & n) == r)) and (t & ~n) == r)
                assert L.ref_should_filter_as_loh(r, t, n) == loh, (r, t, n)
                assert L.ref_should_filter_as_gor(r, t, n) == gor, (r, t, n)
    # spot values of TestAlleleUtil.cpp:119-196
    assert L.ref_should_filter_as_loh(1, 1, 3) == 1 and L.ref_should_filter_as_loh(1, 3, 1) == 0
    assert L.ref_should_filter_as_gor(1, 3, 2) == 1 and L.ref_should_filter_as_gor(1, 1, 3) == 0
