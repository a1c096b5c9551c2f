"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Bar: bit-exact. Every quantity compared is an integer (quantised GLFs, consensus words, Phred scores,
site statistics); the float intermediates only exist in registers and are pinned through them.
"""
import numpy as np
import pytest

from conftest import CONFIGS, oracle_kwargs, ssn_params

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods(built):
    from oracle.loader import Oracle
    from somatic_sniper_b200 import api, synth
    return api, synth, Oracle


def _normalise_oracle_score(score, ref_char, api):
    """The kernel sees only ref_nt16; a lower-case 'n' (nt16 15) reports -1 like 'N' does (no call either way)."""
    s = score.copy()
    s[api.nt16_lut()[ref_char] == 15] = -1
    return s


def check_batch(api, Oracle, cols, cfg="default", scorer=None):
    own = scorer is None
    sc = scorer or api.Scorer(ssn_params(cfg), n_slots=1, max_columns=max(cols["ncol"], 1),
                              max_entries=int(cols["n_entries"]) + 64)
    try:
        sites, dump = sc.score(cols, dump=True)
    finally:
        if own:
            sc.close()
    want = Oracle(**oracle_kwargs(cfg)).score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"],
                                                    want_glf=True)
    bad = np.nonzero((dump[:, :28] != want["glf"]).any(axis=1))[0]
    assert bad.size == 0, "GLF/consensus mismatch in %d columns, first %d:\n got %s\nwant %s" % (
        bad.size, bad[0], dump[bad[0], :28], want["glf"][bad[0]])
    wscore = _normalise_oracle_score(want["score"], cols["ref_char"], api)
    assert np.array_equal(dump[:, 28].astype(np.int32), wscore)
    assert np.array_equal(sites["column"].astype(np.int64), want["emit_col"].astype(np.int64)), "called-site set differs"
    got = api.sites_to_records(sites, cols["ref_char"])
    assert np.array_equal(got, want["recs"]), "site records differ"
    # RMS mapQ travels in the site too (north star: 'RMS mapQ')
    if len(sites):
        assert np.array_equal(sites["tumor"]["rms_mapq"], want["glf"][want["emit_col"], 11])
        assert np.array_equal(sites["normal"]["rms_mapq"], want["glf"][want["emit_col"], 25])
    return len(sites)


def test_known_answer_columns(mods):
    api, synth, Oracle = mods
    cols = synth.kat_columns()
    sc = api.Scorer(ssn_params("default"), n_slots=1, max_columns=8, max_entries=8192)
    sites, dump = sc.score(cols, dump=True)
    sc.close()
    # SURVEY.md App. C.3 (values dumped from the unmodified reference)
    assert list(dump[0, :14]) == [30, 0, 54, 54, 161, 167, 167, 230, 230, 230, 14, 60, 10, 0x313c0333]
    assert list(dump[1, :14]) == [238, 255, 0, 255, 255, 255, 255, 255, 255, 255, 28, 60, 58, 0x513cd311]
    assert list(dump[1, 14:28]) == [0, 90, 90, 90, 255, 255, 255, 255, 255, 255, 0, 60, 30, 0x133c7500]
    assert list(dump[2, :14]) == [0, 155, 255, 255, 255, 255, 255, 255, 255, 255, 255, 52, 1458, 0x1334b649]
    assert list(dump[2, 14:28]) == [0, 255, 255, 255, 255, 255, 255, 255, 255, 255, 0, 60, 1502, 0x123cff00]
    assert list(dump[3, :14]) == [2, 8, 8, 0, 40, 40, 29, 40, 29, 13, 8, 32, 7, 0x18200b0e]
    assert list(dump[3, 14:28]) == [0, 6, 6, 1, 74, 74, 69, 74, 69, 66, 4, 60, 3, 0x193c1c05]
    assert list(sites["column"]) == [1]
    t, n = sites["tumor"][0], sites["normal"][0]
    assert (t["somatic_score"], t["genotype"], n["genotype"], t["consensus_quality"], t["variant_allele_quality"],
            n["consensus_quality"], n["variant_allele_quality"], t["variant_status"]) == (120, 5, 1, 211, 211, 117, 0, 2)
    assert (t["total_mean_mapQ"], n["total_mean_mapQ"]) == (60, 60)
    check_batch(api, Oracle, cols)
    check_batch(api, Oracle, cols, "joint")


@pytest.mark.parametrize("cfg", list(CONFIGS))
def test_random_columns_all_configs(mods, cfg):
    api, synth, Oracle = mods
    cols = synth.make_columns(20000, seed=11, p_som=5e-3, p_het=5e-3, p_loh=2e-3, p_gor=2e-3)
    n = check_batch(api, Oracle, cols, cfg)
    assert n > 0 or cfg == "joint"


@pytest.mark.parametrize("seed,binned", [(1, True), (2, False), (3, True)])
def test_edge_case_columns(mods, seed, binned):
    api, synth, Oracle = mods
    cols = synth.make_columns(20000, seed=seed, binned=binned, edge=1.0, p_som=5e-3, p_het=5e-3, p_loh=2e-3, p_gor=2e-3)
    check_batch(api, Oracle, cols, "nopriors_q0")
    check_batch(api, Oracle, cols, "joint_s")


def test_ragged_and_boundary_depths(mods):
    api, synth, Oracle = mods
    rng = np.random.default_rng(5)
    depths = [1, 2, 3, 31, 32, 33, 63, 64, 65, 127, 128, 129, 255, 256, 257, 511, 512, 513, 1023, 1024, 1025, 2047,
              2048, 2049, 3000, 5000]
    cols = []
    for i, d in enumerate(depths):
        for nn in (1, d, max(1, d // 2)):
            def ent(k, alt_frac):
                out = []
                for _ in range(k):
                    b = "ACGT"[(i + (rng.random() < alt_frac)) & 3]
                    out.append((b, int(rng.choice([2, 12, 23, 37, 41])), int(rng.choice([0, 20, 60, 60, 60])), int(rng.random() < 0.5)))
                return out
            cols.append(("ACGT"[i & 3], ent(d, 0.3), ent(nn, 0.02)))
    batch = synth.columns_from_lists(cols)
    check_batch(api, Oracle, batch, "nopriors_q0")
    check_batch(api, Oracle, batch, "joint_s")


def test_deep_panel_columns(mods):
    api, synth, Oracle = mods
    cols = synth.make_columns(600, depth_t=1500, depth_n=1200, seed=21, binned=False, p_som=2e-2, p_het=2e-2)
    check_batch(api, Oracle, cols, "nopriors_q0")
    cols = synth.make_columns(300, depth_t=2000, depth_n=1000, seed=22, binned=True, edge=0.5, p_som=2e-2)
    check_batch(api, Oracle, cols, "default")


def test_deep_batches_in_flight_on_several_slots(mods):
    """Panel-depth batches (key store in the global scratch) on 4 slots whose kernels overlap: every slot owns its
    scratch, so concurrent launches cannot sort in each other's keys.  Sites and dumps against the oracle."""
    api, synth, Oracle = mods
    batches = [synth.make_columns(400, depth_t=1800 + 100 * i, depth_n=1100, seed=80 + i, binned=False, p_som=3e-2, p_het=2e-2)
               for i in range(8)]
    cap = max(int(b["n_entries"]) for b in batches) + 64
    want = [Oracle(min_somatic_qual=0, use_priors=0).score_batch(b["off"], b["bs"], b["bq"], b["mq"], b["ref_char"], want_glf=True)
            for b in batches]
    for flags in (api.SUBMIT_NO_PREFILTER | api.SUBMIT_COLUMN_DUMP, 0):
        for rounds in range(2):
            sc = api.Scorer(ssn_params("nopriors_q0"), device=0, n_slots=4, max_columns=512, max_entries=cap)
            got, fifo = [], []
            want_dump = bool(flags & api.SUBMIT_COLUMN_DUMP)

            def drain():
                s, _ = sc.results_next()
                b = fifo.pop(0)
                got.append((s, b.column_dump() if want_dump else None))

            for cols in batches:
                if len(fifo) == 4:
                    drain()
                b = sc.acquire(); b.fill(cols); sc.submit(b, flags); fifo.append(b)
            while fifo:
                drain()
            sc.close()
            for (s, d), w, cols in zip(got, want, batches):
                assert np.array_equal(s["column"].astype(np.int64), w["emit_col"].astype(np.int64))
                assert np.array_equal(api.sites_to_records(s, cols["ref_char"]), w["recs"])
                if d is not None:
                    assert np.array_equal(d[:, :28], w["glf"])
    assert sum(len(w["emit_col"]) for w in want) > 10


def test_empty_and_single(mods):
    api, synth, Oracle = mods
    sc = api.Scorer(ssn_params("default"), n_slots=2, max_columns=64, max_entries=4096)
    empty = dict(off=np.zeros(1, np.uint32), bs=np.zeros(0, np.uint8), bq=np.zeros(0, np.uint8), mq=np.zeros(0, np.uint8),
                 ref_char=np.zeros(0, np.uint8), ncol=0, n_entries=0)
    sites, _ = sc.score(empty)
    assert len(sites) == 0
    one = synth.columns_from_lists([("A", [("C", 40, 60, 0)] * 12 + [("C", 40, 60, 1)] * 9, [("A", 40, 60, 0)] * 8 + [("A", 38, 60, 1)] * 9)])
    check_batch(api, Oracle, one, scorer=sc)
    # a column where one sample has no bases scores nothing
    half = synth.columns_from_lists([("A", [("C", 40, 60, 0)] * 5, []), ("A", [], [("A", 40, 60, 0)] * 5)])
    sites, dump = sc.score(half, dump=True)
    assert len(sites) == 0 and list(dump[:, 28].astype(np.int32)) == [-1, -1]
    sc.close()


def test_full_size_properties(mods):
    """BASELINE config #2 size (1 Mb, 30x/60x): size-independent properties + oracle on a sample of tiles."""
    api, synth, Oracle = mods
    cols = synth.make_columns(1_000_000, seed=1234)
    sc = api.Scorer(ssn_params("default"), n_slots=2, max_columns=1_000_000, max_entries=int(cols["n_entries"]) + 64)
    s1, d1 = sc.score(cols, dump=True)
    s2, d2 = sc.score(cols, dump=True)
    assert np.array_equal(d1, d2) and np.array_equal(s1, s2)          # idempotent / deterministic
    # order of read-bases inside a sample is not observable: shuffle each sample segment
    rng = np.random.default_rng(9)
    off = cols["off"].astype(np.int64)
    seg = np.repeat(np.arange(len(off) - 1), np.diff(off))
    perm = np.lexsort((rng.random(len(seg)), seg))
    shuf = dict(cols, bs=cols["bs"][perm], bq=cols["bq"][perm], mq=cols["mq"][perm])
    s3, d3 = sc.score(shuf, dump=True)
    assert np.array_equal(d1, d3) and np.array_equal(s1, s3)
    sc.close()
    # oracle on every 16th block of 4096 columns
    o = Oracle()
    for start in range(0, cols["ncol"], 16 * 4096):
        c0, c1 = start, min(start + 4096, cols["ncol"])
        e0, e1 = int(off[2 * c0]), int(off[2 * c1])
        want = o.score_batch((off[2 * c0:2 * c1 + 1] - e0).astype(np.uint32), cols["bs"][e0:e1], cols["bq"][e0:e1],
                             cols["mq"][e0:e1], cols["ref_char"][c0:c1], want_glf=True)
        assert np.array_equal(d1[c0:c1, :28], want["glf"])
        sel = (s1["column"] >= c0) & (s1["column"] < c1)
        assert np.array_equal(s1["column"][sel] - c0, want["emit_col"])


# ---- committed golden fixtures (outputs of the UNMODIFIED reference, tests/golden/make_golden.py) -------------

import os

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _npz_cols(name):
    c = dict(np.load(os.path.join(GOLD, name)))
    c.pop("pos", None)
    c["ncol"] = len(c["ref_char"])
    c["n_entries"] = len(c["bs"])
    return c


def _check_against_golden(api, cols, want, cfg):
    sc = api.Scorer(ssn_params(cfg), n_slots=1, max_columns=cols["ncol"], max_entries=cols["n_entries"] + 64)
    sites, dump = sc.score(cols, dump=True)
    sc.close()
    assert np.array_equal(dump[:, :28], want["glf"])
    assert np.array_equal(dump[:, 28].astype(np.int32), _normalise_oracle_score(want["score"], cols["ref_char"], api))
    assert np.array_equal(sites["column"].astype(np.int64), want["emit_col"].astype(np.int64))
    assert np.array_equal(api.sites_to_records(sites, cols["ref_char"])[:, :53], want["recs"][:, :53])


@pytest.mark.parametrize("cfg", list(CONFIGS))
def test_reference_golden_columns(mods, cfg):
    api, _, _ = mods
    _check_against_golden(api, _npz_cols("columns_in.npz"), np.load(os.path.join(GOLD, "columns_%s.npz" % cfg)), cfg)


@pytest.mark.parametrize("cfg", ["default", "joint_s", "nopriors_q0"])
def test_reference_integration_columns(mods, cfg):
    """The pileup columns of the reference's integration-test BAM pair; 'default' holds the 3 expected.vcf sites."""
    api, _, _ = mods
    want = np.load(os.path.join(GOLD, "integration_%s.npz" % cfg))
    _check_against_golden(api, _npz_cols("integration.npz"), want, cfg)
    if cfg == "default":
        pos = np.load(os.path.join(GOLD, "integration.npz"))["pos"]
        assert [int(pos[c]) + 1 for c in want["emit_col"]] == [63, 73, 260]


# ---- the exact hom-reference pre-filter (ssn_prefilter.cuh) ------------------------------------

PF_CASES = [
    ("wgs", dict(ncol=400000, seed=41)),
    ("unbinned", dict(ncol=150000, seed=42, binned=False)),
    ("edge", dict(ncol=150000, seed=43, edge=0.5, p_som=5e-3)),
    ("shallow", dict(ncol=150000, depth_t=12, depth_n=6, seed=44, edge=0.1)),
    ("very shallow", dict(ncol=150000, depth_t=3, depth_n=2, seed=45)),
    ("mid", dict(ncol=40000, depth_t=110, depth_n=100, seed=46, binned=False)),
    ("variant rich", dict(ncol=100000, seed=47, p_het=0.05, p_hom=0.02, p_som=0.05, p_loh=0.01, p_gor=0.01)),
    ("deep", dict(ncol=400, depth_t=1500, depth_n=900, seed=48)),
    ("ragged", dict(ncol=100000, depth_t=150, depth_n=20, seed=49, min_depth=0)),   # several staging passes per tile, empty samples
    ("q0", dict(ncol=100000, seed=50)),                                              # quality-0 read-bases (N calls of real BAMs)
]


@pytest.mark.parametrize("name,kw", PF_CASES, ids=[c[0] for c in PF_CASES])
def test_prefilter_audit_and_identical_sites(mods, name, kw):
    """(1) Every column the pre-filter kernel drops has, in the full model's own dump, the homozygous reference as the
    consensus of both samples - so glf_somatic could not have emitted it.  (2) The product path (filter + list-mode
    scoring) returns exactly the sites of the all-columns path.  (3) The kernel drops what the numpy restatement drops."""
    import prefilter_np
    api, synth, Oracle = mods
    kw = dict(kw)
    n = kw.pop("ncol")
    cols = synth.make_columns(n, **kw)
    if name == "q0":
        rng = np.random.default_rng(3)
        zero = ((cols["bs"] & 15) == 15) | (rng.random(len(cols["bq"])) < 2e-3)
        cols["bq"] = np.where(zero, 0, cols["bq"]).astype(np.uint8)
    ent = int(cols["n_entries"]) + 64
    for joint in (0, 1):
        prm = api.default_params(use_joint_priors=joint, min_somatic_qual=0 if joint else 15)
        sc = api.Scorer(prm, device=0, n_slots=1, max_columns=n, max_entries=ent)
        assert sc.prefilter_enabled()
        _, dump = sc.score(cols, dump=True)
        dropped = dump[:, 29] == 1
        tg, ng = prefilter_np.consensus_genotypes(dump)
        ref16 = api.nt16_lut()[cols["ref_char"]]
        bad = dropped & ((tg != ref16) | (ng != ref16))
        assert bad.sum() == 0, "dropped columns with a non-reference consensus: %s" % np.nonzero(bad)[0][:5]
        full, _ = sc.score(cols, flags=api.SUBMIT_NO_PREFILTER)
        assert sc.last_survivors == 0
        fast, _ = sc.score(cols)
        survivors = sc.last_survivors
        sc.close()
        assert fast.tobytes() == full.tobytes(), "the filtered path returned different sites"
        assert not dropped[full["column"]].any()
        if joint == 0:
            want = prefilter_np.decide(cols)
            wanted = (np.diff(cols["off"].astype(np.int64))[0::2] > 0) & (np.diff(cols["off"].astype(np.int64))[1::2] > 0) & (ref16 != 15)
            # a batch of panel depth skips the filter altogether (it could not settle a column): nothing is "handed on"
            assert survivors == (0 if name == "deep" else int((wanted & ~dropped).sum()))
            # float rounding (device: fp32 + MUFU.RCP, numpy: fp64) may flip columns that sit exactly on a bound
            # (the kernel's clean-tile check also sees the neighbouring tiles' alignment bytes, so only compare where every
            # base quality is legal)
            if name in ("wgs", "unbinned", "very shallow", "mid", "variant rich", "deep", "ragged", "q0"):
                assert int((dropped ^ want).sum()) <= 3, (dropped.sum(), want.sum())
            if name in ("wgs", "unbinned", "mid", "ragged", "q0"):
                assert dropped.mean() > 0.95


def test_prefilter_through_the_pipeline(mods):
    """Several batches in flight through acquire/submit/results_next with the filter on: same sites as with it off."""
    api, synth, _ = mods
    batches = [synth.make_columns(50000 + 1000 * i, seed=60 + i, edge=0.02 * i) for i in range(5)]
    cap = max(int(b["n_entries"]) for b in batches) + 64
    out = {}
    for flags in (0, api.SUBMIT_NO_PREFILTER):
        sc = api.Scorer(api.default_params(), device=0, n_slots=3, max_columns=60000, max_entries=cap)
        got, pending = [], 0
        for cols in batches:
            if pending == 3:
                s, h = sc.results_next(); got.append(s); pending -= 1
            b = sc.acquire(); b.fill(cols); sc.submit(b, flags); pending += 1
        while pending:
            s, h = sc.results_next(); got.append(s); pending -= 1
        sc.close()
        out[flags] = got
    for a, b in zip(out[0], out[api.SUBMIT_NO_PREFILTER]):
        assert a.tobytes() == b.tobytes()
    assert sum(len(a) for a in out[0]) > 0


def test_prefilter_off_when_a_precondition_fails(mods):
    """A heterozygote prior so large that the consensus penalty q_r is not positive: the proof does not apply, the ctx runs
    every column through the full model, and the result still matches the oracle."""
    api, synth, Oracle = mods
    cols = synth.make_columns(30000, seed=71, p_som=5e-3, p_het=5e-3)
    sc = api.Scorer(api.default_params(het_rate=0.4), device=0, n_slots=1, max_columns=30000, max_entries=int(cols["n_entries"]) + 64)
    assert not sc.prefilter_enabled()
    sites, _ = sc.score(cols)
    assert sc.last_survivors == 0
    sc.close()
    want = Oracle(het_rate=0.4).score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"])
    assert np.array_equal(sites["column"].astype(np.int64), want["emit_col"].astype(np.int64))
    assert np.array_equal(api.sites_to_records(sites, cols["ref_char"]), want["recs"])


@pytest.mark.parametrize("name,kw", [("default", dict()), ("TNr", dict(theta=0.7, n_hap=3, het_rate=0.01)), ("theta95", dict(theta=0.95))])
def test_prefilter_kernel_on_the_boundary_columns(mods, name, kw):
    """tests/golden/prefilter_boundary.npz: columns hill-climbed (9 million filter evaluations per parameter set,
    tests/test_prefilter_fuzz.py) to within a few hundredths of the filter's decision boundary.  The kernel evaluates its
    bounds in fp32 (MUFU.RCP division, 1e-4 slack): whatever it settles here must be hom-reference in both samples by the
    full model's own dump, the product path must equal the every-column path, and the oracle must agree."""
    import prefilter_np
    api, synth, Oracle = mods
    z = np.load(os.path.join(GOLD, "prefilter_boundary.npz"))
    cols = {k: z["%s_%s" % (name, k)] for k in ("off", "bs", "bq", "mq", "ref_char")}
    cols["ncol"], cols["n_entries"] = len(cols["ref_char"]), len(cols["bs"])
    prm = api.default_params(min_somatic_qual=0, **kw)
    sc = api.Scorer(prm, device=0, n_slots=1, max_columns=cols["ncol"], max_entries=cols["n_entries"] + 64)
    assert sc.prefilter_enabled()
    _, dump = sc.score(cols, dump=True)
    full, _ = sc.score(cols, flags=api.SUBMIT_NO_PREFILTER)
    fast, _ = sc.score(cols)
    sc.close()
    dropped = dump[:, 29] == 1
    tg, ng = prefilter_np.consensus_genotypes(dump)
    ref16 = api.nt16_lut()[cols["ref_char"]]
    assert not (dropped & ((tg != ref16) | (ng != ref16))).any()
    assert fast.tobytes() == full.tobytes()
    assert dropped.mean() > 0.5, "most boundary columns are still inside for the fp32 evaluation (%.3f)" % dropped.mean()
    want = Oracle(min_somatic_qual=0, **kw).score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"], want_glf=True)
    assert np.array_equal(dump[:, :28], want["glf"])
