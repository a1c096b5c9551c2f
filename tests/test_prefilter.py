"""The exact hom-reference pre-filter: CPU checks of the bound tables and of the decision rule against the
oracle.  A dropped column must be one where glf_somatic cannot emit: both consensus genotypes are the homozygous
reference (somatic_sniper.c:156).  The GPU audit of the kernel's own decisions is in tests/test_gpu_parity.py."""
import numpy as np
import pytest

from somatic_sniper_b200 import api, synth
from oracle.loader import Oracle
import prefilter_np


def _audit(cols, orc=None, tables=None):
    drop = prefilter_np.decide(cols, tables, tile_rules=False)
    orc = orc or Oracle(min_somatic_qual=0)
    want = orc.score_batch(cols["off"], cols["bs"], cols["bq"], cols["mq"], cols["ref_char"], want_glf=True)
    tg, ng = prefilter_np.consensus_genotypes(want["glf"])
    ref16 = api.nt16_lut()[cols["ref_char"]]
    bad = drop & ((tg != ref16) | (ng != ref16))
    emitted = np.zeros(cols["ncol"], bool)
    emitted[want["emit_col"]] = True
    return drop, bad, emitted


CASES = [
    ("wgs binned", dict(ncol=60000, seed=11)),
    ("wgs unbinned", dict(ncol=40000, seed=12, binned=False)),
    ("edge cases", dict(ncol=40000, seed=13, edge=0.5, p_som=5e-3)),
    ("shallow", dict(ncol=40000, depth_t=12, depth_n=6, seed=14, edge=0.1)),
    ("very shallow", dict(ncol=40000, depth_t=3, depth_n=2, seed=15)),
    ("mid depth", dict(ncol=15000, depth_t=110, depth_n=100, seed=16, binned=False)),
    ("variant rich", dict(ncol=30000, seed=17, p_het=0.05, p_hom=0.02, p_som=0.05, p_loh=0.01, p_gor=0.01)),
]


@pytest.mark.parametrize("name,kw", CASES, ids=[c[0] for c in CASES])
def test_dropped_columns_are_hom_reference_in_both_samples(name, kw):
    kw = dict(kw)
    cols = synth.make_columns(kw.pop("ncol"), **kw)
    drop, bad, emitted = _audit(cols)
    assert bad.sum() == 0, "pre-filter dropped columns whose consensus is not hom-reference: %s" % np.nonzero(bad)[0][:5]
    assert not (drop & emitted).any()
    if name.startswith("wgs"):
        assert drop.mean() > 0.95, "the filter should settle the bulk of WGS columns (got %.3f)" % drop.mean()


def _adversarial(seed, ncol=6000):
    """Columns built to sit near the filter's decision boundary: low and mixed qualities, heavy strand skew,
    a few high-quality non-reference read-bases, low mapping qualities, depths 1..120."""
    rng = np.random.default_rng(seed)
    cols = []
    acgt = "ACGT"
    for _ in range(ncol):
        ref = acgt[rng.integers(4)]
        ent = []
        for depth_hi in (120, 60):
            n = int(rng.integers(1, depth_hi))
            skew = rng.choice([0.5, 0.5, 0.02, 0.98, 0.2])
            qmode = rng.integers(5)
            n_alt = int(rng.choice([0, 0, 1, 1, 2, 3, 5, n // 4, n // 2]))
            alt = acgt[(acgt.index(ref) + int(rng.integers(1, 4))) & 3]
            e = []
            for i in range(n):
                if qmode == 0: bq = int(rng.choice([2, 12, 23, 37, 0]))
                elif qmode == 1: bq = int(rng.integers(1, 8))
                elif qmode == 2: bq = int(rng.integers(1, 64))
                elif qmode == 3: bq = 40
                else: bq = int(rng.choice([2, 41]))
                mapq = int(rng.choice([60, 60, 60, 0, 1, 3, 20, 37, 255]))
                base = alt if i < n_alt else ref
                if rng.random() < 0.01: base = "N"
                if rng.random() < 0.3 and i < n_alt: base = acgt[rng.integers(4)]
                e.append((base, bq, mapq, int(rng.random() < skew)))
            ent.append(e)
        cols.append((ref, ent[0], ent[1]))
    return synth.columns_from_lists(cols)


@pytest.mark.parametrize("seed", [101, 102])
def test_adversarial_columns(seed):
    cols = _adversarial(seed)
    drop, bad, emitted = _audit(cols)
    assert bad.sum() == 0, np.nonzero(bad)[0][:5]
    assert not (drop & emitted).any()
    assert drop.sum() > 0


@pytest.mark.parametrize("kw", [dict(theta=0.5), dict(theta=0.95, n_hap=4), dict(het_rate=0.01), dict(n_hap=3, het_rate=0.0005)],
                         ids=["theta0.5", "theta0.95-N4", "r0.01", "N3"])
def test_other_model_parameters(kw):
    prm = api.default_params(**{k: v for k, v in kw.items()})
    tables = api.host_prefilter_tables(prm)
    orc = Oracle(theta=kw.get("theta", 0.85), het_rate=kw.get("het_rate", 0.001), n_hap=kw.get("n_hap", 2), min_somatic_qual=0)
    for cols in (synth.make_columns(15000, seed=21, edge=0.05), _adversarial(103, 3000)):
        drop, bad, emitted = _audit(cols, orc, tables)
        assert bad.sum() == 0
        assert not (drop & emitted).any()


def test_tables_are_closed_and_conservative():
    t = api.host_prefilter_tables()
    m = api.host_tables()
    coef = m["coef"].reshape(64, 256, 256)
    ab, h, F = t["ab"], t["h"], t["F"]
    assert t["enabled"]
    assert np.all(np.diff(F) > 0) and F[0] == 0
    assert np.all(F[1:] <= np.cumsum(m["fk"]) * (1 - 0.9e-4))
    # monotone closure: a larger m (more read-bases counted as non-reference) never makes a bound less conservative
    for n in (1, 7, 30, 90, 200, 255):
        assert np.all(np.diff(ab[n, :n + 1, 0]) <= 0)
        assert np.all(np.diff(ab[n, :n + 1, 1]) >= 0)
        assert np.all(np.diff(ab[n, :n + 1, 2]) <= 0)
        for mm in (0, 1, n // 2, n):
            ks = np.arange(max(n - mm, 1), n + 1)
            assert ab[n, mm, 0] <= min(0.0, coef[4:64, n, ks].min()) + 1e-4
            if mm:
                assert ab[n, mm, 1] >= coef[4:64, n, 1:mm + 1].max() - 1e-4
    # vt: upper bound of p[RR] = esum + coef[e][n][k] for every (e, k <= m) with esum <= min(s, (e + 0.5) k)
    vt = t["vt"]
    for n in (5, 22, 30, 90, 200):
        for mm in (1, 2, 5, 8):
            if mm > n:
                continue
            assert np.all(np.diff(vt[n, mm - 1]) >= 0)                      # monotone in the quality sum
            if mm > 1:
                assert np.all(vt[n, mm - 1] >= vt[n, mm - 2])                # and in m
            for sv in (0, 4, 37, 80, 255):
                for k in range(1, mm + 1):
                    for e in (4, 20, 37, 63):
                        assert vt[n, mm - 1, sv] >= min(sv, (e + 0.5) * k) + coef[e, n, k] - 1e-3
    lhet = m["lhet"].reshape(256, 256)
    for a, b in ((30, 0), (29, 1), (88, 2), (5, 5), (200, 55)):
        assert h[a, b] <= -4.343 * max(lhet[a, b], lhet[b, a]) + 1e-3
        if a + b < 255:
            assert h[a, b] <= h[a + 1, b]
        if b:
            assert h[a, b] <= h[a, b - 1]


def test_disabled_when_a_precondition_fails():
    # a heterozygote prior so large that the consensus penalty q_r is not positive: the proof does not apply
    t = api.host_prefilter_tables(api.default_params(het_rate=0.4))
    assert not t["enabled"]
    cols = synth.make_columns(2000, seed=5)
    assert not prefilter_np.decide(cols, t).any()
