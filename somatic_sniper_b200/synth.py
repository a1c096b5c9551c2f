"""Deterministic synthetic pileup-column batches in the packer's SoA layout.

A batch is what the host packer hands to the scorer (include/ssn.h, ``ssn_batch``):
    off      uint32[2*ncol+1]  off[2c]..off[2c+1] = tumor entries, off[2c+1]..off[2c+2] = normal entries
    bs       uint8[n]          nt16 base code | strand<<4
    bq       uint8[n]          base quality
    mq       uint8[n]          mapping quality
    ref_char uint8[ncol]       reference base character (ASCII, case preserved)

The generator draws columns directly (no reads): Poisson depth around the nominal coverage, an
Illumina-like base-quality mixture (binned {2,12,23,37} or unbinned 2..41), the mapQ mixture of
SURVEY.md §8(d), substitution errors at 10^(-Q/10), germline het/hom sites in both samples, somatic
SNVs at tumor VAF in {0.1,0.25,0.4}, LOH and gain-of-reference sites so -L/-G and SS=3 fire.
``edge`` > 0 additionally sprinkles the reference's corner cases (N / '=' / IUPAC read bases, mapQ 0,
baseQ 0 and >= 64, reference N and lower case, strand-skewed and single-read columns).
"""
import numpy as np

NT16 = {"A": 1, "C": 2, "G": 4, "T": 8, "N": 15, "=": 0}
_ACGT = np.array([1, 2, 4, 8], np.uint8)
_ACGT_CH = np.frombuffer(b"ACGT", np.uint8)


def make_columns(ncol, depth_t=60.0, depth_n=30.0, seed=1234, binned=True, edge=0.0,
                 p_het=1e-3, p_hom=1e-4, p_som=2e-4, p_loh=2e-5, p_gor=2e-5, min_depth=1, max_depth=None):
    """Returns dict(off, bs, bq, mq, ref_char, ncol, n_entries, truth)."""
    rng = np.random.default_rng(seed)
    nT = np.maximum(rng.poisson(depth_t, ncol), min_depth).astype(np.int64)
    nN = np.maximum(rng.poisson(depth_n, ncol), min_depth).astype(np.int64)
    if max_depth is not None:
        nT = np.minimum(nT, max_depth)
        nN = np.minimum(nN, max_depth)
    if edge > 0:
        k = rng.random(ncol) < edge * 0.2
        nT[k] = rng.integers(1, 4, k.sum())
        k = rng.random(ncol) < edge * 0.2
        nN[k] = rng.integers(1, 4, k.sum())
    cnt = np.empty(2 * ncol, np.int64)
    cnt[0::2] = nT
    cnt[1::2] = nN
    off = np.zeros(2 * ncol + 1, np.int64)
    np.cumsum(cnt, out=off[1:])
    n = int(off[-1])
    assert n < 2 ** 32
    seg = np.repeat(np.arange(2 * ncol), cnt)          # entry -> (column, sample) segment
    col = seg >> 1
    is_t = (seg & 1) == 0

    ref_idx = rng.integers(0, 4, ncol)
    alt_idx = (ref_idx + rng.integers(1, 4, ncol)) & 3
    # per-column genotype model: alt allele fraction in tumor / normal
    u = rng.random(ncol)
    vaf_t = np.zeros(ncol)
    vaf_n = np.zeros(ncol)
    kind = np.zeros(ncol, np.int8)
    edges = np.cumsum([p_het, p_hom, p_som, p_loh, p_gor])
    het = u < edges[0]
    hom = (u >= edges[0]) & (u < edges[1])
    som = (u >= edges[1]) & (u < edges[2])
    loh = (u >= edges[2]) & (u < edges[3])
    gor = (u >= edges[3]) & (u < edges[4])
    vaf_t[het] = 0.5; vaf_n[het] = 0.5; kind[het] = 1
    vaf_t[hom] = 1.0; vaf_n[hom] = 1.0; kind[hom] = 2
    vaf_t[som] = rng.choice([0.1, 0.25, 0.4], som.sum()); kind[som] = 3
    vaf_t[loh] = 1.0; vaf_n[loh] = 0.5; kind[loh] = 4
    vaf_t[gor] = 0.5; vaf_n[gor] = 1.0; kind[gor] = 5

    vaf = np.where(is_t, vaf_t[col], vaf_n[col])
    base_idx = np.where(rng.random(n) < vaf, alt_idx[col], ref_idx[col])
    if binned:
        bq = rng.choice(np.array([2, 12, 23, 37], np.uint8), n, p=[0.02, 0.05, 0.13, 0.80])
    else:
        bq = np.clip(rng.normal(34, 6, n).round(), 2, 41).astype(np.uint8)
    err = rng.random(n) < np.power(10.0, -bq.astype(np.float64) / 10.0)
    base_idx = np.where(err, (base_idx + rng.integers(1, 4, n)) & 3, base_idx)
    nt = _ACGT[base_idx]
    strand = (rng.random(n) < 0.5).astype(np.uint8)
    r = rng.random(n)
    mq = np.full(n, 60, np.uint8)
    m = (r >= 0.85) & (r < 0.93)
    mq[m] = rng.integers(40, 60, m.sum())
    m = (r >= 0.93) & (r < 0.98)
    mq[m] = rng.integers(1, 40, m.sum())
    mq[r >= 0.98] = 0
    # N read bases at Q2 (0.05 %)
    m = rng.random(n) < 5e-4
    nt[m] = 15
    bq[m] = 2
    ref_char = _ACGT_CH[ref_idx].copy()

    if edge > 0:
        def pick(frac):
            return rng.random(n) < edge * frac
        m = pick(0.02); nt[m] = 0                                   # '=' read bases
        m = pick(0.02); nt[m] = rng.choice([3, 5, 6, 7, 9, 10, 11, 12, 13, 14], m.sum())  # IUPAC codes
        m = pick(0.03); bq[m] = 0
        m = pick(0.02); bq[m] = rng.choice([64, 65, 93, 128, 192, 200, 255], m.sum())
        m = pick(0.03); mq[m] = rng.choice([0, 1, 2, 3, 4, 61, 100, 127, 128, 200, 255], m.sum())
        m = pick(0.03); bq[m] = rng.integers(0, 6, m.sum())
        cm = rng.random(ncol) < edge * 0.05
        ref_char[cm] = rng.choice(np.frombuffer(b"NnacgtRYMKswBDHV=x", np.uint8), cm.sum())
        # strand-skewed columns
        cm = rng.random(ncol) < edge * 0.1
        strand[cm[col]] = 1
        # low-complexity quality columns (many exact ties)
        cm = rng.random(ncol) < edge * 0.1
        bq[cm[col]] = 30
        mq[cm[col]] = 60

    bs = (nt | (strand << 4)).astype(np.uint8)
    return dict(off=off.astype(np.uint32), bs=bs, bq=bq.astype(np.uint8), mq=mq, ref_char=ref_char,
                ncol=ncol, n_entries=n, truth=kind)


def algorithmic_bytes(batch):
    """B_col = 3*(n_T+n_N) + 8 + 1, summed over the batch."""
    return 3 * int(batch["n_entries"]) + 9 * int(batch["ncol"])


def columns_from_lists(cols):
    """cols: list of (ref_char, tumor_entries, normal_entries); entries = list of (base_char|nt16, baseQ, mapQ, strand)."""
    off = [0]
    bs, bq, mq, ref = [], [], [], []
    for ref_c, t, nn in cols:
        for ent in (t, nn):
            for b, q, m, s in ent:
                code = NT16[b] if isinstance(b, str) else int(b)
                bs.append(code | (int(s) << 4)); bq.append(q); mq.append(m)
            off.append(len(bs))
        ref.append(ord(ref_c) if isinstance(ref_c, str) else int(ref_c))
    return dict(off=np.array(off, np.uint32), bs=np.array(bs, np.uint8), bq=np.array(bq, np.uint8),
                mq=np.array(mq, np.uint8), ref_char=np.array(ref, np.uint8), ncol=len(cols), n_entries=len(bs))


def kat_columns():
    """Known-answer columns KAT0..KAT3, reference base A."""
    A, Cc, G, T, N, EQ = "A", "C", "G", "T", "N", "="
    kat0 = [((A if i < 8 else Cc), 30, 60, i & 1) for i in range(10)]
    kat1_t = [(A, 30, 60, 0)] * 20 + [(A, 30, 60, 1)] * 20 + [(G, 35, 60, 0)] * 10 + [(G, 35, 60, 1)] * 8
    kat1_n = [(A, 30, 60, 0)] * 15 + [(A, 30, 60, 1)] * 15
    kat2_t = ([(A, 37, 60, 0)] * 700 + [(A, 25, 40, 1)] * 600 + [(Cc, 37, 60, 0)] * 90 + [(Cc, 12, 29, 1)] * 60
              + [(N, 2, 60, 0)] * 5 + [(T, 0, 60, 1)] * 3)
    kat2_n = [(A, 37, 60, 0)] * 800 + [(A, 30, 60, 1)] * 700 + [(Cc, 20, 60, 0)] * 2
    kat3_t = [(T, 2, 0, 0)] * 3 + [(T, 40, 3, 1)] * 2 + [(N, 2, 60, 0)] + [(EQ, 30, 60, 1)]
    kat3_n = [(A, 40, 60, 0)] * 2 + [(T, 3, 60, 1)]
    return columns_from_lists([("A", kat0, kat0), ("A", kat1_t, kat1_n), ("A", kat2_t, kat2_n), ("A", kat3_t, kat3_n)])
