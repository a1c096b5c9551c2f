"""Synthetic reads and the read-window layout the device pileup takes (ssn_batch_submit_reads, include/ssn.h).

Test and benchmark input only: `make_reads` fabricates coordinate-sorted reads of one sample over a reference,
`build_read_windows` files both samples' reads under windows of 8192 positions the way the BAM reader of the CLI does
(csrc/host/ssn_bamread.c: a read appears in every window it overlaps, in file order) and lays the records out as
ssn_read_rec + cigar + 4-bit sequence + qualities.  No pileup logic lives here."""
import numpy as np

from .api import READ_WINDOW, READ_WINDOW_DTYPE

W_SHIFT = 13
OP_M, OP_I, OP_D, OP_N, OP_S, OP_H, OP_P, OP_EQ, OP_X = range(9)


def _ranges(lens):
    """0..len-1 for every len, concatenated."""
    lens = np.asarray(lens, np.int64)
    total = int(lens.sum())
    start = np.cumsum(lens) - lens
    return np.arange(total, dtype=np.int64) - np.repeat(start, lens)


def make_reference(seed, length, n_frac=0.002):
    """nt16 codes (1, 2, 4, 8; 15 = N) of a random reference."""
    rng = np.random.default_rng(seed)
    ref = (1 << rng.integers(0, 4, length)).astype(np.uint8)
    ref[rng.random(length) < n_frac] = 15
    return ref


def make_reads(seed, ref, depth, read_len=150, start=0, length=None, complex_cigars=False, err=0.01, q0_frac=0.0,
               low_mapq_frac=0.05):
    """Coordinate-sorted reads of one sample: dict(beg, mapq, strand, l_qseq, cig_off, cig, seq_off, seq, qual)."""
    rng = np.random.default_rng(seed)
    length = len(ref) - start if length is None else length
    n = max(1, int(depth * length / read_len))
    beg = np.sort(rng.integers(max(0, start - read_len + 1), start + length, n)).astype(np.int32)
    kind = rng.integers(0, 10, n) if complex_cigars is True else np.zeros(n, np.int64)
    a = rng.integers(10, read_len - 20, n)
    cigs, lq = [], np.zeros(n, np.uint32)
    for i in range(n):
        k, L, x = int(kind[i]), read_len, int(a[i])
        if complex_cigars == "random":
            # any operator, any order, lengths from 0: what a hostile or merely odd BAM can hold
            ops = rng.integers(0, 9, int(rng.integers(1, 9)))
            c = [(int(op), int(rng.integers(0, 41))) for op in ops]
            k = 7 if rng.random() < 0.3 else -1          # l_qseq unrelated to the CIGAR in 30 % of the reads
            L = int(rng.integers(0, 120))
        elif k == 1:
            c = [(OP_S, 5), (OP_M, L - 5)]
        elif k == 2:
            c = [(OP_M, x), (OP_I, 3), (OP_M, L - x - 3)]
        elif k == 3:
            c = [(OP_M, x), (OP_D, 4), (OP_M, L - x)]
        elif k == 4:
            c = [(OP_M, x), (OP_N, 200), (OP_M, L - x)]
        elif k == 5:
            c = [(OP_H, 10), (OP_M, L), (OP_H, 3)]
        elif k == 6:
            c = [(OP_M, x), (OP_EQ, 20), (OP_X, 2), (OP_P, 1), (OP_M, L - x)]     # =, X, P advance nothing (QUIRK B6)
        elif k == 7:
            c = [(OP_M, L + 10)]                                                   # CIGAR longer than the read
        elif k == 8:
            c = [(OP_S, 3), (OP_M, x), (OP_D, 1), (OP_M, 7), (OP_I, 1), (OP_M, L - x - 11), (OP_S, 0)]
        else:
            c = [(OP_M, L)]
        cigs.append(c)
        lq[i] = L if k == 7 else sum(l for op, l in c if op in (OP_M, OP_I, OP_S))
    ncig = np.array([len(c) for c in cigs], np.int64)
    cig = np.array([(l << 4) | op for c in cigs for op, l in c], np.uint32)
    cig_off = np.concatenate([[0], np.cumsum(ncig)]).astype(np.int64)
    seq_off = np.concatenate([[0], np.cumsum(lq.astype(np.int64))])
    total = int(seq_off[-1])
    # bases: the reference under an ungapped placement (good enough for a pileup test), with errors
    ridx = np.repeat(beg.astype(np.int64), lq.astype(np.int64)) + _ranges(lq)
    seq = ref[np.minimum(ridx, len(ref) - 1)].copy()
    flip = rng.random(total) < err
    seq[flip] = (1 << rng.integers(0, 4, int(flip.sum()))).astype(np.uint8)
    seq[rng.random(total) < 0.001] = 15
    qual = rng.integers(2, 42, total).astype(np.uint8)
    if q0_frac:
        qual[rng.random(total) < q0_frac] = 0
    mapq = np.where(rng.random(n) < low_mapq_frac, rng.integers(0, 60, n), 60).astype(np.uint8)
    strand = rng.integers(0, 2, n).astype(np.uint8)
    return dict(beg=beg, mapq=mapq, strand=strand, l_qseq=lq, cig_off=cig_off, cig=cig, seq_off=seq_off, seq=seq, qual=qual)


def calend(reads):
    """bam_calend: start + reference length of the M/D/N runs (samtools-0.1.6 bam.c)."""
    op, ln = reads["cig"] & 0xf, (reads["cig"] >> 4).astype(np.int64)
    adv = np.where((op == OP_M) | (op == OP_D) | (op == OP_N), ln, 0)
    cs = np.concatenate([[0], np.cumsum(adv)])
    return reads["beg"].astype(np.int64) + cs[reads["cig_off"][1:]] - cs[reads["cig_off"][:-1]]


def _pack_seq(reads):
    lq = reads["l_qseq"].astype(np.int64)
    plen = (lq + 1) >> 1
    poff = np.cumsum(plen) - plen
    j = _ranges(lq)
    dst = np.repeat(poff, lq) + (j >> 1)
    out = np.zeros(int(plen.sum()), np.uint8)
    even = (j & 1) == 0
    out[dst[even]] = reads["seq"][even] << 4
    out[dst[~even]] |= reads["seq"][~even] & 15
    return out, poff, plen


def build_read_windows(tumor, normal, ref, tid=0, first_window=None, n_windows=None):
    """dict(recs u8, rec_off u32, win READ_WINDOW_DTYPE, ref_nt16 u8[w, 8192], entries_bound)."""
    samples = (tumor, normal)
    ends = [calend(s) for s in samples]
    w0s = [s["beg"].astype(np.int64) >> W_SHIFT for s in samples]
    w1s = [np.maximum(e - 1, s["beg"].astype(np.int64)) >> W_SHIFT for s, e in zip(samples, ends)]
    if first_window is None:
        first_window = int(min(w.min() for w in w0s))
    if n_windows is None:
        n_windows = int(max(w.max() for w in w1s)) - first_window + 1
    ent_w, ent_s, ent_r = [], [], []
    for s in range(2):
        cnt = w1s[s] - w0s[s] + 1
        r = np.repeat(np.arange(cnt.size, dtype=np.int64), cnt)
        w = w0s[s][r] + _ranges(cnt) - first_window
        keep = (w >= 0) & (w < n_windows)
        ent_w.append(w[keep]); ent_r.append(r[keep]); ent_s.append(np.full(int(keep.sum()), s, np.int64))
    ew, es, er = np.concatenate(ent_w), np.concatenate(ent_s), np.concatenate(ent_r)
    order = np.lexsort((er, es, ew))                       # window, sample, file order
    ew, es, er = ew[order], es[order], er[order]
    n_ent = ew.size
    f = lambda name, dt: np.concatenate([samples[0][name], samples[1][name]]).astype(dt)
    nT = samples[0]["beg"].size
    g = er + es * nT                                       # index into the concatenated per-read arrays
    beg, mapq, strand, lq = f("beg", np.int64), f("mapq", np.int64), f("strand", np.int64), f("l_qseq", np.int64)
    end = np.concatenate(ends)
    ncig = np.concatenate([np.diff(s["cig_off"]) for s in samples]).astype(np.int64)
    mlen = []                                              # read-bases a read can add: the length of its M runs
    for s in samples:
        m = np.where((s["cig"] & 0xf) == OP_M, (s["cig"] >> 4).astype(np.int64), 0)
        cs = np.concatenate([[0], np.cumsum(m)])
        mlen.append(cs[s["cig_off"][1:]] - cs[s["cig_off"][:-1]])
    mlen = np.concatenate(mlen)
    cig_start = np.concatenate([samples[0]["cig_off"][:-1], samples[1]["cig_off"][:-1] + samples[0]["cig"].size])
    cig = np.concatenate([samples[0]["cig"], samples[1]["cig"]])
    packed, poff, plen = [], [], []
    base = 0
    for s in samples:
        pk, po, pl = _pack_seq(s)
        packed.append(pk); poff.append(po + base); plen.append(pl)
        base += pk.size
    packed, poff, plen = np.concatenate(packed), np.concatenate(poff), np.concatenate(plen)
    qual = np.concatenate([samples[0]["qual"], samples[1]["qual"]])
    qoff = np.concatenate([samples[0]["seq_off"][:-1], samples[1]["seq_off"][:-1] + samples[0]["qual"].size])
    nbytes = (20 + 4 * ncig[g] + plen[g] + lq[g] + 3) & ~3
    rec_byte = np.cumsum(nbytes) - nbytes
    recs = np.zeros(int(nbytes.sum()), np.uint8)
    words = recs.view("<u4")
    wo = rec_byte >> 2
    words[wo] = beg[g].astype(np.int32).view(np.uint32) if n_ent else 0
    words[wo + 1] = end[g].astype(np.int32).view(np.uint32) if n_ent else 0
    words[wo + 2] = lq[g]
    words[wo + 3] = ncig[g] | (mapq[g] << 16) | (strand[g] << 24)
    words[wo + 4] = nbytes
    k = _ranges(ncig[g])
    words[np.repeat(wo + 5, ncig[g]) + k] = cig[np.repeat(cig_start[g], ncig[g]) + k]
    sdst = rec_byte + 20 + 4 * ncig[g]
    k = _ranges(plen[g])
    recs[np.repeat(sdst, plen[g]) + k] = packed[np.repeat(poff[g], plen[g]) + k]
    k = _ranges(lq[g])
    recs[np.repeat(sdst + plen[g], lq[g]) + k] = qual[np.repeat(qoff[g], lq[g]) + k]
    win = np.zeros(n_windows, READ_WINDOW_DTYPE)
    key = ew * 2 + es
    first = np.searchsorted(key, np.arange(2 * n_windows))
    last = np.searchsorted(key, np.arange(2 * n_windows), side="right")
    win["rec_first"] = first.reshape(n_windows, 2)
    win["n_reads"] = (last - first).reshape(n_windows, 2)
    win["lo"] = (np.arange(n_windows) + first_window) << W_SHIFT
    win["tid"] = tid
    wref = np.full((n_windows, READ_WINDOW), 15, np.uint8)
    lo0 = first_window << W_SHIFT
    seg = ref[lo0:lo0 + n_windows * READ_WINDOW]
    wref.reshape(-1)[:seg.size] = seg
    return dict(recs=recs, rec_off=wo.astype(np.uint32), win=win, ref_nt16=wref, entries_bound=int(np.maximum(lq, mlen)[g].sum()) + 64)


def pileup_numpy(tumor, normal, ref, first_window, n_windows):
    """Straight restatement of the column construction for small inputs (tests only): a read-base per position of
    every M run (sniper_pileup.c:57-104), columns where both samples have one and the reference is callable.
    Returns the column dict of hostapi.pack_read_windows."""
    lo0, hi0 = first_window << W_SHIFT, (first_window + n_windows) << W_SHIFT
    per = []
    for s in (tumor, normal):
        cols = {}
        for r in range(s["beg"].size):
            x, y = int(s["beg"][r]), 0
            l_q, so = int(s["l_qseq"][r]), int(s["seq_off"][r])
            for c in s["cig"][s["cig_off"][r]:s["cig_off"][r + 1]]:
                op, l = int(c) & 15, int(c) >> 4
                if op == OP_M:
                    for p in range(max(x, lo0), min(x + l, hi0)):
                        q = y + p - x
                        b = (int(s["seq"][so + q]) if q < l_q else 15) | (int(s["strand"][r]) << 4)
                        cols.setdefault(p, []).append((b, int(s["qual"][so + q]) if q < l_q else 0, int(s["mapq"][r])))
                    x += l; y += l
                elif op in (OP_D, OP_N):
                    x += l
                elif op in (OP_I, OP_S):
                    y += l
        per.append(cols)
    off, ent, refc, pos = [0], [], [], []
    for p in sorted(set(per[0]) & set(per[1])):
        if p >= len(ref) or ref[p] == 15:
            continue
        ent += per[0][p]
        off.append(len(ent))
        ent += per[1][p]
        off.append(len(ent))
        refc.append(int(ref[p])); pos.append(p - lo0)
    e = np.array(ent, np.uint8).reshape(-1, 3)
    return dict(ncol=len(pos), n_entries=len(ent), off=np.array(off, np.uint32), bs=e[:, 0].copy(), bq=e[:, 1].copy(),
                mq=e[:, 2].copy(), ref_nt16=np.array(refc, np.uint8), pos=np.array(pos, np.uint32))
