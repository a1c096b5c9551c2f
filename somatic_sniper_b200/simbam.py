"""Deterministic synthetic tumor/normal BAM pairs written with a from-scratch BAM/BGZF writer.

    ref = make_reference([("chr1", 200_000), ("chr2", 50_000)], seed=1234)     # i.i.d. ACGT, optional N runs
    write_fasta(path, ref)
    truth = make_variants(ref, seed)                                           # germline het/hom, somatic, LOH, GOR
    simulate_bam(path, ref, truth, sample="tumor", depth=60, seed=...)         # 2x150 pairs, coordinate sorted

Reads: paired 2x150, insert N(400,50), flags 99/147/83/163; 0.5 % carry a 1-5 bp insertion or deletion, 2 % a soft
clip, 0.2 % an N skip, 1 % are duplicates (0x400), 0.5 % QC-fail or secondary (all filtered by the reference's mask);
mapQ mixture {60: 85 %, 40-59: 8 %, 1-39: 5 %, 0: 2 %}; base qualities binned {2,12,23,37}; substitution errors at
10^(-Q/10); N bases (Q2) at 0.05 %.  Only the CIGAR operators samtools-0.1.6 knows (M I D N S H).
"""
import struct
import zlib

import numpy as np

_ACGT = np.frombuffer(b"ACGT", np.uint8)
_NT16 = np.full(256, 15, np.uint8)
for _i, _c in enumerate(b"=ACMGRSVTWYHKDBN"):
    _NT16[_c] = _i
    _NT16[ord(chr(_c).lower())] = _i
READ_LEN = 150


def make_reference(contigs, seed=1234, n_runs=0):
    """contigs: list of (name, length) -> list of (name, uint8 array of ASCII bases)."""
    rng = np.random.default_rng(seed)
    out = []
    for name, length in contigs:
        seq = _ACGT[rng.integers(0, 4, length)].copy()
        for _ in range(n_runs):
            a = int(rng.integers(0, max(1, length - 200)))
            seq[a:a + int(rng.integers(20, 200))] = ord("N")
        out.append((name, seq))
    return out


def write_fasta(path, ref, width=60):
    with open(path, "wb") as f:
        for name, seq in ref:
            f.write(b">" + name.encode() + b"\n")
            n = len(seq)
            full = n // width * width
            if full:
                body = np.empty((full // width, width + 1), np.uint8)
                body[:, :width] = seq[:full].reshape(-1, width)
                body[:, width] = 10
                f.write(body.tobytes())
            if n > full:
                f.write(seq[full:].tobytes() + b"\n")


def make_variants(ref, seed=99, het_every=1000, hom_every=10000, som_every=5000, loh_every=50000, gor_every=50000):
    """Per contig: dict(pos, alt (ASCII), vaf_t, vaf_n) arrays."""
    rng = np.random.default_rng(seed)
    out = []
    for name, seq in ref:
        n = len(seq)
        pos_l, vt_l, vn_l = [], [], []
        for every, vt, vn in ((het_every, 0.5, 0.5), (hom_every, 1.0, 1.0), (som_every, None, 0.0), (loh_every, 1.0, 0.5), (gor_every, 0.5, 1.0)):
            k = max(1, n // every) if n >= every // 4 else 0
            p = rng.integers(0, n, k)
            pos_l.append(p)
            vt_l.append(rng.choice([0.1, 0.25, 0.4], k) if vt is None else np.full(k, vt))
            vn_l.append(np.full(k, vn))
        pos = np.concatenate(pos_l)
        pos, idx = np.unique(pos, return_index=True)
        vt = np.concatenate(vt_l)[idx]
        vn = np.concatenate(vn_l)[idx]
        refb = seq[pos]
        alt = _ACGT[(np.searchsorted(_ACGT, np.where(refb == ord("N"), ord("A"), refb)) + rng.integers(1, 4, len(pos))) & 3]
        out.append(dict(pos=pos, alt=alt, vaf_t=vt, vaf_n=vn))
    return out


class BgzfWriter:
    def __init__(self, path, level=6):
        self.f = open(path, "wb")
        self.buf = bytearray()
        self.level = level

    def write(self, b):
        self.buf += b
        while len(self.buf) >= 0xff00:
            self._block(bytes(self.buf[:0xff00]))
            del self.buf[:0xff00]

    def _block(self, data):
        c = zlib.compressobj(self.level, zlib.DEFLATED, -15)
        comp = c.compress(data) + c.flush()
        bsize = len(comp) + 25
        self.f.write(struct.pack("<4BI2BH2BHH", 31, 139, 8, 4, 0, 0, 255, 6, 66, 67, 2, bsize))
        self.f.write(comp)
        self.f.write(struct.pack("<II", zlib.crc32(data) & 0xffffffff, len(data)))

    def close(self, eof_marker=True):
        if self.buf:
            self._block(bytes(self.buf))
            self.buf = bytearray()
        if eof_marker:
            self._block(b"")
        self.f.close()


def _reg2bin(beg, end):
    end -= 1
    if beg >> 14 == end >> 14:
        return 4681 + (beg >> 14)
    if beg >> 17 == end >> 17:
        return 585 + (beg >> 17)
    if beg >> 20 == end >> 20:
        return 73 + (beg >> 20)
    if beg >> 23 == end >> 23:
        return 9 + (beg >> 23)
    if beg >> 26 == end >> 26:
        return 1 + (beg >> 26)
    return 0


_OPS = {"M": 0, "I": 1, "D": 2, "N": 3, "S": 4, "H": 5, "P": 6, "=": 7, "X": 8}


def bam_record(tid, pos, mapq, flag, cigar, seq_codes, qual, name=b"r", mtid=-1, mpos=-1, isize=0):
    """cigar: list of (op char, len); seq_codes: uint8 nt16 codes; qual: uint8."""
    l = len(seq_codes)
    codes = np.asarray(seq_codes, np.uint8)
    if l & 1:
        codes = np.concatenate([codes, np.zeros(1, np.uint8)])
    packed = ((codes[0::2] << 4) | codes[1::2]).astype(np.uint8).tobytes()
    end = pos + sum(n for op, n in cigar if op in "MDN")
    name0 = name + b"\0"
    core = struct.pack("<iiIIiiii", tid, pos, _reg2bin(pos, max(end, pos + 1)) << 16 | mapq << 8 | len(name0),
                       flag << 16 | len(cigar), l, mtid, mpos, isize)
    cg = b"".join(struct.pack("<I", n << 4 | _OPS[op]) for op, n in cigar)
    body = core + name0 + cg + packed + np.asarray(qual, np.uint8).tobytes()
    return struct.pack("<i", len(body)) + body


def bam_header(contigs, text=None):
    if text is None:
        text = "@HD\tVN:1.0\tSO:coordinate\n" + "".join("@SQ\tSN:%s\tLN:%d\n" % (n, l) for n, l in contigs)
    t = text.encode()
    h = b"BAM\1" + struct.pack("<i", len(t)) + t + struct.pack("<i", len(contigs))
    for n, l in contigs:
        nb = n.encode() + b"\0"
        h += struct.pack("<i", len(nb)) + nb + struct.pack("<i", l)
    return h


def write_bam(path, contigs, records, level=6, eof_marker=True):
    """records: iterable of already encoded records in coordinate order."""
    w = BgzfWriter(path, level)
    w.write(bam_header(contigs))
    for r in records:
        w.write(r)
    w.close(eof_marker)


def simulate_bam(path, ref, variants, sample="tumor", depth=60.0, seed=1, level=6, special=True):
    """Writes a coordinate-sorted BAM of simulated pairs over every contig of `ref`.  Returns the number of records."""
    rng = np.random.default_rng(seed)
    contigs = [(n, len(s)) for n, s in ref]
    w = BgzfWriter(path, level)
    w.write(bam_header(contigs))
    total = 0
    ar = np.arange(READ_LEN)
    for tid, (name, seq) in enumerate(ref):
        L = len(seq)
        if L < 2 * READ_LEN + 200:
            continue
        npairs = int(depth * L / (2 * READ_LEN))
        ins = np.clip(rng.normal(400, 50, npairs).round().astype(np.int64), READ_LEN, None)
        start = rng.integers(0, np.maximum(1, L - ins), npairs)
        ins = np.minimum(ins, L - start)
        ins = np.maximum(ins, READ_LEN)
        start = np.minimum(start, L - ins)
        flip = rng.random(npairs) < 0.5
        pos = np.concatenate([start, start + ins - READ_LEN])
        n = 2 * npairs
        # flags: first-in-pair forward / second reverse (99/147) or flipped (163/83)
        flag = np.concatenate([np.where(flip, 163, 99), np.where(flip, 83, 147)]).astype(np.int64)
        mpos = np.concatenate([start + ins - READ_LEN, start])
        isize = np.concatenate([ins, -ins])
        r = rng.random(npairs)
        mq = np.full(npairs, 60, np.int64)
        m = (r >= 0.85) & (r < 0.93); mq[m] = rng.integers(40, 60, m.sum())
        m = (r >= 0.93) & (r < 0.98); mq[m] = rng.integers(1, 40, m.sum())
        mq[r >= 0.98] = 0
        mq = np.concatenate([mq, mq])
        r = rng.random(n)
        flag[r < 0.01] |= 0x400
        flag[(r >= 0.01) & (r < 0.0125)] |= 0x200
        flag[(r >= 0.0125) & (r < 0.015)] |= 0x100
        order = np.argsort(pos, kind="stable")
        pos, flag, mpos, isize, mq = pos[order], flag[order], mpos[order], isize[order], mq[order]
        # bases: reference, then variants by allele fraction, then sequencing errors
        idx = pos[:, None] + ar[None, :]
        bases = seq[idx]
        v = variants[tid]
        if len(v["pos"]):
            vaf = v["vaf_t"] if sample == "tumor" else v["vaf_n"]
            lo = np.searchsorted(v["pos"], pos, "left")
            hi = np.searchsorted(v["pos"], pos + READ_LEN, "left")
            for k in range(int((hi - lo).max()) if len(lo) else 0):
                sel = np.nonzero(lo + k < hi)[0]
                vi = lo[sel] + k
                take = rng.random(len(sel)) < vaf[vi]
                bases[sel[take], v["pos"][vi[take]] - pos[sel[take]]] = v["alt"][vi[take]]
        qual = rng.choice(np.array([2, 12, 23, 37], np.uint8), (n, READ_LEN), p=[0.02, 0.05, 0.13, 0.80])
        err = rng.random((n, READ_LEN)) < np.power(10.0, -qual.astype(np.float64) / 10.0)
        sub = _ACGT[rng.integers(0, 4, (n, READ_LEN))]
        bases = np.where(err, sub, bases)
        nmask = rng.random((n, READ_LEN)) < 5e-4
        bases[nmask] = ord("N")
        qual[nmask] = 2
        codes = _NT16[bases]
        packed = ((codes[:, 0::2] << 4) | codes[:, 1::2]).astype(np.uint8)
        kind = rng.random(n) if special else np.ones(n)
        name0 = b"r\0"
        for i in range(n):
            p = int(pos[i])
            k = kind[i]
            if k < 0.03 and p + READ_LEN + 8 < L:
                # special CIGARs are rebuilt from the reference so that the bases stay consistent
                q = qual[i]
                if k < 0.0025:
                    il = 1 + int(k * 2000) % 5
                    cigar = [("M", 70), ("I", il), ("M", READ_LEN - 70 - il)]
                    c = np.concatenate([_NT16[seq[p:p + 70]], codes[i, 70:70 + il], _NT16[seq[p + 70:p + READ_LEN - il]]])
                elif k < 0.005:
                    dl = 1 + int(k * 2000) % 5
                    cigar = [("M", 80), ("D", dl), ("M", READ_LEN - 80)]
                    c = np.concatenate([_NT16[seq[p:p + 80]], _NT16[seq[p + 80 + dl:p + dl + READ_LEN]]])
                elif k < 0.007:
                    cigar = [("M", 60), ("N", 300), ("M", READ_LEN - 60)] if p + 300 + READ_LEN < L else [("M", READ_LEN)]
                    c = codes[i] if len(cigar) == 1 else np.concatenate([_NT16[seq[p:p + 60]], _NT16[seq[p + 360:p + 300 + READ_LEN]]])
                elif k < 0.017:
                    cigar = [("S", 10), ("M", READ_LEN - 10)]
                    c = codes[i]
                elif k < 0.027:
                    cigar = [("M", READ_LEN - 12), ("S", 12)]
                    c = codes[i]
                else:
                    cigar = [("H", 5), ("M", READ_LEN), ("H", 3)]
                    c = codes[i]
                rec = bam_record(tid, p, int(mq[i]), int(flag[i]), cigar, c, q, b"r", tid, int(mpos[i]), int(isize[i]))
            else:
                core = struct.pack("<iiIIiiii", tid, p, _reg2bin(p, p + READ_LEN) << 16 | int(mq[i]) << 8 | 2,
                                   int(flag[i]) << 16 | 1, READ_LEN, tid, int(mpos[i]), int(isize[i]))
                body = core + name0 + struct.pack("<I", READ_LEN << 4) + packed[i].tobytes() + qual[i].tobytes()
                rec = struct.pack("<i", len(body)) + body
            w.write(rec)
        total += n
    w.close()
    return total
