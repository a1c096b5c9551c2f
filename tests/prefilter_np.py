"""numpy restatement of the hom-reference pre-filter's decision (somatic_sniper_b200/csrc/ssn_prefilter.cuh), evaluated with
the bound tables the product's host code builds (ssn_prefilter_tables.c).  Test infrastructure: lets the CPU suite check
the filter's proof obligations against the oracle without a GPU, and gives the GPU tests an independent expectation of
how many columns the kernel should drop."""
import numpy as np

from somatic_sniper_b200 import api

RAW_BYTES = 3200      # kRawBytes: staged read-bases per array of one 32-column tile


def decide(cols, tables=None, tile_rules=True):
    """Returns bool[ncol]: True where the pre-filter drops the column (both samples provably call hom-reference).
    tile_rules: also apply the kernel's per-tile conditions (clean base qualities, depth and staging limits)."""
    t = tables or api.host_prefilter_tables()
    ab, h, F = t["ab"].astype(np.float64), t["h"].astype(np.float64), t["F"].astype(np.float64)
    ncol = int(cols["ncol"])
    if not t["enabled"] or ncol == 0:
        return np.zeros(ncol, bool)
    off = cols["off"].astype(np.int64)
    bs, bq, mq = cols["bs"], cols["bq"].astype(np.int64), cols["mq"].astype(np.int64)
    ref16 = api.nt16_lut()[cols["ref_char"]].astype(np.int64)
    n_all = np.diff(off)
    seg = np.repeat(np.arange(2 * ncol), n_all)
    col = seg >> 1
    live = bq != 0                      # quality-0 read-bases contribute nothing to the model (sniper_maqcns.c:165-167)
    n = np.bincount(seg, live, 2 * ncol).astype(np.int64)
    nonref = ((bs & 15) != ref16[col]) & live
    rev = ((bs >> 4) & 1) == 1
    q = np.minimum(bq, mq & 0x7f)
    cnt = lambda w: np.bincount(seg, w, 2 * ncol)
    m = cnt(nonref).astype(np.int64)
    S = cnt(nonref * bq)
    n1 = cnt(~nonref & rev & live).astype(np.int64)
    q1 = cnt((~nonref & rev & live) * q)
    q0 = cnt((~nonref & ~rev & live) * q)
    sample_ok = (n_all <= 255) & (sample_margin(t, n, m, S, n1, q1, q0) >= 0.0)
    nT, nN = n_all[0::2], n_all[1::2]
    wanted = (nT > 0) & (nN > 0) & (ref16 != 15)
    drop = wanted & np.isin(ref16, [1, 2, 4, 8]) & sample_ok[0::2] & sample_ok[1::2]
    if tile_rules:
        bad_bq = np.bincount(col, bq > 63, ncol) > 0
        ntile = (ncol + 31) // 32
        tile = np.arange(ncol) >> 5
        pad = ntile * 32 - ncol
        tile_bad = np.bincount(tile, bad_bq, ntile) > 0
        # (tiles deeper than the staging area are staged in several passes of consecutive columns; the clean check
        # is then per pass, i.e. the kernel may drop columns of a tile this restatement gives up on)
        # slack bytes of the neighbouring tiles are scanned too (conservative in the kernel); ignore that here
        drop &= ~tile_bad[tile]
    return drop


def sample_margin(t, n, m, S, n1, q1, q0):
    """How far one sample is from not being settled: min(t1 - SU, t1 + cmin_all - U, h + g2_run - U) with the quantities of
    ssn_prefilter.cuh's header (fp64 here, fp32 in the kernel); the filter needs t1 > SU and the other two >= 0, so a
    sample is settled iff the value is >= 0 and t1 > SU (a margin of exactly 0 in the first term does not settle: it is
    returned as -1e-9).  n: live read-bases (<= 255), m: non-reference ones, S: their base qualities summed, n1 / q1:
    reference read-bases on the reverse strand and their q summed, q0: the forward strand's q sum."""
    ab, h, F = t["ab"].astype(np.float64), t["h"].astype(np.float64), t["F"].astype(np.float64)
    vt = t["vt"].astype(np.float64)
    n, m, n1 = (np.asarray(x).astype(np.int64) for x in (n, m, n1))
    n0 = n - m - n1
    nn, mm = np.minimum(n, 255), np.minimum(m, 255)
    t1 = F[np.minimum(n1, 256)] * q1 / np.maximum(n1 + m, 1) + F[np.minimum(n0, 256)] * q0 / np.maximum(n0 + m, 1)
    SU = (S + 3 * m) * 1.0001
    U = np.maximum(SU + ab[nn, mm, 1], 0.0) + 1.1
    tight = (m >= 1) & (m <= 8) & (SU <= 255.0)
    U = np.where(tight, np.minimum(U, vt[nn, np.clip(m, 1, 8) - 1, np.minimum(np.ceil(SU), 255).astype(np.int64)] + 1.1), U)
    hs = h[np.minimum(n - m, 255), mm]
    first = np.where(t1 > SU, t1 - SU, -1e-9)
    return np.minimum(first, np.minimum(t1 + ab[nn, mm, 0] - U, hs + ab[nn, mm, 2] - U))


def consensus_genotypes(glf):
    """(tumor, normal) best genotype masks from the oracle's / the kernel's 28-word column dump."""
    return glf[:, 13] >> 28, glf[:, 27] >> 28
