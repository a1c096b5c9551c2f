"""CPU tests of the region shards (csrc/host/ssn_bamseek.c, ssn_shard.c): index-free seeking into a BAM against a
brute-force record list, and the union of the shards' column batches against the unmodified reference's single
front-to-back walk (bam_sspileup_file, sniper_pileup.c:226-266, through oracle/_ref) — including the keep rule of :216
(QUIRK B1) across shard edges and reads that reach across them."""
import ctypes as C
import os
import struct
import zlib

import numpy as np
import pytest

from test_packer import _check_pair, _packed_columns, _simple_reads, _write_pair, host, sim  # noqa: F401  (fixtures)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _all_records(path):
    """[(voff, tid, pos, flag, mapq)] of every record, by a plain sequential parse; plus the virtual offset of the end."""
    raw = open(path, "rb").read()
    blocks, at = [], 0          # (coff, inflated bytes)
    while at < len(raw):
        xlen = struct.unpack_from("<H", raw, at + 10)[0]
        bsize = struct.unpack_from("<H", raw, at + 16)[0] + 1
        data = zlib.decompress(raw[at + 12 + xlen:at + bsize - 8], -15)
        if not data:
            break
        blocks.append((at, data))
        at += bsize
    stream = b"".join(d for _, d in blocks)
    starts = np.cumsum([0] + [len(d) for _, d in blocks])

    def voff(u):
        b = int(np.searchsorted(starts, u, "right") - 1)
        if b >= len(blocks):
            return at << 16
        return blocks[b][0] << 16 | (u - int(starts[b]))

    l_text = struct.unpack_from("<i", stream, 4)[0]
    u = 8 + l_text
    n_ref = struct.unpack_from("<i", stream, u)[0]
    u += 4
    for _ in range(n_ref):
        ln = struct.unpack_from("<i", stream, u)[0]
        u += 4 + ln + 4
    recs = []
    while u + 4 <= len(stream):
        bs = struct.unpack_from("<i", stream, u)[0]
        if u + 4 + bs > len(stream):
            break
        tid, pos, x2, x3 = struct.unpack_from("<iiII", stream, u + 4)
        recs.append((voff(u), tid, pos, x3 >> 16, (x2 >> 8) & 0xff))
        u += 4 + bs
    return recs, voff(u)


@pytest.fixture(scope="module")
def seeklib(host):
    lib = host.load_library()

    class Point(C.Structure):
        _fields_ = [("voff", C.c_uint64), ("tid", C.c_int32), ("pos", C.c_int32)]
    lib.ssn_bamfile_open.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
    lib.ssn_bamfile_open.restype = C.c_void_p
    lib.ssn_bamfile_close.argtypes = [C.c_void_p]
    lib.ssn_bamfile_first_record.argtypes = [C.c_void_p]
    lib.ssn_bamfile_first_record.restype = C.c_uint64
    lib.ssn_bamfile_probe.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(Point)]
    lib.ssn_bamfile_lower_bound.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(Point)]
    lib.ssn_bamfile_prev_passing.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    return lib, Point


@pytest.fixture(scope="module")
def big_pair(built, tmp_path_factory):
    from somatic_sniper_b200 import cli_tools, simbam
    d = str(tmp_path_factory.mktemp("shards"))
    contigs = [("chr1", 70000), ("chr2", 26000), ("chrEmpty", 100), ("chr3", 41000), ("chr4", 9000), ("chr5", 17000)]
    pair = cli_tools.make_pair(d, contigs, 24, 14, seed=77, n_runs=2,
                               variant_kw=dict(het_every=500, hom_every=2000, som_every=1000, loh_every=5000, gor_every=5000))
    pair["ref"] = simbam.make_reference(contigs, seed=77, n_runs=2)
    pair["contigs"] = contigs
    return pair


def _key(tid, pos):
    return ((0x7fffffff if tid < 0 else tid) << 32) | max(pos, 0)


def test_seek_against_a_brute_force_record_list(seeklib, big_pair):
    lib, Point = seeklib
    path = big_pair["tumor"]
    recs, end_voff = _all_records(path)
    assert len(recs) > 10000
    err = C.create_string_buffer(256)
    f = lib.ssn_bamfile_open(os.fsencode(path), err, 256)
    assert f, err.value
    try:
        assert lib.ssn_bamfile_first_record(f) == recs[0][0]
        rng = np.random.default_rng(3)
        size = os.path.getsize(path)
        voffs = np.array([r[0] for r in recs], np.uint64)
        keys = np.array([_key(r[1], r[2]) for r in recs], np.uint64)
        block_starts = np.unique(voffs >> np.uint64(16))
        # probe: the first record that starts in a block at or after the offset
        for coff in [int(x) for x in rng.integers(recs[0][0] >> 16, size, 60)] + [int(block_starts[5]), int(block_starts[5]) + 1, size - 1, size - 29]:
            p = Point()
            rc = lib.ssn_bamfile_probe(f, coff, C.byref(p))
            later = block_starts[block_starts >= coff]
            if len(later) == 0:
                assert rc == 0
                continue
            i = int(np.searchsorted(voffs, np.uint64(int(later[0]) << 16), "left"))
            assert rc == 1 and (p.voff, p.tid, p.pos) == recs[i][:3], coff
        # lower_bound: the first record at or behind (tid, pos)
        targets = [(0, 0), (0, 1), (5, 16999), (5, 17000), (6, 0), (2, 0), (2, 50), (3, 0), (1, 25999)]
        targets += [(int(t), int(p)) for t, p in zip(rng.integers(0, 6, 80), rng.integers(0, 70000, 80))]
        for tid, pos in targets:
            p = Point()
            rc = lib.ssn_bamfile_lower_bound(f, tid, pos, C.byref(p))
            i = int(np.searchsorted(keys, np.uint64(_key(tid, pos)), "left"))
            if i == len(recs):
                assert rc == 0
            else:
                assert rc == 1 and (p.voff, p.tid, p.pos) == recs[i][:3], (tid, pos)
        # prev_passing: the last record in front of a position that passes the read filter
        for mask, thr in ((0x704, 0), (0x704, 30)):
            passing = np.array([(r[3] & mask) == 0 and r[4] >= thr for r in recs])
            for i in [0, 1, 2, len(recs) - 1] + [int(x) for x in rng.integers(0, len(recs), 60)]:
                t, q = C.c_int32(-7), C.c_int32(-7)
                rc = lib.ssn_bamfile_prev_passing(f, recs[i][0], mask, thr, C.byref(t), C.byref(q))
                before = np.nonzero(passing[:i])[0]
                if len(before) == 0:
                    assert rc == 0
                else:
                    assert rc == 1 and (t.value, q.value) == recs[before[-1]][1:3], i
            t, q = C.c_int32(), C.c_int32()
            assert lib.ssn_bamfile_prev_passing(f, end_voff, mask, thr, C.byref(t), C.byref(q)) == 1
            assert (t.value, q.value) == recs[int(np.nonzero(passing)[0][-1])][1:3]
    finally:
        lib.ssn_bamfile_close(f)


@pytest.mark.parametrize("stride,lookback,mapq,threads", [(1, 1, 0, 3), (2, 1, 20, 0), (3, 2, 1, 6), (0, 0, 0, 4)])
def test_shards_reproduce_the_reference_walk(host, sim, big_pair, stride, lookback, mapq, threads):
    """The union of the shards' batches, in shard order, equals the unmodified reference's one walk over the pair:
    intervals of 1-3 windows (so that edges fall inside targets, at target starts and in empty targets), the automatic
    cut, reads that reach across edges (all within the look-back), and QUIRK B1 at every target start."""
    from somatic_sniper_b200 import api
    p = big_pair
    ss = host.ShardSet(p["tumor"], p["normal"], p["fasta"], mapq, threads=threads, stride_windows=stride, lookback_windows=lookback,
                       max_columns=5000, max_entries=300000)
    n = ss.n_shards()
    rg = ss.ranges()
    assert rg[0][:2] == (0, 0) and rg[-1][2] == 0x7fffffff
    assert all(rg[i][2:] == rg[i + 1][:2] for i in range(n - 1)) and all(r[1] % 8192 == 0 for r in rg)
    if stride:
        assert n >= sum(l for _, l in p["contigs"]) // (stride * 8192)
    batches = list(ss)
    assert not ss.long_span()
    st = ss.stats()
    ss.close()
    got = _packed_columns(batches)
    # against the single front-to-back packer ...
    pk = host.Packer(p["tumor"], p["normal"], p["fasta"], mapq, threads=2)
    serial = _packed_columns(list(pk))
    pk.close()
    assert got == serial
    assert st.columns == len(got)
    # ... which test_packer pins against the reference walker; do it here directly as well
    seqs = [s for _, s in p["ref"]]
    from test_packer import _expected_columns, _reference_walk
    rc, re_ = _reference_walk(p["tumor"], p["normal"], mapq)
    want = _expected_columns(rc, re_, lambda t, q: seqs[t][q] if t < len(seqs) and q < len(seqs[t]) else ord("N"), api.nt16_lut())
    assert got == want and len(got) > 100000


def test_read_route_over_shards_equals_the_serial_read_route(host, big_pair):
    p = big_pair
    ss = host.ShardSet(p["tumor"], p["normal"], p["fasta"], 0, threads=3, stride_windows=2, lookback_windows=1)
    got = []
    for _, pk in ss.shards():
        for rb in pk.read_batches(max_windows=3):
            got.append(host.pack_read_windows(rb))
    ss.close()
    pk = host.Packer(p["tumor"], p["normal"], p["fasta"], 0, threads=2)
    want = [host.pack_read_windows(rb) for rb in pk.read_batches(max_windows=3)]
    pk.close()
    cat = lambda bs, k: np.concatenate([b[k] for b in bs])
    for k in ("bs", "bq", "mq", "ref_nt16"):
        assert np.array_equal(cat(got, k), cat(want, k)), k
    assert sum(b["ncol"] for b in got) == sum(b["ncol"] for b in want) > 100000


@pytest.mark.parametrize("last_start,kept", [(9, True), (10, False)])
def test_keep_rule_state_crosses_a_shard_edge(host, sim, tmp_path, last_start, kept):
    """QUIRK B1 with the target change exactly on a shard edge: chr1 is one window long (8192), so with one-window
    intervals the shard that starts chr2 must learn chr1's last passing read start from the backward scan.  Filtered
    reads (duplicate, mapQ below the threshold) between the two must not count."""
    contigs = [("chr1", 8192), ("chr2", 20000)]
    ref = [("chr1", np.frombuffer(b"ACGT" * 2048, np.uint8)), ("chr2", np.frombuffer(b"TGCA" * 5000, np.uint8))]
    mk = lambda: (_simple_reads(0, [0, 5, last_start]) + _simple_reads(0, [40, 41], flag=0x400) + _simple_reads(0, [50], mq=3)
                  + _simple_reads(1, [0], length=10) + _simple_reads(1, [0, 0, 4], length=10)
                  + _simple_reads(1, [8190, 8200, 16380, 16390], length=30))
    t, n, fa = _write_pair(sim, str(tmp_path), contigs, ref, mk(), mk())
    ss = host.ShardSet(t, n, fa, 10, threads=2, stride_windows=1, lookback_windows=1)
    assert ss.n_shards() == 4
    got = _packed_columns(list(ss))
    assert not ss.long_span()
    ss.close()
    want = _check_pair(host, sim, t, n, fa, ref, mapq=10)          # serial packer == reference walker
    assert got == want
    assert [len(g[3]) for g in got if g[0] == 1 and g[1] == 0] == [3 if kept else 2]
    assert [len(g[3]) for g in got if g[0] == 1 and g[1] in (8192, 8205)] == [1, 2]   # the reads across chr2's window edge


def test_long_read_across_an_edge_raises_the_flag(host, sim, tmp_path):
    """A read that begins more than the look-back in front of a shard edge and reaches behind it (a 20 kb N skip) is
    invisible to the next shard: the set reports it, and the caller's second pass as one interval is exact."""
    contigs = [("chr1", 60000)]
    ref = [("chr1", np.frombuffer(b"ACGT" * 15000, np.uint8))]
    long_read = [(0, 100, 60, 0, "10M20000N10M", "A" * 20, [30] * 20)]
    mk = lambda: _simple_reads(0, [50, 90]) + long_read + _simple_reads(0, [200, 8300, 20100, 20105, 20110, 30000], length=30)
    t, n, fa = _write_pair(sim, str(tmp_path), contigs, ref, mk(), mk())
    ss = host.ShardSet(t, n, fa, 0, threads=2, stride_windows=1, lookback_windows=1)
    sharded = _packed_columns(list(ss))
    assert ss.long_span()
    ss.close()
    ss = host.ShardSet(t, n, fa, 0, threads=2, serial=1)
    assert ss.n_shards() == 1
    serial = _packed_columns(list(ss))
    assert not ss.long_span()
    ss.close()
    want = _check_pair(host, sim, t, n, fa, ref)
    assert serial == want and sharded != want
    # with a look-back that covers the read, the shards see it
    ss = host.ShardSet(t, n, fa, 0, threads=2, stride_windows=1, lookback_windows=3)
    assert _packed_columns(list(ss)) == want and not ss.long_span()
    ss.close()


def test_many_small_targets_share_the_sequence_cache(host, sim, tmp_path):
    """More targets inside one group of windows than the sequence cache keeps idle (the old 4-slot cache freed
    a sequence a pending window still pointed at)."""
    from somatic_sniper_b200 import cli_tools
    contigs = [("c%d" % i, 2000 + 100 * i) for i in range(9)]
    pair = cli_tools.make_pair(str(tmp_path), contigs, 12, 8, seed=31, n_runs=1)
    ref = sim.make_reference(contigs, seed=31, n_runs=1)
    got = _check_pair(host, sim, pair["tumor"], pair["normal"], pair["fasta"], ref, threads=3)
    assert len({g[0] for g in got}) == 9
    pk = host.Packer(pair["tumor"], pair["normal"], pair["fasta"], 0, threads=2)
    n = sum(host.pack_read_windows(rb)["ncol"] for rb in pk.read_batches(max_windows=16))
    pk.close()
    assert n == len(got)


def test_mixed_routes_windows_too_deep_for_a_read_batch(host, big_pair):
    """The CLI's fall-back: ssn_packer_next_reads refuses a window that alone exceeds the read batch (SSN_ERR_OVERFLOW,
    nothing consumed), ssn_packer_next packs exactly that window into columns (one-window mode), then the reads route
    resumes.  With a capacity between the shallow and the deep windows of the pair both routes alternate; the union of the
    columns equals the column route's."""
    from somatic_sniper_b200 import api
    p = big_pair
    lib = host.load_library()
    pk = host.Packer(p["tumor"], p["normal"], p["fasta"], 0, threads=2)
    want = _packed_columns(list(pk))
    pk.close()
    cap_entries = 8192 * 38 * 150 // 150          # windows at full depth (24x + 14x) exceed it; ragged ones at contig ends fit
    pk = host.Packer(p["tumor"], p["normal"], p["fasta"], 0, threads=2, max_columns=1 << 14, max_entries=cap_entries)
    got, n_reads_batches, n_col_batches = [], 0, 0
    it_reads = pk.read_batches(max_rec_bytes=1 << 24, max_reads=1 << 18, max_windows=2)
    cols_mode = False
    while True:
        if not cols_mode:
            try:
                rb = next(it_reads)
            except StopIteration:
                break
            except RuntimeError as e:
                assert "capacity" in str(e)
                cols_mode = True
                lib.ssn_packer_one_window(pk.h, 1)
                it_reads = pk.read_batches(max_rec_bytes=1 << 24, max_reads=1 << 18, max_windows=2)
                continue
            b = host.pack_read_windows(rb)
            # positions come back as window << 13 | offset: rebuild tid / pos / ref_char for the comparison
            w, o = b["pos"] >> 13, b["pos"] & 8191
            b["tid"] = rb["win"]["tid"][w]; b["pos"] = rb["win"]["lo"][w] + o; b["ref_char"] = rb["ref_char"][w, o]
            got += _packed_columns([b]); n_reads_batches += 1
            continue
        b = next(iter(pk), None)
        if b is None:
            break
        got += _packed_columns([b]); n_col_batches += 1
        if not lib.ssn_packer_partial(pk.h):
            cols_mode = False
            lib.ssn_packer_one_window(pk.h, 0)
    pk.close()
    assert n_reads_batches >= 2 and n_col_batches >= 10
    assert got == want


def test_several_consumers_take_distinct_shards(host, big_pair):
    """One feeder per GPU takes shards from the same set.  Callers that arrive while an earlier one still waits for its
    shard's readers must get the next shard, not the same one (a claim race handed one packer to two feeders once: they
    then freed each other's windows).  Four threads, tiny intervals so that callers outrun the readers; the union of what
    they pack equals the single consumer's, and every shard is handed out exactly once."""
    import threading
    p = big_pair
    def run(n_consumers):
        ss = host.ShardSet(p["tumor"], p["normal"], p["fasta"], 0, threads=6, stride_windows=1, lookback_windows=1, max_reading=2, max_open=3)
        got, seen, lock, errs = {}, [], threading.Lock(), []
        def work():
            try:
                for idx, pk in ss.shards():
                    cols = 0
                    sig = 0
                    for rb in pk.read_batches(max_windows=2):
                        b = host.pack_read_windows(rb)
                        cols += b["ncol"]
                        sig += int(b["bs"].astype(np.int64).sum()) + 3 * int(b["bq"].astype(np.int64).sum())
                    with lock:
                        seen.append(idx)
                        got[idx] = (cols, sig)
            except Exception as e:          # noqa: BLE001
                errs.append(e)
        th = [threading.Thread(target=work) for _ in range(n_consumers)]
        [t.start() for t in th]
        [t.join() for t in th]
        n = ss.n_shards()
        assert not ss.long_span()
        ss.close()
        assert not errs, errs
        assert sorted(seen) == list(range(n))
        return got
    one, four = run(1), run(4)
    assert one == four and sum(c for c, _ in one.values()) > 100000


@pytest.mark.parametrize("direct", [0, 1, 2])
def test_readers_through_an_inflate_service(host, big_pair, direct):
    """The readers' GPU-inflate plumbing without a GPU: a service whose `run` decodes every block with the DEVICE decoder's
    source compiled for the host (ssn_inflate_dev_host) and refuses every seventh block (the reader must then inflate that
    one itself).  Lanes are handed out and given back; the columns equal the host-inflate run's.  direct = 1: the service
    also offers host_alloc / host_free / run_to, so the readers own their output buffers and the service writes there.
    direct = 2: lanes whose input buffer holds 70 000 bytes, so a batch ends where the next payload would not fit."""
    import threading
    from somatic_sniper_b200 import api
    p = big_pair
    dev = api.load_library()
    dev.ssn_inflate_dev_host_shift.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32]
    dev.ssn_inflate_dev_host_shift.restype = C.c_int
    lib = host.load_library()

    class Block(C.Structure):
        _fields_ = [("in_off", C.c_uint32), ("in_len", C.c_uint32), ("out_off", C.c_uint32), ("out_len", C.c_uint32)]

    ACQ = C.CFUNCTYPE(C.c_int32, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32))
    REL = C.CFUNCTYPE(None, C.c_void_p, C.c_int32)
    RUN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.POINTER(Block), C.c_uint32, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint8))

    ALLOC = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_uint64)
    FREE = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p)
    RUNTO = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.POINTER(Block), C.c_uint32, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint8), C.c_void_p)

    class Service(C.Structure):
        _fields_ = [("ctx", C.c_void_p), ("lane_acquire", ACQ), ("lane_release", REL), ("run", RUN),
                    ("host_alloc", ALLOC), ("host_free", FREE), ("run_to", RUNTO)]

    MAXB, lock = 24, threading.Lock()
    lanes = [dict(busy=False, i=np.zeros(MAXB * 65600 + 64, np.uint8), o=np.zeros(MAXB * 65536 + 64, np.uint8)) for _ in range(8)]
    stats = dict(acquired=0, released=0, blocks=0, refused=0, none_free=0)

    def acquire(ctx, pin, pin_cap, pout, pout_cap, pmax):
        with lock:
            for k, L in enumerate(lanes):
                if not L["busy"]:
                    L["busy"] = True
                    stats["acquired"] += 1
                    pin[0] = L["i"].ctypes.data; pin_cap[0] = 70000 if direct == 2 else MAXB * 65600
                    pout[0] = L["o"].ctypes.data; pout_cap[0] = MAXB * 65536
                    pmax[0] = MAXB
                    return k
            stats["none_free"] += 1
            return -1

    def release(ctx, k):
        with lock:
            lanes[k]["busy"] = False
            stats["released"] += 1

    owned = {}

    def host_alloc(ctx, nbytes):
        a = np.zeros(nbytes + 64, np.uint8)
        with lock:
            owned[a.ctypes.data] = a
            stats["allocs"] = stats.get("allocs", 0) + 1
        return a.ctypes.data

    def host_free(ctx, ptr):
        with lock:
            del owned[ptr]

    def run(ctx, k, blk, n, in_bytes, out_bytes, refused, dst=None):
        L = lanes[k]
        if direct == 2:
            assert in_bytes <= 70000
            with lock:
                stats["runs"] = stats.get("runs", 0) + 1
        out_base = dst if dst else L["o"].ctypes.data
        if dst:
            assert dst in owned
            with lock:
                stats["direct_runs"] = stats.get("direct_runs", 0) + 1
        for j in range(n):
            b = blk[j]
            with lock:
                stats["blocks"] += 1
                nth = stats["blocks"]
            if nth % 7 == 0:
                refused[j] = 1
                with lock:
                    stats["refused"] += 1
                continue
            rc = dev.ssn_inflate_dev_host_shift(L["i"].ctypes.data + b.in_off, b.in_len, out_base + b.out_off, b.out_len, 0)
            refused[j] = 1 if rc else 0
        return 0

    if direct == 1:
        svc = Service(None, ACQ(acquire), REL(release), RUN(run), ALLOC(host_alloc), FREE(host_free), RUNTO(run))
    else:
        svc = Service(None, ACQ(acquire), REL(release), RUN(run), ALLOC(), FREE(), RUNTO())
    lib.ssn_shardset_set_inflate_service.argtypes = [C.c_void_p, C.POINTER(Service)]
    lib.ssn_shardset_set_inflate_service.restype = None
    ss = host.ShardSet(p["tumor"], p["normal"], p["fasta"], 0, threads=4, stride_windows=2, lookback_windows=1, max_reading=3, max_open=4,
                       max_columns=5000, max_entries=300000)
    lib.ssn_shardset_set_inflate_service(ss.h, C.byref(svc))
    got = _packed_columns(list(ss))
    assert not ss.long_span()
    ss.close()
    pk = host.Packer(p["tumor"], p["normal"], p["fasta"], 0, threads=2)
    want = _packed_columns(list(pk))
    pk.close()
    assert got == want
    assert stats["blocks"] > 200 and stats["refused"] > 20 and stats["acquired"] == stats["released"] > 10
    if direct == 2:
        assert stats["runs"] * 8 > stats["blocks"]           # a handful of blocks per run, not 24
    if direct == 1:
        assert stats["allocs"] >= 2 and stats["direct_runs"] > 10 and not owned      # every buffer came back
