"""ctypes mirror of include/ssn_host.h (libssn_host.so: the BAM -> SoA column packer and the line writers).
Holds no logic of its own; tests and tools drive the native code through it."""
import ctypes as C
import os

import numpy as np

from . import api

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
CLI = os.path.join(_HERE, "bin", "bam-somaticsniper")


class ColMap(C.Structure):
    _fields_ = [("tid", C.POINTER(C.c_int32)), ("pos", C.POINTER(C.c_uint32)), ("ref_char", C.POINTER(C.c_uint8))]


class PackerStats(C.Structure):
    _fields_ = [("records", C.c_uint64 * 2), ("columns", C.c_uint64), ("entries", C.c_uint64), ("batches", C.c_uint64),
                ("t_wait", C.c_double), ("t_count", C.c_double), ("t_fill", C.c_double)]


class ShardOpts(C.Structure):
    _fields_ = [("n_threads", C.c_int), ("stride_windows", C.c_uint32), ("lookback_windows", C.c_uint32), ("max_reading", C.c_int),
                ("max_open", C.c_int), ("serial", C.c_int)]


HOST_SYMBOLS = ["ssn_shardset_open", "ssn_shardset_close", "ssn_shardset_n_shards", "ssn_shardset_n_targets", "ssn_shardset_target_name",
                "ssn_shardset_range", "ssn_shardset_next", "ssn_shardset_done", "ssn_shardset_long_span", "ssn_shardset_get_stats", "ssn_shardset_set_inflate_service",
                "ssn_packer_one_window", "ssn_packer_partial", "ssn_packer_open", "ssn_packer_close", "ssn_packer_n_targets", "ssn_packer_target_name", "ssn_packer_next",
                "ssn_packer_get_stats", "ssn_pack_read_windows", "ssn_packer_next_reads", "ssn_host_last_error", "ssn_writer_create", "ssn_writer_header", "ssn_writer_site",
                "ssn_writer_destroy", "ssn_writer_n_formats", "ssn_writer_format_name"]


def load_library():
    global _LIB
    if _LIB is not None:
        return _LIB
    path = os.environ.get("SSN_HOST_LIB") or os.path.join(_HERE, "lib", "libssn_host.so")   # override: sanitizer builds (tools/sanitize_host.sh)
    if not os.path.exists(path):
        raise RuntimeError("libssn_host.so is not built (run python __graft_entry__.py)")
    lib = C.CDLL(path)
    lib.ssn_packer_open.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    lib.ssn_packer_open.restype = C.c_int
    lib.ssn_packer_close.argtypes = [C.c_void_p]
    lib.ssn_packer_close.restype = None
    lib.ssn_packer_n_targets.argtypes = [C.c_void_p, C.c_int]
    lib.ssn_packer_n_targets.restype = C.c_int32
    lib.ssn_packer_target_name.argtypes = [C.c_void_p, C.c_int, C.c_int32]
    lib.ssn_packer_target_name.restype = C.c_char_p
    lib.ssn_packer_next.argtypes = [C.c_void_p, C.POINTER(api.BatchView), C.POINTER(ColMap), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]
    lib.ssn_packer_next.restype = C.c_int
    lib.ssn_packer_get_stats.argtypes = [C.c_void_p, C.POINTER(PackerStats)]
    lib.ssn_packer_get_stats.restype = None
    lib.ssn_host_last_error.restype = C.c_char_p
    lib.ssn_packer_next_reads.argtypes = [C.c_void_p, C.POINTER(api.ReadsView), C.c_uint32, C.c_uint64, C.POINTER(C.c_uint8),
                                          C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]
    lib.ssn_packer_next_reads.restype = C.c_int
    lib.ssn_pack_read_windows.argtypes = [C.POINTER(api.ReadsView), C.c_uint32, C.POINTER(api.BatchView), C.POINTER(C.c_uint32),
                                          C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]
    lib.ssn_pack_read_windows.restype = C.c_int
    lib.ssn_shardset_open.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.POINTER(ShardOpts), C.POINTER(C.c_void_p)]
    lib.ssn_shardset_open.restype = C.c_int
    lib.ssn_shardset_close.argtypes = [C.c_void_p]
    lib.ssn_shardset_close.restype = None
    lib.ssn_shardset_n_shards.argtypes = [C.c_void_p]
    lib.ssn_shardset_n_shards.restype = C.c_int32
    lib.ssn_shardset_n_targets.argtypes = [C.c_void_p, C.c_int]
    lib.ssn_shardset_n_targets.restype = C.c_int32
    lib.ssn_shardset_target_name.argtypes = [C.c_void_p, C.c_int, C.c_int32]
    lib.ssn_shardset_target_name.restype = C.c_char_p
    lib.ssn_shardset_range.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_uint32), C.POINTER(C.c_int32), C.POINTER(C.c_uint32)]
    lib.ssn_shardset_range.restype = C.c_int
    lib.ssn_shardset_next.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_void_p)]
    lib.ssn_shardset_next.restype = C.c_int
    lib.ssn_shardset_done.argtypes = [C.c_void_p, C.c_int32]
    lib.ssn_shardset_done.restype = C.c_int
    lib.ssn_shardset_long_span.argtypes = [C.c_void_p]
    lib.ssn_shardset_long_span.restype = C.c_int
    lib.ssn_shardset_get_stats.argtypes = [C.c_void_p, C.POINTER(PackerStats)]
    lib.ssn_shardset_get_stats.restype = None
    lib.ssn_packer_one_window.argtypes = [C.c_void_p, C.c_int]
    lib.ssn_packer_one_window.restype = None
    lib.ssn_packer_partial.argtypes = [C.c_void_p]
    lib.ssn_packer_partial.restype = C.c_int
    lib.ssn_writer_n_formats.restype = C.c_int
    lib.ssn_writer_format_name.argtypes = [C.c_int]
    lib.ssn_writer_format_name.restype = C.c_char_p
    _LIB = lib
    return lib


class Packer:
    """Iterates the SoA batches of a tumor/normal BAM pair (host arrays owned by numpy; no GPU involved)."""

    def __init__(self, tumor, normal, fasta, min_mapq=0, threads=2, max_columns=1 << 16, max_entries=1 << 23, handle=None):
        self.lib = load_library()
        self.max_columns, self.max_entries = max_columns, max_entries
        self.h = C.c_void_p()
        self.owned = handle is None
        if handle is not None:          # a region shard's packer (owned by its ShardSet)
            self.h = handle
            return
        st = self.lib.ssn_packer_open(os.fsencode(tumor), os.fsencode(normal), os.fsencode(fasta) if fasta else None,
                                      int(min_mapq), int(threads), C.byref(self.h))
        if st != 0:
            raise RuntimeError("ssn_packer_open: %s" % self.lib.ssn_host_last_error().decode())
        self.max_columns, self.max_entries = max_columns, max_entries

    def close(self):
        if self.h and self.owned:
            self.lib.ssn_packer_close(self.h)
        self.h = C.c_void_p()

    def target_names(self, which=0):
        n = self.lib.ssn_packer_n_targets(self.h, which)
        return [self.lib.ssn_packer_target_name(self.h, which, i).decode() for i in range(n)]

    def stats(self):
        s = PackerStats()
        self.lib.ssn_packer_get_stats(self.h, C.byref(s))
        return s

    def __iter__(self):
        mc, me = self.max_columns, self.max_entries
        while True:
            off = np.zeros(2 * mc + 1, np.uint32)
            bs = np.zeros(me + 64, np.uint8); bq = np.zeros(me + 64, np.uint8); mq = np.zeros(me + 64, np.uint8)
            ref16 = np.zeros(mc + 64, np.uint8)
            tid = np.zeros(mc, np.int32); pos = np.zeros(mc, np.uint32); refc = np.zeros(mc, np.uint8)
            view = api.BatchView()
            view.off = off.ctypes.data_as(C.POINTER(C.c_uint32))
            view.bs = bs.ctypes.data_as(C.POINTER(C.c_uint8)); view.bq = bq.ctypes.data_as(C.POINTER(C.c_uint8))
            view.mq = mq.ctypes.data_as(C.POINTER(C.c_uint8)); view.ref_nt16 = ref16.ctypes.data_as(C.POINTER(C.c_uint8))
            view.max_columns = mc; view.max_entries = me
            cm = ColMap(tid.ctypes.data_as(C.POINTER(C.c_int32)), pos.ctypes.data_as(C.POINTER(C.c_uint32)),
                        refc.ctypes.data_as(C.POINTER(C.c_uint8)))
            ncol, nent = C.c_uint32(0), C.c_uint64(0)
            st = self.lib.ssn_packer_next(self.h, C.byref(view), C.byref(cm), C.byref(ncol), C.byref(nent))
            if st != 0:
                raise RuntimeError("ssn_packer_next: %s" % self.lib.ssn_host_last_error().decode())
            n, e = ncol.value, nent.value
            if n == 0:
                return
            yield dict(off=off[:2 * n + 1].copy(), bs=bs[:e].copy(), bq=bq[:e].copy(), mq=mq[:e].copy(), ref_nt16=ref16[:n].copy(),
                       ref_char=refc[:n].copy(), tid=tid[:n].copy(), pos=pos[:n].copy(), ncol=n, n_entries=e)


def _packer_read_batches(self, max_rec_bytes=1 << 24, max_reads=1 << 18, max_windows=4):
    """Iterates read batches (ssn_packer_next_reads): dicts like reads.build_read_windows' plus ref_char[w, 8192], tid, lo."""
    nw = max_windows
    while True:
        recs = np.zeros(max_rec_bytes, np.uint8)
        rec_off = np.zeros(max_reads, np.uint32)
        win = np.zeros(nw, api.READ_WINDOW_DTYPE)
        wref = np.zeros((nw, api.READ_WINDOW), np.uint8)
        refc = np.zeros((nw, api.READ_WINDOW), np.uint8)
        rv = api.ReadsView()
        rv.recs = recs.ctypes.data_as(C.POINTER(C.c_uint8))
        rv.rec_off = rec_off.ctypes.data_as(C.POINTER(C.c_uint32))
        rv.win = win.ctypes.data
        rv.ref_nt16 = wref.ctypes.data_as(C.POINTER(C.c_uint8))
        rv.max_rec_bytes, rv.max_reads, rv.max_windows = max_rec_bytes, max_reads, nw
        n_w, n_r, n_b = C.c_uint32(), C.c_uint32(), C.c_uint64()
        st = self.lib.ssn_packer_next_reads(self.h, C.byref(rv), self.max_columns, self.max_entries,
                                            refc.ctypes.data_as(C.POINTER(C.c_uint8)), C.byref(n_w), C.byref(n_r), C.byref(n_b))
        if st != 0:
            raise RuntimeError("ssn_packer_next_reads: %s" % self.lib.ssn_host_last_error().decode())
        k = int(n_w.value)
        if k == 0:
            return
        yield dict(recs=recs[:int(n_b.value)].copy(), rec_off=rec_off[:int(n_r.value)].copy(), win=win[:k].copy(),
                   ref_nt16=wref[:k].copy(), ref_char=refc[:k].copy(), entries_bound=self.max_entries)


Packer.read_batches = _packer_read_batches


def pack_read_windows(rw, max_columns=None, max_entries=None):
    """Whole read windows (reads.build_read_windows) through the host packer's count/fill code (ssn_pack_read_windows).
    Returns a column dict: ncol, n_entries, off, bs, bq, mq, ref_nt16, pos (window << 13 | offset)."""
    lib = load_library()
    nw = int(rw["win"].size)
    max_columns = max_columns or max(1, nw * api.READ_WINDOW)
    max_entries = max_entries or max(64, int(rw["entries_bound"]))
    recs = np.ascontiguousarray(rw["recs"], np.uint8)
    rec_off = np.ascontiguousarray(rw["rec_off"], np.uint32)
    win = np.ascontiguousarray(rw["win"])
    wref = np.ascontiguousarray(rw["ref_nt16"], np.uint8)
    rv = api.ReadsView()
    rv.recs = recs.ctypes.data_as(C.POINTER(C.c_uint8))
    rv.rec_off = rec_off.ctypes.data_as(C.POINTER(C.c_uint32))
    rv.win = win.ctypes.data
    rv.ref_nt16 = wref.ctypes.data_as(C.POINTER(C.c_uint8))
    rv.max_rec_bytes, rv.max_reads, rv.max_windows = recs.size, rec_off.size, nw
    off = np.zeros(2 * max_columns + 1, np.uint32)
    bs, bq, mq = (np.zeros(max_entries + 64, np.uint8) for _ in range(3))
    ref = np.zeros(max_columns + 64, np.uint8)
    pos = np.zeros(max_columns, np.uint32)
    bv = api.BatchView()
    bv.off = off.ctypes.data_as(C.POINTER(C.c_uint32))
    bv.bs, bv.bq, bv.mq, bv.ref_nt16 = (a.ctypes.data_as(C.POINTER(C.c_uint8)) for a in (bs, bq, mq, ref))
    bv.max_columns, bv.max_entries = max_columns, max_entries
    ncol, nent = C.c_uint32(), C.c_uint64()
    st = lib.ssn_pack_read_windows(C.byref(rv), nw, C.byref(bv), pos.ctypes.data_as(C.POINTER(C.c_uint32)), C.byref(ncol), C.byref(nent))
    if st != 0:
        raise RuntimeError("ssn_pack_read_windows: %s" % lib.ssn_host_last_error().decode())
    n, e = int(ncol.value), int(nent.value)
    return dict(ncol=n, n_entries=e, off=off[:2 * n + 1].copy(), bs=bs[:e].copy(), bq=bq[:e].copy(), mq=mq[:e].copy(),
                ref_nt16=ref[:n].copy(), pos=pos[:n].copy())


class ShardSet:
    """Region shards of a pair (ssn_shardset): iterating yields, shard after shard in genome order, what a Packer over the
    whole pair yields.  `route` = "columns" (ssn_packer_next) or "reads" (ssn_packer_next_reads)."""

    def __init__(self, tumor, normal, fasta, min_mapq=0, threads=2, stride_windows=0, lookback_windows=0, max_reading=0, max_open=0,
                 serial=0, max_columns=1 << 16, max_entries=1 << 23):
        self.lib = load_library()
        self.h = C.c_void_p()
        o = ShardOpts(int(threads), int(stride_windows), int(lookback_windows), int(max_reading), int(max_open), int(serial))
        st = self.lib.ssn_shardset_open(os.fsencode(tumor), os.fsencode(normal), os.fsencode(fasta) if fasta else None, int(min_mapq),
                                        C.byref(o), C.byref(self.h))
        if st != 0:
            raise RuntimeError("ssn_shardset_open: %s" % self.lib.ssn_host_last_error().decode())
        self.max_columns, self.max_entries = max_columns, max_entries

    def close(self):
        if self.h:
            self.lib.ssn_shardset_close(self.h)
            self.h = C.c_void_p()

    def n_shards(self):
        return int(self.lib.ssn_shardset_n_shards(self.h))

    def ranges(self):
        out = []
        for i in range(self.n_shards()):
            a, b, c, d = C.c_int32(), C.c_uint32(), C.c_int32(), C.c_uint32()
            assert self.lib.ssn_shardset_range(self.h, i, C.byref(a), C.byref(b), C.byref(c), C.byref(d)) == 0
            out.append((a.value, b.value, c.value, d.value))
        return out

    def long_span(self):
        return bool(self.lib.ssn_shardset_long_span(self.h))

    def stats(self):
        s = PackerStats()
        self.lib.ssn_shardset_get_stats(self.h, C.byref(s))
        return s

    def shards(self):
        """Yields (index, Packer) in genome order; the packer is handed back when the consumer asks for the next one."""
        while True:
            idx, pk = C.c_int32(), C.c_void_p()
            st = self.lib.ssn_shardset_next(self.h, C.byref(idx), C.byref(pk))
            if st != 0:
                raise RuntimeError("ssn_shardset_next: %s" % self.lib.ssn_host_last_error().decode())
            if idx.value < 0:
                return
            try:
                yield idx.value, Packer(None, None, None, max_columns=self.max_columns, max_entries=self.max_entries, handle=pk)
            finally:
                st = self.lib.ssn_shardset_done(self.h, idx.value)
            if st != 0:
                raise RuntimeError("ssn_shardset_done: %s" % self.lib.ssn_host_last_error().decode())

    def __iter__(self):
        for _, pk in self.shards():
            for b in pk:
                yield b
