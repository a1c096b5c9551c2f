"""ctypes binding of libssn.so (include/ssn.h) — the host-side mirror used by tests and bench.py.

The product's host side is C (the packer / CLI link libssn.so directly); this module exists so the
parity tests can drive the same C ABI from Python.  It contains no scoring logic and NO fallback: if
libssn.so is missing or no CUDA device is usable, construction raises.

Names follow the reference's vocabulary: a *column* is one reference position with the tumor and the
normal pileup; a *site* is a column that passed glf_somatic's gate (somatic_sniper.c:225-227).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SSN_LIB_PATH") or os.path.join(HERE, "lib", "libssn.so")   # override: kernel experiments only
COLUMN_DUMP_WORDS = 30     # [29] = 1 if the hom-reference pre-filter would have dropped the column
SUBMIT_COLUMN_DUMP = 1
SUBMIT_NO_PREFILTER = 2    # every column through the full model (identical results; measurement only)


class Params(C.Structure):
    """ssn_params (include/ssn.h); defaults = main.c:69-78 + sniper_maqcns_init."""
    _fields_ = [("theta", C.c_float), ("het_rate", C.c_float), ("eta", C.c_float),
                ("n_hap", C.c_int32), ("cap_mapq", C.c_int32), ("min_somatic_qual", C.c_int32),
                ("include_loh", C.c_int32), ("include_gor", C.c_int32), ("use_priors", C.c_int32),
                ("use_joint_priors", C.c_int32), ("somatic_mutation_rate", C.c_double)]


class Tables(C.Structure):
    _fields_ = [("coef", C.POINTER(C.c_double)), ("lhet", C.POINTER(C.c_double)), ("fk", C.POINTER(C.c_double)),
                ("q_r", C.c_float), ("prior", C.POINTER(C.c_int32)), ("jointprior", C.POINTER(C.c_int32)),
                ("qadd", C.POINTER(C.c_int32))]


class PrefilterView(C.Structure):
    """ssn_prefilter_view: the pre-filter's bound tables (ssn_prefilter_tables.c)."""
    _fields_ = [("ab", C.POINTER(C.c_float)), ("h", C.POINTER(C.c_float)), ("F", C.POINTER(C.c_float)),
                ("vt", C.POINTER(C.c_float)), ("n", C.c_int32), ("enabled", C.c_int32)]


class BatchView(C.Structure):
    _fields_ = [("off", C.POINTER(C.c_uint32)), ("bs", C.POINTER(C.c_uint8)), ("bq", C.POINTER(C.c_uint8)),
                ("mq", C.POINTER(C.c_uint8)), ("ref_nt16", C.POINTER(C.c_uint8)),
                ("max_columns", C.c_uint32), ("max_entries", C.c_uint64)]


class ReadsView(C.Structure):
    """ssn_reads_view: pinned arrays of one read batch (device pileup)."""
    _fields_ = [("recs", C.POINTER(C.c_uint8)), ("rec_off", C.POINTER(C.c_uint32)), ("win", C.c_void_p),
                ("ref_nt16", C.POINTER(C.c_uint8)), ("max_rec_bytes", C.c_uint64), ("max_reads", C.c_uint32),
                ("max_windows", C.c_uint32)]


READ_WINDOW = 1 << 13
READ_REC_DTYPE = np.dtype([("beg", "<i4"), ("end", "<i4"), ("l_qseq", "<u4"), ("n_cigar", "<u2"), ("mapq", "u1"),
                           ("strand", "u1"), ("bytes", "<u4")])
READ_WINDOW_DTYPE = np.dtype([("rec_first", "<u4", 2), ("n_reads", "<u4", 2), ("lo", "<u4"), ("tid", "<i4")])
assert READ_REC_DTYPE.itemsize == 20 and READ_WINDOW_DTYPE.itemsize == 24

SAMPLE_DTYPE = np.dtype([
    ("genotype", "<i4"), ("joint_genotype", "<i4"), ("joint_consensus_quality", "<i4"),
    ("consensus_quality", "<i4"), ("variant_allele_quality", "<i4"), ("somatic_score", "<i4"),
    ("variant_status", "<i4"), ("mean_baseQ", "<u4", 4), ("mean_mapQ", "<u4", 4), ("base_occ", "<u4", 4),
    ("dp4", "<u4", 4), ("total_depth", "<u4"), ("total_mean_mapQ", "<u4"), ("rms_mapq", "<i4")])
SITE_DTYPE = np.dtype([("column", "<u4"), ("ref_base4", "<i4"), ("tumor", SAMPLE_DTYPE), ("normal", SAMPLE_DTYPE)])
assert SITE_DTYPE.itemsize == 216

# every function include/ssn.h declares: name -> (restype, argtypes)
_VP, _U32, _U64, _I = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
SIGNATURES = {
    "ssn_abi_version": (_I, []),
    "ssn_last_error": (C.c_char_p, []),
    "ssn_params_default": (_I, [C.POINTER(Params)]),
    "ssn_create": (_I, [C.POINTER(Params), _I, _U32, _U32, _U64, C.POINTER(_VP)]),
    "ssn_destroy": (None, [_VP]),
    "ssn_get_tables": (_I, [_VP, C.POINTER(Tables)]),
    "ssn_host_tables_create": (_I, [C.POINTER(Params), C.POINTER(Tables), C.POINTER(_VP)]),
    "ssn_host_tables_destroy": (None, [_VP]),
    "ssn_nt16_of_char": (_I, [_I]),
    "ssn_batch_acquire": (_I, [_VP, C.POINTER(_VP)]),
    "ssn_batch_view_get": (_I, [_VP, C.POINTER(BatchView)]),
    "ssn_batch_set_size": (_I, [_VP, _U32, _U64]),
    "ssn_batch_submit": (_I, [_VP, _VP, _U32]),
    "ssn_results_next": (_I, [_VP, C.POINTER(_VP), C.POINTER(_U32), C.POINTER(_VP)]),
    "ssn_batch_column_dump": (_I, [_VP, C.POINTER(_VP)]),
    "ssn_batch_release": (_I, [_VP, _VP]),
    "ssn_batch_upload": (_I, [_VP, _VP]),
    "ssn_batch_score_resident": (_I, [_VP, _VP, _U32]),
    "ssn_batch_fetch": (_I, [_VP, _VP, C.POINTER(_VP), C.POINTER(_U32)]),
    "ssn_timer_start": (_I, [_VP]),
    "ssn_timer_stop": (_I, [_VP, C.POINTER(C.c_float)]),
    "ssn_sync": (_I, [_VP]),
    "ssn_kernel_launches": (_U64, [_VP]),
    "ssn_kernel_time": (_I, [_VP, C.POINTER(C.c_float), C.POINTER(_U64), _I]),
    "ssn_flush_l2": (_I, [_VP]),
    "ssn_kernel_time_split": (_I, [_VP, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(_U64), _I]),
    "ssn_prefilter_enabled": (_I, [_VP]),
    "ssn_batch_survivors": (_I, [_VP, C.POINTER(_U32)]),
    "ssn_host_prefilter_create": (_I, [C.POINTER(Params), C.POINTER(PrefilterView), C.POINTER(_VP)]),
    "ssn_host_prefilter_destroy": (None, [_VP]),
    "ssn_reads_reserve": (_I, [_VP, _U64, _U32, _U32]),
    "ssn_batch_reads_view": (_I, [_VP, _VP, C.POINTER(ReadsView)]),
    "ssn_batch_submit_reads": (_I, [_VP, _VP, _U32, _U32, _U64, _U32]),
    "ssn_batch_column_positions": (_I, [_VP, C.POINTER(_VP), C.POINTER(_U32), C.POINTER(_U64)]),
    "ssn_batch_download_columns": (_I, [_VP, _VP]),
    "ssn_pileup_time": (_I, [_VP, C.POINTER(C.c_float), C.POINTER(_U64), _I]),
    "ssn_inflater_create": (_I, [_I, _U32, _U64, _U64, _U32, C.POINTER(_VP)]),
    "ssn_inflater_destroy": (None, [_VP]),
    "ssn_inflater_limits": (_I, [_VP, C.POINTER(_U64), C.POINTER(_U64), C.POINTER(_U32)]),
    "ssn_inflater_lane_acquire": (_I, [_VP, C.POINTER(C.c_int32), C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.POINTER(C.c_uint8))]),
    "ssn_inflater_lane_release": (None, [_VP, C.c_int32]),
    "ssn_inflater_run": (_I, [_VP, C.c_int32, _VP, _U32, _U64, _U64, C.POINTER(C.c_uint8)]),
    "ssn_inflater_stats": (_I, [_VP, C.POINTER(_U64), C.POINTER(_U64), C.POINTER(_U64), C.POINTER(_U64), C.POINTER(_U64)]),
    "ssn_inflater_run_to": (_I, [_VP, C.c_int32, _VP, _U32, _U64, _U64, C.POINTER(C.c_uint8), _VP]),
    "ssn_inflater_host_alloc": (_I, [_VP, _U64, C.POINTER(_VP)]),
    "ssn_inflater_host_free": (None, [_VP, _VP]),
    "ssn_inflate_static": (_I, [C.c_int, _VP, _U64, _VP, _U32, _VP, _U64, _VP, _VP]),
}

_lib = None


class SsnError(RuntimeError):
    pass


def load_library():
    """dlopen libssn.so and type every entry point; raises if the extension was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SsnError("libssn.so is not built (%s); run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "- there is no CPU fallback" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(lib, name)
            f.restype = res
            f.argtypes = args
        _lib = lib
    return _lib


def _check(st):
    if st != 0:
        raise SsnError("libssn status %d: %s" % (st, load_library().ssn_last_error().decode()))


def default_params(**over):
    p = Params()
    _check(load_library().ssn_params_default(C.byref(p)))
    for k, v in over.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def _tables_to_numpy(t):
    arr = np.ctypeslib.as_array
    return dict(coef=arr(t.coef, shape=(64 * 65536,)).copy(), lhet=arr(t.lhet, shape=(65536,)).copy(),
                fk=arr(t.fk, shape=(256,)).copy(), q_r=np.float32(t.q_r),
                prior=arr(t.prior, shape=(16, 10)).copy(), jointprior=arr(t.jointprior, shape=(16, 10, 10)).copy(),
                qadd=arr(t.qadd, shape=(1024,)).copy())


def host_tables(params=None):
    """Model tables built by the product's host code, no device needed."""
    lib = load_library()
    p = params or default_params()
    t, h = Tables(), _VP()
    _check(lib.ssn_host_tables_create(C.byref(p), C.byref(t), C.byref(h)))
    try:
        return _tables_to_numpy(t)
    finally:
        lib.ssn_host_tables_destroy(h)


def host_prefilter_tables(params=None):
    """The pre-filter's bound tables built by the product's host code (no device needed):
    dict(ab float32[N,N,4], h float32[N,N], F float32[N+1], vt float32[N,8,256], enabled bool)."""
    lib = load_library()
    p = params or default_params()
    v, h = PrefilterView(), _VP()
    _check(lib.ssn_host_prefilter_create(C.byref(p), C.byref(v), C.byref(h)))
    try:
        n = int(v.n)
        arr = np.ctypeslib.as_array
        return dict(ab=arr(v.ab, shape=(n, n, 4)).copy(), h=arr(v.h, shape=(n, n)).copy(),
                    F=arr(v.F, shape=(n + 1,)).copy(), vt=arr(v.vt, shape=(n, 8, 256)).copy(), enabled=bool(v.enabled))
    finally:
        lib.ssn_host_prefilter_destroy(h)


def nt16_of_chars(chars):
    lib = load_library()
    return np.array([lib.ssn_nt16_of_char(int(c)) for c in np.asarray(chars, np.uint8)], np.uint8)


_NT16_LUT = None


def nt16_lut():
    global _NT16_LUT
    if _NT16_LUT is None:
        _NT16_LUT = nt16_of_chars(np.arange(256, dtype=np.uint8))
    return _NT16_LUT


class Batch:
    """One pinned SoA slot of a Scorer (ssn_batch)."""

    def __init__(self, scorer, handle):
        self.scorer, self.h = scorer, handle
        v = BatchView()
        _check(scorer.lib.ssn_batch_view_get(handle, C.byref(v)))
        self.max_columns, self.max_entries = int(v.max_columns), int(v.max_entries)
        arr = np.ctypeslib.as_array
        self.off = arr(v.off, shape=(2 * self.max_columns + 1,))
        self.bs = arr(v.bs, shape=(self.max_entries,))
        self.bq = arr(v.bq, shape=(self.max_entries,))
        self.mq = arr(v.mq, shape=(self.max_entries,))
        self.ref_nt16 = arr(v.ref_nt16, shape=(self.max_columns,))
        self.ncol = 0
        self.nent = 0

    def fill(self, cols):
        """cols: dict from synth.make_columns (off, bs, bq, mq, ref_char)."""
        ncol, n = int(cols["ncol"]), int(cols["n_entries"])
        if ncol > self.max_columns or n > self.max_entries:
            raise SsnError("batch of %d columns / %d entries exceeds slot capacity" % (ncol, n))
        self.off[:2 * ncol + 1] = cols["off"]
        self.bs[:n] = cols["bs"]
        self.bq[:n] = cols["bq"]
        self.mq[:n] = cols["mq"]
        self.ref_nt16[:ncol] = nt16_lut()[cols["ref_char"]]
        self.set_size(ncol, n)

    def set_size(self, ncol, n):
        _check(self.scorer.lib.ssn_batch_set_size(self.h, ncol, n))
        self.ncol, self.nent = ncol, n

    def survivors(self):
        """Columns the pre-filter handed to the full model (valid once the batch's results were drained)."""
        n = _U32()
        _check(self.scorer.lib.ssn_batch_survivors(self.h, C.byref(n)))
        return int(n.value)

    def column_dump(self):
        p = _VP()
        _check(self.scorer.lib.ssn_batch_column_dump(self.h, C.byref(p)))
        a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), shape=(self.ncol, COLUMN_DUMP_WORDS))
        return a.copy()


def _sites_to_numpy(ptr, n):
    if n == 0:
        return np.zeros(0, SITE_DTYPE)
    buf = C.string_at(ptr, n * SITE_DTYPE.itemsize)
    return np.frombuffer(buf, SITE_DTYPE).copy()


class Scorer:
    """ssn_ctx: the drop-in for bam_sspileup_file(..., glf_somatic, ...) at batch granularity."""

    def __init__(self, params=None, device=0, n_slots=2, max_columns=1 << 16, max_entries=1 << 23):
        self.lib = load_library()
        self.params = params or default_params()
        self.ctx = _VP()
        _check(self.lib.ssn_create(C.byref(self.params), device, n_slots, max_columns, max_entries, C.byref(self.ctx)))

    def close(self):
        if self.ctx:
            self.lib.ssn_destroy(self.ctx)
            self.ctx = _VP()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def tables(self):
        t = Tables()
        _check(self.lib.ssn_get_tables(self.ctx, C.byref(t)))
        return _tables_to_numpy(t)

    # -- pipeline ---------------------------------------------------------------------------------
    def acquire(self):
        h = _VP()
        _check(self.lib.ssn_batch_acquire(self.ctx, C.byref(h)))
        return Batch(self, h)

    def submit(self, batch, flags=0):
        _check(self.lib.ssn_batch_submit(self.ctx, batch.h, flags))

    def results_next(self):
        sites, n, b = _VP(), _U32(), _VP()
        _check(self.lib.ssn_results_next(self.ctx, C.byref(sites), C.byref(n), C.byref(b)))
        return _sites_to_numpy(sites, n.value), b.value

    def release(self, batch):
        _check(self.lib.ssn_batch_release(self.ctx, batch.h))

    def score(self, cols, dump=False, flags=0):
        """Convenience: one batch through acquire/fill/submit/results_next. Returns (sites, dump|None)."""
        b = self.acquire()
        b.fill(cols)
        self.submit(b, flags | (SUBMIT_COLUMN_DUMP if dump else 0))
        sites, _ = self.results_next()
        d = b.column_dump() if dump else None
        self.last_survivors = b.survivors()
        self.release(b)
        return sites, d

    def prefilter_enabled(self):
        return bool(self.lib.ssn_prefilter_enabled(self.ctx))

    # -- read batches: the device builds the columns (ssn_pileup.cuh) -------------------------------
    def reads_reserve(self, max_rec_bytes, max_reads, max_windows):
        _check(self.lib.ssn_reads_reserve(self.ctx, max_rec_bytes, max_reads, max_windows))

    def reads_view(self, batch):
        """numpy views of the slot's pinned read arrays: dict(recs u8, rec_off u32, win READ_WINDOW_DTYPE, ref_nt16 u8[w, 8192])."""
        v = ReadsView()
        _check(self.lib.ssn_batch_reads_view(self.ctx, batch.h, C.byref(v)))
        arr = np.ctypeslib.as_array
        nw = int(v.max_windows)
        win = arr(C.cast(v.win, C.POINTER(C.c_uint8)), shape=(nw * READ_WINDOW_DTYPE.itemsize,)).view(READ_WINDOW_DTYPE)
        return dict(recs=arr(v.recs, shape=(int(v.max_rec_bytes),)), rec_off=arr(v.rec_off, shape=(int(v.max_reads),)),
                    win=win, ref_nt16=arr(v.ref_nt16, shape=(nw, READ_WINDOW)))

    def fill_reads(self, batch, rw):
        """rw: dict from reads.build_read_windows. Copies it into the slot's pinned arrays; returns the submit arguments."""
        v = self.reads_view(batch)
        nb, nr, nw = rw["recs"].size, rw["rec_off"].size, rw["win"].size
        if nb > v["recs"].size or nr > v["rec_off"].size or nw > v["win"].size:
            raise SsnError("read batch of %d bytes / %d reads / %d windows exceeds the reserved capacity" % (nb, nr, nw))
        v["recs"][:nb] = rw["recs"]
        v["rec_off"][:nr] = rw["rec_off"]
        v["win"][:nw] = rw["win"]
        v["ref_nt16"][:nw] = rw["ref_nt16"]
        return nw, nr, nb

    def submit_reads(self, batch, n_windows, n_reads, rec_bytes, flags=0):
        _check(self.lib.ssn_batch_submit_reads(self.ctx, batch.h, n_windows, n_reads, rec_bytes, flags))

    def column_positions(self, batch):
        """(pos u32[ncol] = window << 13 | offset, n_entries) of a finished read batch."""
        p, n, e = _VP(), _U32(), _U64()
        _check(self.lib.ssn_batch_column_positions(batch.h, C.byref(p), C.byref(n), C.byref(e)))
        pos = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), shape=(int(n.value),)).copy() if n.value else np.zeros(0, np.uint32)
        return pos, int(e.value)

    def download_columns(self, batch):
        """The column batch the device built from a read batch, as a dict like synth.make_columns' (minus ref_char)."""
        pos, nent = self.column_positions(batch)
        _check(self.lib.ssn_batch_download_columns(self.ctx, batch.h))
        ncol = pos.size
        return dict(ncol=ncol, n_entries=nent, off=batch.off[:2 * ncol + 1].copy(), bs=batch.bs[:nent].copy(),
                    bq=batch.bq[:nent].copy(), mq=batch.mq[:nent].copy(), ref_nt16=batch.ref_nt16[:ncol].copy(), pos=pos)

    def pileup_time(self, reset=False):
        ms, n = C.c_float(), _U64()
        _check(self.lib.ssn_pileup_time(self.ctx, C.byref(ms), C.byref(n), 1 if reset else 0))
        return ms.value, int(n.value)

    # -- resident path ------------------------------------------------------------------------------
    def upload(self, batch):
        _check(self.lib.ssn_batch_upload(self.ctx, batch.h))

    def score_resident(self, batch, flags=0):
        _check(self.lib.ssn_batch_score_resident(self.ctx, batch.h, flags))

    def fetch(self, batch):
        sites, n = _VP(), _U32()
        _check(self.lib.ssn_batch_fetch(self.ctx, batch.h, C.byref(sites), C.byref(n)))
        return _sites_to_numpy(sites, n.value)

    # -- timing ---------------------------------------------------------------------------------------
    def timer_start(self):
        _check(self.lib.ssn_timer_start(self.ctx))

    def timer_stop(self):
        ms = C.c_float()
        _check(self.lib.ssn_timer_stop(self.ctx, C.byref(ms)))
        return ms.value

    def sync(self):
        _check(self.lib.ssn_sync(self.ctx))

    def kernel_launches(self):
        return int(self.lib.ssn_kernel_launches(self.ctx))

    def kernel_time(self, reset=False):
        ms, n = C.c_float(), _U64()
        _check(self.lib.ssn_kernel_time(self.ctx, C.byref(ms), C.byref(n), 1 if reset else 0))
        return ms.value, int(n.value)

    def kernel_time_split(self, reset=False):
        """(pre-filter ms, scoring ms, batches) summed over the batches timed since the last reset."""
        a, b, n = C.c_float(), C.c_float(), _U64()
        _check(self.lib.ssn_kernel_time_split(self.ctx, C.byref(a), C.byref(b), C.byref(n), 1 if reset else 0))
        return a.value, b.value, int(n.value)

    def flush_l2(self):
        _check(self.lib.ssn_flush_l2(self.ctx))


def sites_to_records(sites, ref_chars):
    """ssn_site array -> the oracle's flattened sniper_output_t layout (int32[k,54]) for comparison."""
    k = len(sites)
    rec = np.zeros((k, 54), np.int32)
    if k == 0:
        return rec
    rec[:, 0] = np.asarray(ref_chars, np.uint8)[sites["column"]]
    rec[:, 1] = sites["ref_base4"]
    for base, s in ((2, sites["tumor"]), (28, sites["normal"])):
        for i, f in enumerate(["genotype", "joint_genotype", "joint_consensus_quality", "consensus_quality",
                               "variant_allele_quality", "somatic_score", "variant_status"]):
            rec[:, base + i] = s[f]
        rec[:, base + 7:base + 11] = s["mean_baseQ"]
        rec[:, base + 11:base + 15] = s["mean_mapQ"]
        rec[:, base + 15:base + 19] = s["base_occ"]
        rec[:, base + 19:base + 23] = s["dp4"]
        rec[:, base + 23] = s["total_depth"]
        rec[:, base + 24] = s["total_mean_mapQ"]
    return rec


INFLATE_BLOCK_DTYPE = np.dtype([("in_off", np.uint32), ("in_len", np.uint32), ("out_off", np.uint32), ("out_len", np.uint32)])


INFLATE_STATIC_MAX_BLOCKS = 4096          # include/ssn.h: SSN_INFLATE_STATIC_MAX_BLOCKS


def inflate_static(device, d_in, in_bytes, blocks, d_out, out_bytes, d_status, stream=0):
    """ssn_inflate_static: d_in / d_out / d_status are device addresses (ints) the caller owns; nothing is allocated."""
    lib = load_library()
    blocks = np.ascontiguousarray(blocks, INFLATE_BLOCK_DTYPE)
    _check(lib.ssn_inflate_static(device, d_in, in_bytes, blocks.ctypes.data, len(blocks), d_out, out_bytes, d_status, stream))


class Inflater:
    """ssn_inflater: BGZF blocks inflated on the GPU (csrc/ssn_inflate_api.cu); mirror for tests and tools."""

    def __init__(self, device=0, n_lanes=4, in_cap=8 << 20, out_cap=16 << 20, max_blocks=256):
        self.lib = load_library()
        self.h = _VP()
        _check(self.lib.ssn_inflater_create(device, n_lanes, in_cap, out_cap, max_blocks, C.byref(self.h)))
        self.in_cap, self.out_cap, self.max_blocks = in_cap, out_cap, max_blocks

    def close(self):
        if self.h:
            self.lib.ssn_inflater_destroy(self.h)
            self.h = _VP()

    def acquire(self):
        lane, pi, po = C.c_int32(), C.POINTER(C.c_uint8)(), C.POINTER(C.c_uint8)()
        _check(self.lib.ssn_inflater_lane_acquire(self.h, C.byref(lane), C.byref(pi), C.byref(po)))
        return lane.value, np.ctypeslib.as_array(pi, shape=(self.in_cap,)), np.ctypeslib.as_array(po, shape=(self.out_cap,))

    def release(self, lane):
        self.lib.ssn_inflater_lane_release(self.h, lane)

    def run(self, lane, blocks, in_bytes, out_bytes):
        blocks = np.ascontiguousarray(blocks, INFLATE_BLOCK_DTYPE)
        refused = np.zeros(len(blocks), np.uint8)
        _check(self.lib.ssn_inflater_run(self.h, lane, blocks.ctypes.data, len(blocks), in_bytes, out_bytes,
                                         refused.ctypes.data_as(C.POINTER(C.c_uint8))))
        return refused

    def inflate_many(self, payloads, sizes):
        """payloads: list of raw deflate streams, sizes: their inflated sizes -> (list of bytes or None where refused)"""
        lane, hin, hout = self.acquire()
        try:
            out = []
            i = 0
            while i < len(payloads):
                blk, a, b = [], 0, 0
                j = i
                while j < len(payloads) and len(blk) < self.max_blocks and a + len(payloads[j]) + 4 <= self.in_cap and b + sizes[j] <= self.out_cap:
                    hin[a:a + len(payloads[j])] = np.frombuffer(payloads[j], np.uint8)
                    blk.append((a, len(payloads[j]), b, sizes[j]))
                    a += len(payloads[j]); b += sizes[j]
                    j += 1
                assert j > i, "a block does not fit the lane"
                refused = self.run(lane, np.array(blk, INFLATE_BLOCK_DTYPE), a, b)
                for (ao, al, bo, bl), r in zip(blk, refused):
                    out.append(None if r else hout[bo:bo + bl].tobytes())
                i = j
            return out
        finally:
            self.release(lane)
