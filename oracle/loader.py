"""ctypes loaders for the CPU oracle.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package never does.

* ``Oracle``     — oracle/_build/libssn_oracle.so, the plain-C restatement (oracle/ssn_oracle.c).
* ``Reference``  — oracle/_ref/libref_harness.so, the UNMODIFIED reference compiled from
                   /root/reference by oracle/build_ref.sh (prebuilt file travels to the GPU box).

Both expose the same calls over the product's SoA column layout:
    off  uint32[2*ncol+1]   off[2c]..off[2c+1] tumor entries, off[2c+1]..off[2c+2] normal entries
    bs   uint8[n]           nt16 | strand<<4
    bq   uint8[n]           base quality
    mq   uint8[n]           mapping quality
    ref  uint8[ncol]        reference character (ASCII)
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "_build", "libssn_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_harness.so")
REF_CLI = os.path.join(HERE, "_ref", "bam-somaticsniper")
REF_SAMTOOLS = os.path.join(HERE, "_ref", "samtools")
REC_INTS = 54

# field names of the flattened sniper_output_t record (oracle/ref_harness.c)
SAMPLE_FIELDS = (
    ["genotype", "joint_genotype", "joint_consensus_quality", "consensus_quality",
     "variant_allele_quality", "somatic_score", "variant_status"]
    + ["mean_baseQ%d" % i for i in range(4)] + ["mean_mapQ%d" % i for i in range(4)]
    + ["base_occ%d" % i for i in range(4)] + ["dp4_%d" % i for i in range(4)]
    + ["total_depth", "total_mean_mapQ", "pad"])


def build_oracle():
    """Compile the oracle port (and the reference when /root/reference is present)."""
    subprocess.run(["make", "-s", "-C", HERE, "_build/libssn_oracle.so"], check=True)
    if os.path.isdir("/root/reference/src/lib/sniper"):
        subprocess.run([os.path.join(HERE, "build_ref.sh")], check=True, stdout=subprocess.DEVNULL)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _u8(a):
    return np.ascontiguousarray(a, dtype=np.uint8)


class _Common:
    def score_batch(self, off, bs, bq, mq, ref_char, max_recs=None, want_glf=False):
        """Returns dict(score int32[ncol], emit_col int32[k], recs int32[k,54], glf uint32[ncol,28]|None)."""
        off = np.ascontiguousarray(off, dtype=np.uint32)
        ncol = (len(off) - 1) // 2
        bs, bq, mq, ref_char = _u8(bs), _u8(bq), _u8(mq), _u8(ref_char)
        if max_recs is None:
            max_recs = ncol
        score = np.zeros(ncol, np.int32)
        emit_col = np.zeros(max(max_recs, 1), np.int32)
        recs = np.zeros((max(max_recs, 1), REC_INTS), np.int32)
        glf = np.zeros((ncol, 28), np.uint32) if want_glf else None
        n = self._score_batch(ncol, off, bs, bq, mq, ref_char, score, emit_col, recs, max_recs, glf)
        n = min(n, max_recs)
        return dict(score=score, emit_col=emit_col[:n].copy(), recs=recs[:n].copy(), glf=glf, n_emit=n)


class Oracle(_Common):
    kind = "port"

    def __init__(self, theta=0.85, het_rate=0.001, n_hap=2, use_priors=1, use_joint=0,
                 somatic_rate=0.01, min_somatic_qual=15, include_loh=1, include_gor=1):
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        L = self.L = C.CDLL(ORACLE_SO)
        L.orc_model_new.restype = C.c_void_p
        L.orc_model_new.argtypes = [C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_double,
                                    C.c_int, C.c_int, C.c_int]
        L.orc_model_free.argtypes = [C.c_void_p]
        for name, ty in (("orc_coef", C.c_double), ("orc_lhet", C.c_double), ("orc_fk", C.c_double),
                         ("orc_prior", C.c_int), ("orc_jointprior", C.c_int), ("orc_qadd", C.c_int)):
            f = getattr(L, name)
            f.restype = C.POINTER(ty)
            f.argtypes = [C.c_void_p]
        L.orc_q_r.restype = C.c_float
        L.orc_q_r.argtypes = [C.c_void_p]
        L.orc_glfgen_flat.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_void_p, C.c_void_p]
        L.orc_dqstats.argtypes = [C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_uint32, C.c_void_p]
        L.orc_score_batch.restype = C.c_int
        L.orc_score_batch.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 9 + [C.c_int, C.c_void_p]
        L.orc_nt16_of_char.argtypes = [C.c_int]
        self.m = L.orc_model_new(theta, het_rate, n_hap, use_priors, use_joint, somatic_rate,
                                 min_somatic_qual, include_loh, include_gor)

    def __del__(self):
        try:
            self.L.orc_model_free(self.m)
        except Exception:
            pass

    def tables(self):
        L, m = self.L, self.m
        arr = np.ctypeslib.as_array
        return dict(coef=arr(L.orc_coef(m), shape=(64 * 256 * 256,)).copy(),
                    lhet=arr(L.orc_lhet(m), shape=(65536,)).copy(),
                    fk=arr(L.orc_fk(m), shape=(256,)).copy(),
                    q_r=np.float32(L.orc_q_r(m)),
                    prior=arr(L.orc_prior(m), shape=(16, 10)).copy(),
                    jointprior=arr(L.orc_jointprior(m), shape=(16, 10, 10)).copy(),
                    qadd=arr(L.orc_qadd(m), shape=(1024,)).copy())

    def nt16(self, chars):
        return np.array([self.L.orc_nt16_of_char(int(c)) for c in _u8(chars)], np.uint8)

    def glfgen(self, bs, bq, mq, ref_nt16, floats=False):
        bs, bq, mq = _u8(bs), _u8(bq), _u8(mq)
        out = np.zeros(14, np.uint32)
        fl = np.zeros(24, np.float32)
        self.L.orc_glfgen_flat(self.m, len(bs), _p(bs), _p(bq), _p(mq), int(ref_nt16), _p(out), _p(fl))
        return (out, fl) if floats else out

    def dqstats(self, bs, bq, mq, ref_base4, wanted):
        bs, bq, mq = _u8(bs), _u8(bq), _u8(mq)
        out = np.zeros(18, np.uint32)
        self.L.orc_dqstats(len(bs), _p(bs), _p(bq), _p(mq), int(ref_base4), int(wanted), _p(out))
        return out

    def _score_batch(self, ncol, off, bs, bq, mq, ref_char, score, emit_col, recs, max_recs, glf):
        nt16 = _LUT_NT16[ref_char]
        return self.L.orc_score_batch(self.m, ncol, _p(off), _p(bs), _p(bq), _p(mq), _p(ref_char), _p(nt16),
                                      _p(score), _p(emit_col), _p(recs), max_recs,
                                      _p(glf) if glf is not None else None)


def _make_nt16_lut():
    lut = np.full(256, 15, np.uint8)
    for i, ch in enumerate("=ACMGRSVTWYHKDBN"):
        lut[ord(ch)] = i
        if ch.isalpha():
            lut[ord(ch.lower())] = i
    for i in range(4):
        lut[ord("0") + i] = 1 << i
    return lut


_LUT_NT16 = _make_nt16_lut()


def nt16_lut():
    """bam_nt16_table (samtools-0.1.6 bam_import.c:23-40) as a numpy LUT."""
    return _LUT_NT16.copy()


class Reference(_Common):
    """The unmodified reference (global state: one model at a time per process)."""
    kind = "reference"

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def __init__(self, theta=0.85, het_rate=0.001, n_hap=2, use_priors=1, use_joint=0,
                 somatic_rate=0.01, min_somatic_qual=15, include_loh=1, include_gor=1):
        L = self.L = C.CDLL(REF_SO)
        L.ref_init.argtypes = [C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int]
        for name, ty in (("ref_coef", C.c_double), ("ref_lhet", C.c_double), ("ref_fk", C.c_double),
                         ("ref_prior", C.c_int), ("ref_jointprior", C.c_int), ("ref_qadd", C.c_int)):
            getattr(L, name).restype = C.POINTER(ty)
        L.ref_q_r.restype = C.c_float
        L.ref_glfgen.argtypes = [C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_void_p]
        L.ref_dqstats.argtypes = [C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_uint32, C.c_void_p]
        L.ref_score_batch.restype = C.c_int
        L.ref_score_batch.argtypes = [C.c_int] + [C.c_void_p] * 8 + [C.c_int]
        L.ref_init(theta, het_rate, n_hap, use_priors, use_joint, somatic_rate, min_somatic_qual,
                   include_loh, include_gor)

    def tables(self):
        L = self.L
        arr = np.ctypeslib.as_array
        return dict(coef=arr(L.ref_coef(), shape=(64 * 256 * 256,)).copy(),
                    lhet=arr(L.ref_lhet(), shape=(65536,)).copy(),
                    fk=arr(L.ref_fk(), shape=(256,)).copy(),
                    q_r=np.float32(L.ref_q_r()),
                    prior=arr(L.ref_prior(), shape=(16, 10)).copy(),
                    jointprior=arr(L.ref_jointprior(), shape=(16, 10, 10)).copy(),
                    qadd=arr(L.ref_qadd(), shape=(1024,)).copy())

    def nt16(self, chars):
        return np.array([self.L.ref_nt16_of_char(int(c)) for c in _u8(chars)], np.uint8)

    def glfgen(self, bs, bq, mq, ref_nt16):
        bs, bq, mq = _u8(bs), _u8(bq), _u8(mq)
        out = np.zeros(14, np.uint32)
        self.L.ref_glfgen(len(bs), _p(bs), _p(bq), _p(mq), int(ref_nt16), _p(out))
        return out

    def dqstats(self, bs, bq, mq, ref_base4, wanted):
        bs, bq, mq = _u8(bs), _u8(bq), _u8(mq)
        out = np.zeros(18, np.uint32)
        self.L.ref_dqstats(len(bs), _p(bs), _p(bq), _p(mq), int(ref_base4), int(wanted), _p(out))
        return out

    def _score_batch(self, ncol, off, bs, bq, mq, ref_char, score, emit_col, recs, max_recs, glf):
        if glf is not None:
            nt16 = _LUT_NT16[ref_char]
            for c in range(ncol):
                t0, n0, e = int(off[2 * c]), int(off[2 * c + 1]), int(off[2 * c + 2])
                glf[c, :14] = self.glfgen(bs[t0:n0], bq[t0:n0], mq[t0:n0], nt16[c])
                glf[c, 14:] = self.glfgen(bs[n0:e], bq[n0:e], mq[n0:e], nt16[c])
        return self.L.ref_score_batch(ncol, _p(off), _p(bs), _p(bq), _p(mq), _p(ref_char),
                                      _p(score), _p(emit_col), _p(recs), max_recs)
