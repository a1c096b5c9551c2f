"""Reads -> pileup columns.  CPU: the host packer's count/fill code over whole read windows (ssn_pack_read_windows, the CPU
side of the device pileup's parity) against a straight numpy restatement of the column construction
(src/lib/sniper/sniper_pileup.c:57-104,186-205).  GPU: the device pileup (csrc/ssn_pileup.cuh) against the host packer,
byte for byte, and the sites scored from both."""
import numpy as np
import pytest

from somatic_sniper_b200 import reads as R

COL_KEYS = ("off", "bs", "bq", "mq", "ref_nt16", "pos")


def make_case(seed, length, depth_t, depth_n, complex_cigars, start=0, q0=0.0, read_len=100):
    ref = R.make_reference(seed, start + length + 4 * read_len)
    ref_t = ref.copy()                                   # the tumor carries a variant every 97 positions
    ref_t[::97] = np.where(ref_t[::97] == 1, 2, 1)
    t = R.make_reads(seed + 1, ref_t, depth_t, read_len, start, length, complex_cigars, q0_frac=q0)
    n = R.make_reads(seed + 2, ref, depth_n, read_len, start, length, complex_cigars, q0_frac=q0)
    return ref, t, n


def same_columns(a, b):
    assert a["ncol"] == b["ncol"] and a["n_entries"] == b["n_entries"]
    for k in COL_KEYS:
        assert np.array_equal(a[k], b[k]), k


CASES = {
    "plain": dict(seed=1, length=3000, depth_t=6, depth_n=4, complex_cigars=False),
    "cigars": dict(seed=2, length=3000, depth_t=8, depth_n=5, complex_cigars=True),
    "window edge": dict(seed=3, length=2500, depth_t=5, depth_n=5, complex_cigars=True, start=8192 - 1200),
    "second window, quality 0": dict(seed=4, length=2000, depth_t=4, depth_n=7, complex_cigars=True, start=3 * 8192 + 17, q0=0.05),
    "sparse normal": dict(seed=5, length=4000, depth_t=5, depth_n=0.3, complex_cigars=True),
    "random cigars": dict(seed=6, length=3000, depth_t=12, depth_n=9, complex_cigars="random", start=8192 - 1500),
    "random cigars, quality 0": dict(seed=7, length=2500, depth_t=9, depth_n=14, complex_cigars="random", q0=0.1),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_host_windows_against_numpy(built, name):
    from somatic_sniper_b200 import hostapi
    ref, t, n = make_case(**CASES[name])
    rw = R.build_read_windows(t, n, ref)
    fw = int(rw["win"]["lo"][0]) >> 13
    want = R.pileup_numpy(t, n, ref, fw, rw["win"].size)
    got = hostapi.pack_read_windows(rw)
    assert want["ncol"] > 100
    same_columns(got, want)


def test_host_windows_reject_scattered_records(built):
    from somatic_sniper_b200 import hostapi
    ref, t, n = make_case(seed=9, length=1000, depth_t=3, depth_n=3, complex_cigars=False)
    rw = R.build_read_windows(t, n, ref)
    rw["rec_off"] = rw["rec_off"].copy()
    rw["rec_off"][1] += 1
    with pytest.raises(RuntimeError, match="not contiguous"):
        hostapi.pack_read_windows(rw)


# ---- device ---------------------------------------------------------------------------------------------------------

GPU_CASES = dict(CASES)
GPU_CASES["wgs 60x/30x, 5 windows"] = dict(seed=11, length=5 * 8192 - 300, depth_t=60, depth_n=30, complex_cigars=False, read_len=150)
GPU_CASES["cigars 40x/40x"] = dict(seed=12, length=2 * 8192, depth_t=40, depth_n=40, complex_cigars=True, start=8192 - 77, read_len=150)
GPU_CASES["deep 600x"] = dict(seed=13, length=3000, depth_t=600, depth_n=300, complex_cigars=True, start=5000)
NT16_CHARS = np.frombuffer(b"=ACMGRSVTWYHKDBN", np.uint8)


def device_pileup(api, rw, flags=0, max_columns=None, max_entries=None):
    nw = rw["win"].size
    sc = api.Scorer(n_slots=2, max_columns=max_columns or nw * 8192, max_entries=max_entries or rw["entries_bound"])
    sc.reads_reserve(rw["recs"].size + 64, rw["rec_off"].size + 1, nw)
    b = sc.acquire()
    sc.submit_reads(b, *sc.fill_reads(b, rw), flags=flags)
    sites, _ = sc.results_next()
    return sc, b, sites


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(GPU_CASES))
def test_device_pileup_identical_to_host_packer(built, name):
    from somatic_sniper_b200 import api, hostapi
    ref, t, n = make_case(**GPU_CASES[name])
    rw = R.build_read_windows(t, n, ref)
    want = hostapi.pack_read_windows(rw)
    sc, b, sites = device_pileup(api, rw)
    got = sc.download_columns(b)
    same_columns(got, want)
    sc.release(b)
    # the same batch packed on the host, through the ordinary submit
    cols = dict(want, ref_char=NT16_CHARS[want["ref_nt16"]])
    sites_host, _ = sc.score(cols)
    assert sites.size == sites_host.size and sites.size > 5
    assert sites.tobytes() == sites_host.tobytes()
    assert sc.pileup_time()[1] == 1
    sc.close()


@pytest.mark.gpu
def test_device_pileup_errors(built):
    from somatic_sniper_b200 import api
    ref, t, n = make_case(seed=21, length=3000, depth_t=10, depth_n=10, complex_cigars=False)
    rw = R.build_read_windows(t, n, ref)
    with pytest.raises(api.SsnError, match="exceed the slot capacity"):
        device_pileup(api, rw, max_columns=1000)
    with pytest.raises(api.SsnError, match="exceed the slot capacity"):
        device_pileup(api, rw, max_entries=20000)
    # swap the first two tumor reads' starts: no longer sorted
    bad = dict(rw, recs=rw["recs"].copy())
    w = bad["recs"].view("<u4")
    o = bad["rec_off"]
    w[o[0]], w[o[5]] = w[o[5]], w[o[0]]
    if w[o[0]] != w[o[5]]:
        with pytest.raises(api.SsnError, match="not sorted"):
            device_pileup(api, bad)
    # an index entry past the buffer, a record whose sequence would run past it
    bad = dict(rw, rec_off=rw["rec_off"].copy())
    bad["rec_off"][3] = rw["recs"].size // 4 + 7
    with pytest.raises(api.SsnError, match="outside the record buffer"):
        device_pileup(api, bad)
    bad = dict(rw, recs=rw["recs"].copy())
    bad["recs"].view("<u4")[rw["rec_off"][-1] + 2] = 1 << 30
    with pytest.raises(api.SsnError, match="outside the record buffer"):
        device_pileup(api, bad)
    sc = api.Scorer(n_slots=1)
    b = sc.acquire()
    with pytest.raises(api.SsnError, match="ssn_reads_reserve"):
        sc.reads_view(b)
    sc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n_slots", [3, 6])
def test_device_pileup_pipeline_of_batches(built, n_slots):
    """Several read batches in flight on their slot streams; results come back in submission order.  Up to 4 slots the
    scoring kernels are queued with the pileup (from the device-side column count), beyond that when the batch is drained."""
    from somatic_sniper_b200 import api, hostapi
    batches = []
    for k in range(5):
        ref, t, n = make_case(seed=30 + k, length=6000 + 500 * k, depth_t=20, depth_n=12, complex_cigars=bool(k & 1))
        rw = R.build_read_windows(t, n, ref)
        batches.append((rw, hostapi.pack_read_windows(rw)))
    sc = api.Scorer(n_slots=n_slots, max_columns=2 * 8192, max_entries=1 << 20)
    sc.reads_reserve(max(rw["recs"].size for rw, _ in batches) + 64, max(rw["rec_off"].size for rw, _ in batches) + 1, 2)
    pending, done = [], []
    for rw, want in batches:
        if len(pending) == min(n_slots, 3):
            sites, _ = sc.results_next()
            b0, want0 = pending.pop(0)
            same_columns(sc.download_columns(b0), want0)
            done.append(sites)
            sc.release(b0)
        b = sc.acquire()
        sc.submit_reads(b, *sc.fill_reads(b, rw))
        pending.append((b, want))
    while pending:
        sites, _ = sc.results_next()
        b0, want0 = pending.pop(0)
        same_columns(sc.download_columns(b0), want0)
        done.append(sites)
        sc.release(b0)
    for (rw, want), sites in zip(batches, done):
        host_sites, _ = sc.score(dict(want, ref_char=NT16_CHARS[want["ref_nt16"]]))
        assert sites.tobytes() == host_sites.tobytes()
    sc.close()
