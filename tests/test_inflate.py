"""CPU tests of the BGZF block decoder (csrc/host/ssn_inflate.c) against zlib: every DEFLATE block type and coding
strategy, matches at every short distance, block-size edge cases, whole BAM files block by block, and damaged streams
(which must be refused, never crash: the readers then fall back to zlib)."""
import ctypes as C
import os
import struct
import zlib

import numpy as np
import pytest

SLACK = 32


@pytest.fixture(scope="module")
def inflate(built):
    from somatic_sniper_b200 import hostapi
    lib = hostapi.load_library()
    lib.ssn_inflate_raw.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    lib.ssn_inflate_raw.restype = C.c_int

    def run(comp, ulen):
        src = C.create_string_buffer(comp + b"\xa5" * SLACK, len(comp) + SLACK)
        dst = C.create_string_buffer(ulen + 16)
        dst.raw = b"\x5a" * (ulen + 16)
        rc = lib.ssn_inflate_raw(src, len(comp), dst, ulen)
        assert dst.raw[ulen:] == b"\x5a" * 16, "wrote past the end of the output"
        return rc, dst.raw[:ulen]
    return run


def _deflate(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, mem=8):
    c = zlib.compressobj(level, zlib.DEFLATED, -15, mem, strategy)
    return c.compress(data) + c.flush()


def _samples():
    rng = np.random.default_rng(12)
    out = {"empty": b"", "one": b"x", "two": b"ab", "zeros": bytes(65280), "random": rng.integers(0, 256, 65280, dtype=np.uint8).tobytes()}
    for period in (1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 16, 17, 255, 258, 259):
        out["period%d" % period] = (rng.integers(0, 256, period, dtype=np.uint8).tobytes() * (70000 // period + 1))[:65280]
    out["quals"] = rng.choice(np.array([2, 12, 23, 37], np.uint8), 65280, p=[0.02, 0.05, 0.13, 0.8]).tobytes()
    out["nibbles"] = rng.choice(np.array([0x11, 0x12, 0x14, 0x18, 0x21, 0x22, 0x24, 0x28, 0x41, 0x42, 0x44, 0x48, 0x81, 0x82, 0x84, 0x88], np.uint8), 65280).tobytes()
    text = (b"@SQ\tSN:chr%d\tLN:%d\n" * 1)
    out["text"] = b"".join(text % (i % 25, 1000 * i) for i in range(2500))[:65280]
    out["skewed"] = rng.choice(np.arange(256, dtype=np.uint8), 65280, p=np.r_[[0.5], np.full(255, 0.5 / 255)]).tobytes()   # long and short codes
    out["mixed"] = out["quals"][:20000] + out["random"][:10000] + out["zeros"][:15000] + out["text"][:20000]
    return out


SAMPLES = _samples()


@pytest.mark.parametrize("name", sorted(SAMPLES))
def test_matches_zlib_on_every_block_type_and_strategy(inflate, name):
    data = SAMPLES[name]
    for level in (0, 1, 4, 6, 9):
        for strategy in (zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE, zlib.Z_FILTERED):
            for mem in (1, 8):
                comp = _deflate(data, level, strategy, mem)
                rc, got = inflate(comp, len(data))
                assert rc == 0 and got == data, (name, level, strategy, mem)


def test_several_blocks_in_one_stream_and_sync_flushes(inflate):
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    parts = [SAMPLES["quals"][:9000], SAMPLES["random"][:3000], b"", SAMPLES["text"][:20000], SAMPLES["zeros"][:100]]
    comp = b""
    for i, p in enumerate(parts):
        comp += c.compress(p) + c.flush(zlib.Z_FULL_FLUSH if i % 2 else zlib.Z_SYNC_FLUSH)     # empty stored blocks in between
    comp += c.flush()
    data = b"".join(parts)
    rc, got = inflate(comp, len(data))
    assert rc == 0 and got == data


def test_damaged_streams_are_refused(inflate):
    rng = np.random.default_rng(5)
    data = SAMPLES["mixed"]
    comp = _deflate(data, 6)
    rc, _ = inflate(comp, len(data) - 1)                 # stated size too small
    assert rc != 0
    rc, _ = inflate(comp, len(data) + 1)                 # ... too large
    assert rc != 0
    for cut in (1, 2, 3, 10, len(comp) // 2, len(comp) - 1):
        rc, _ = inflate(comp[:cut], len(data))           # truncated
        assert rc != 0
    refused = 0
    for _ in range(300):                                 # flipped bits: refused or, rarely, still a valid stream of the same size
        b = bytearray(comp)
        for k in rng.integers(0, len(b), int(rng.integers(1, 4))):
            b[k] ^= 1 << int(rng.integers(0, 8))
        rc, got = inflate(bytes(b), len(data))
        if rc != 0:
            refused += 1
        else:
            assert zlib.decompress(bytes(b), -15) == got
    assert refused > 50
    for _ in range(200):                                 # random bytes as a stream
        junk = rng.integers(0, 256, int(rng.integers(1, 400)), dtype=np.uint8).tobytes()
        rc, got = inflate(junk, 1000)
        if rc == 0:
            assert zlib.decompress(junk, -15)[:1000] == got


def test_every_block_of_the_integration_and_simulated_bams(inflate, tmp_path):
    from somatic_sniper_b200 import cli_tools
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pair = cli_tools.make_pair(str(tmp_path), [("chr1", 20000)], 20, 10, seed=3)
    fast = cli_tools.make_pair_fast(str(tmp_path / "fast"), [("chr1", 200000)], 30, 15, seed=4, level=6)
    n = 0
    for path in (os.path.join(root, "tests", "golden", "integration", "t-small.bam"), pair["tumor"], pair["normal"], fast["tumor"]):
        raw = open(path, "rb").read()
        at = 0
        while at < len(raw):
            xlen = struct.unpack_from("<H", raw, at + 10)[0]
            bsize = struct.unpack_from("<H", raw, at + 16)[0] + 1
            comp = raw[at + 12 + xlen:at + bsize - 8]
            crc, isize = struct.unpack_from("<II", raw, at + bsize - 8)
            rc, got = inflate(comp, isize)
            assert rc == 0 and zlib.crc32(got) & 0xffffffff == crc
            at += bsize
            n += 1
    assert n > 100


# ---- the DEVICE decoder's source (csrc/ssn_inflate_dev.cuh), compiled for the host with its lane loops as real loops ----

@pytest.fixture(scope="module")
def inflate_dev(built):
    from somatic_sniper_b200 import api
    lib = api.load_library()
    lib.ssn_inflate_dev_host_shift.argtypes = [C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32, C.c_uint32]
    lib.ssn_inflate_dev_host_shift.restype = C.c_int

    def run(comp, ulen, shift=0):
        dst = C.create_string_buffer(ulen + 16)
        dst.raw = b"\x5a" * (ulen + 16)
        rc = lib.ssn_inflate_dev_host_shift(comp, len(comp), dst, ulen, shift)
        assert dst.raw[ulen:] == b"\x5a" * 16, "wrote past the end of the output"
        return rc, dst.raw[:ulen]
    return run


@pytest.mark.parametrize("name", sorted(SAMPLES))
def test_device_decoder_source_matches_zlib(inflate_dev, name):
    data = SAMPLES[name]
    k = 0
    for level in (0, 1, 6, 9):
        for strategy in (zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE):
            comp = _deflate(data, level, strategy, 8 if k % 2 else 1)
            rc, got = inflate_dev(comp, len(data), shift=k % 4)
            k += 1
            assert rc == 0 and got == data, (name, level, strategy, rc)


def test_device_decoder_source_on_bam_blocks_and_damage(inflate_dev, tmp_path):
    from somatic_sniper_b200 import cli_tools
    fast = cli_tools.make_pair_fast(str(tmp_path / "fast"), [("chr1", 150000)], 30, 15, seed=4, level=6)
    raw = open(fast["tumor"], "rb").read()
    at, n, first = 0, 0, None
    while at < len(raw):
        xlen = struct.unpack_from("<H", raw, at + 10)[0]
        bsize = struct.unpack_from("<H", raw, at + 16)[0] + 1
        comp = raw[at + 12 + xlen:at + bsize - 8]
        crc, isize = struct.unpack_from("<II", raw, at + bsize - 8)
        rc, got = inflate_dev(comp, isize, shift=n % 4)
        assert rc == 0 and zlib.crc32(got) & 0xffffffff == crc
        if first is None and isize > 30000:
            first = (comp, isize, got)
        at += bsize
        n += 1
    assert n > 20
    comp, isize, data = first
    assert inflate_dev(comp, isize - 1)[0] != 0 and inflate_dev(comp, isize + 1)[0] != 0
    for cut in (1, 2, 7, len(comp) // 2, len(comp) - 1):
        assert inflate_dev(comp[:cut], isize)[0] != 0
    rng = np.random.default_rng(8)
    for _ in range(200):
        b = bytearray(comp)
        for j in rng.integers(0, len(b), int(rng.integers(1, 4))):
            b[j] ^= 1 << int(rng.integers(0, 8))
        rc, got = inflate_dev(bytes(b), isize)
        if rc == 0:
            assert zlib.decompress(bytes(b), -15) == got


# ---- the kernel itself ------------------------------------------------------------------------------------------------

@pytest.mark.gpu
def test_gpu_inflater_matches_zlib(built, tmp_path):
    """ssn_inflate_kernel through the lane service: every sample x level x strategy in one batch, whole BAM files block
    by block, damaged streams refused (flagged for the host) without disturbing their neighbours."""
    from somatic_sniper_b200 import api, cli_tools
    inf = api.Inflater(n_lanes=2, in_cap=8 << 20, out_cap=24 << 20, max_blocks=512)
    try:
        payloads, want = [], []
        for name in sorted(SAMPLES):
            for level in (0, 1, 6, 9):
                for strategy in (zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE):
                    payloads.append(_deflate(SAMPLES[name], level, strategy))
                    want.append(SAMPLES[name])
        got = inf.inflate_many(payloads, [len(w) for w in want])
        assert all(g == w for g, w in zip(got, want)), [i for i, (g, w) in enumerate(zip(got, want)) if g != w][:5]
        fast = cli_tools.make_pair_fast(str(tmp_path), [("chr1", 1500000)], 60, 30, seed=4, level=6)
        for path in (fast["tumor"], fast["normal"]):
            raw = open(path, "rb").read()
            at, pl, sz, crc = 0, [], [], []
            while at < len(raw):
                xlen = struct.unpack_from("<H", raw, at + 10)[0]
                bsize = struct.unpack_from("<H", raw, at + 16)[0] + 1
                c, n = struct.unpack_from("<II", raw, at + bsize - 8)
                if n:
                    pl.append(raw[at + 12 + xlen:at + bsize - 8]); sz.append(n); crc.append(c)
                at += bsize
            got = inf.inflate_many(pl, sz)
            assert len(got) > 500 and all(g is not None and zlib.crc32(g) & 0xffffffff == c for g, c in zip(got, crc))
        # damage: flipped bits, truncation, wrong sizes -- refused or (rarely) still a valid stream; neighbours untouched
        rng = np.random.default_rng(3)
        good = _deflate(SAMPLES["mixed"], 6)
        batch, sizes, kind = [], [], []
        for k in range(120):
            b = bytearray(good)
            if k % 3 == 0:
                for j in rng.integers(0, len(b), 2):
                    b[j] ^= 1 << int(rng.integers(0, 8))
                batch.append(bytes(b)); sizes.append(len(SAMPLES["mixed"])); kind.append("flip")
            elif k % 3 == 1:
                batch.append(bytes(b[:int(rng.integers(1, len(b) - 1))])); sizes.append(len(SAMPLES["mixed"])); kind.append("cut")
            else:
                batch.append(good); sizes.append(len(SAMPLES["mixed"])); kind.append("good")
        got = inf.inflate_many(batch, sizes)
        for g, c, k in zip(got, batch, kind):
            if k == "good":
                assert g == SAMPLES["mixed"]
            elif k == "cut":
                assert g is None
            elif g is not None:
                assert zlib.decompress(c, -15) == g
    finally:
        inf.close()
