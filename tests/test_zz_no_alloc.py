"""ssn_inflate_static (include/ssn.h): the inflate kernel with its scratch — block table and work counter — in a file-scope
__device__ array, for callers that cannot allocate.  Checked through the C ABI against the existing entry point
(ssn_inflater_run) byte for byte, on two calls in a row with different batches (the array keeps the first call's contents: a
stale counter or table would show in the second), plus the size limit.  Collected last on purpose (pytest -x)."""
import zlib

import numpy as np
import pytest


def _deflate(data, level):
    c = zlib.compressobj(level, zlib.DEFLATED, -15)
    return c.compress(data) + c.flush()


def _batch(seed, n):
    rng = np.random.default_rng(seed)
    payloads, sizes = [], []
    for k in range(n):
        size = int(rng.integers(1, 65281))
        kind = k % 4
        if kind == 0:
            raw = rng.choice(np.array([2, 12, 23, 37], np.uint8), size, p=[0.02, 0.05, 0.13, 0.8]).tobytes()
        elif kind == 1:
            raw = (rng.integers(0, 256, int(rng.integers(1, 300)), dtype=np.uint8).tobytes() * (size + 1))[:size]
        elif kind == 2:
            raw = rng.integers(0, 256, size, dtype=np.uint8).tobytes()
        else:
            raw = rng.choice(np.frombuffer(b"ACGTN=\t\n0123456789", np.uint8), size).tobytes()
        payloads.append(_deflate(raw, (0, 1, 6, 9)[(k // 4) % 4])); sizes.append(size)
    # two damaged streams: both entry points must refuse (or accept) them alike and leave the neighbours alone
    payloads[5] = payloads[5][:max(1, len(payloads[5]) // 2)]
    b = bytearray(payloads[9]); b[len(b) // 3] ^= 0x10; payloads[9] = bytes(b)
    return payloads, sizes


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_the_lane_service(built):
    import torch
    from somatic_sniper_b200 import api
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    inf = api.Inflater(n_lanes=1, in_cap=24 << 20, out_cap=24 << 20, max_blocks=512)
    # the caller's buffers (this test's, made once, before any call of the entry point under test)
    cap = 24 << 20
    d_in = torch.zeros(cap + 64, dtype=torch.uint8, device=dev)
    d_out = torch.zeros(cap, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(api.INFLATE_STATIC_MAX_BLOCKS, dtype=torch.int32, device=dev)
    h_in = np.zeros(cap, np.uint8)
    stream = torch.cuda.Stream(device=dev)
    try:
        for call, (seed, n) in enumerate(((101, 331), (202, 77))):          # second batch smaller: a stale table tail must not matter
            payloads, sizes = _batch(seed, n)
            want = inf.inflate_many(payloads, sizes)                          # existing entry point: bytes, or None where refused
            blk, a, b = [], 0, 0
            for p, s in zip(payloads, sizes):
                h_in[a:a + len(p)] = np.frombuffer(p, np.uint8)
                blk.append((a, len(p), b, s)); a += len(p); b += s
            d_in[:a].copy_(torch.from_numpy(h_in[:a]))
            d_out.fill_(0xEE); d_status.fill_(-1)
            torch.cuda.synchronize()
            api.inflate_static(0, d_in.data_ptr(), a, np.array(blk, api.INFLATE_BLOCK_DTYPE), d_out.data_ptr(), b, d_status.data_ptr(),
                               stream.cuda_stream)
            stream.synchronize()
            status = d_status[:n].cpu().numpy()
            got = d_out[:b].cpu().numpy()
            assert (status >= 0).all(), "call %d: blocks the kernel never reached" % call
            assert [w is None for w in want] == [bool(x) for x in status], "call %d: the two entry points refuse different blocks" % call
            assert sum(w is None for w in want) >= 1
            for (ao, al, bo, bl), w in zip(blk, want):
                if w is not None:
                    assert got[bo:bo + bl].tobytes() == w, "call %d: block at %d differs" % (call, bo)
        # beyond the stated maximum: refused with a message that names the limit, nothing launched
        too_many = np.zeros(api.INFLATE_STATIC_MAX_BLOCKS + 1, api.INFLATE_BLOCK_DTYPE)
        with pytest.raises(Exception) as e:
            api.inflate_static(0, d_in.data_ptr(), 0, too_many, d_out.data_ptr(), 0, d_status.data_ptr(), stream.cuda_stream)
        assert "4096" in str(e.value) and "SSN_INFLATE_STATIC_MAX_BLOCKS" in str(e.value)
        # a block outside the batch is refused on the host
        with pytest.raises(Exception):
            api.inflate_static(0, d_in.data_ptr(), 10, np.array([(0, 11, 0, 5)], api.INFLATE_BLOCK_DTYPE), d_out.data_ptr(), 5, d_status.data_ptr(), 0)
    finally:
        inf.close()
