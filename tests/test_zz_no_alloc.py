"""fxg_count_device_noalloc: the allocation-free literal count (scratch tables in a file-scope __device__ array of
libfxgpu.so instead of cudaMalloc'ed memory) against the existing device-resident entry point
(fxg_scan_device_async / _finish) on the same inputs.  Integer work: the counts must be EQUAL.

Collected last on purpose (the file name sorts behind every other test file): a
failure here must not hide the other tests.  Passes alone in a fresh process and after the rest of the suite."""
import ctypes as C
import random

import pytest

import helpers

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fx(built):
    import fxtract_b200
    return fxtract_b200


def _device_copy(torch, data):
    """input batch in HBM: 16-byte aligned (torch allocations are), readable 64 bytes past its end"""
    t = torch.zeros(len(data) + 64, dtype=torch.uint8, device="cuda:0")
    t[:len(data)] = torch.frombuffer(bytearray(data), dtype=torch.uint8).to("cuda:0")
    return t


def _existing(ctx, t, n):
    ctx.scan_device_async(t.data_ptr(), n)
    r = ctx.scan_device_finish()
    return dict(n_lines=r.n_lines, n_records=r.n_records, n_matched=r.n_matched, violation=r.violation,
                canonical_ok=r.canonical_ok, first_bad=r.first_bad_offset)


def _noalloc(torch, lib, ctx, t, n, out):
    out.fill_(0x5a)                                   # stale contents must not survive
    torch.cuda.synchronize()
    rc = ctx.count_device_noalloc(t.data_ptr(), n, out.data_ptr())
    assert rc == 0, ctx.last_error()
    torch.cuda.synchronize()                          # the call itself never synchronises
    c = lib.NoAllocCounts.from_buffer_copy(bytes(out.cpu().numpy().tobytes()))
    return dict(n_lines=c.n_lines, n_records=c.n_records, n_matched=c.n_matched, violation=c.violation,
                canonical_ok=c.canonical_ok, first_bad=c.first_bad_offset)


def test_noalloc_count_equals_existing_entry_point(fx):
    import torch
    from fxtract_b200 import lib
    rng = random.Random(2024)
    big = helpers.make_fastq(rng, 28000, seqlen=(60, 200))          # ~7.6 MiB: near the limit, ~300 tiles
    small = helpers.make_fastq(rng, 700, seqlen=(20, 90))
    assert len(big) <= lib.NOALLOC_MAX_BYTES
    dirty = small[:5000] + b"@x\nAC\x7fGT\n+\nIIIII\n" + small[5000:]   # rejected by both (and cut mid-record)
    out = torch.zeros(C.sizeof(lib.NoAllocCounts), dtype=torch.uint8, device="cuda:0")
    for plen, target in ((20, 0), (6, 0), (3, 1), (0, 0)):
        pat = helpers.pick_substring(rng, big, lo=plen, hi=plen) if plen else b""
        if target == 1:
            pat = b"r1"
        ctx = fx.Context(fx.MODE_LITERAL, target=target, count_only=True)
        ctx.set_patterns([pat])
        # two calls in a row on the shared scratch: a large batch, then a smaller one with other counts (everything
        # the first call left behind must be cleared or overwritten), then the large one again
        for data in (big, small, big, dirty, small):
            t = _device_copy(torch, data)
            want = _existing(ctx, t, len(data))
            got = _noalloc(torch, lib, ctx, t, len(data), out)
            assert got == want, (plen, target, len(data), got, want)
        assert want["canonical_ok"] == 1 and want["n_records"] == 700
        if plen in (0, 3):
            assert want["n_matched"] > 0
        ctx.close()


def test_noalloc_refuses_what_it_cannot_serve(fx):
    import torch
    from fxtract_b200 import lib
    out = torch.zeros(C.sizeof(lib.NoAllocCounts), dtype=torch.uint8, device="cuda:0")
    t = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda:0")
    ctx = fx.Context(fx.MODE_LITERAL, count_only=True)
    ctx.set_patterns([b"ACGTACGTAC"])
    # beyond the stated maximum: refused with a clear error before anything is enqueued (the pointer is never read)
    rc = ctx.count_device_noalloc(t.data_ptr(), lib.NOALLOC_MAX_BYTES + 1, out.data_ptr())
    assert rc == -1 and "FXG_NOALLOC_MAX_BYTES" in ctx.last_error(), (rc, ctx.last_error())
    rc = ctx.count_device_noalloc(t.data_ptr() + 4, 1024, out.data_ptr())
    assert rc == -1 and "aligned" in ctx.last_error()
    # an empty batch is fine: zero counts
    got = _noalloc(torch, lib, ctx, t, 0, out)
    assert got["n_lines"] == 0 and got["n_records"] == 0 and got["canonical_ok"] == 1
    ctx.close()
    multi = fx.Context(fx.MODE_MULTI, count_only=True)
    multi.set_patterns([b"ACGTACGTACGTACGTACGT", b"TTTTTTTTTTTTTTTTTTTT"])
    rc = multi.count_device_noalloc(t.data_ptr(), 1024, out.data_ptr())
    assert rc == -5 and "LITERAL" in multi.last_error()
    multi.close()
