"""clrt_scan_inclusive_u32_noalloc (include/clrt.h): the scan for callers that cannot allocate — descriptors in a static device array
inside the library, no context, the caller's stream.  Checked through the C ABI against the existing entry point
clrt_scan_inclusive_u32 on the same inputs: integer work, so the outputs must be bit-identical (the tolerance of every scan test here).

This file is collected LAST on purpose: it passes alone in a fresh process and after the other tests
(its first call does the per-device set-up whatever ran before)."""
import ctypes as C
import numpy as np
import pytest
from openclraytracer_b200 import api, scenes, _lib

pytestmark = pytest.mark.gpu
MAX_ELEMENTS = 1 << 28          # CLRT_SCAN_NOALLOC_MAX_ELEMENTS


def values(n, seed):
    return (scenes.splitmix64(np.arange(n, dtype=np.uint64) + np.uint64(seed)) & np.uint64(7)).astype(np.uint32)


def scan_both(ctx, lib, a, in_place=False, offset_words=0):
    """(existing entry point's output, no-alloc entry point's output) for the same input; buffers come from the CALLER (api.Buffer)."""
    n = len(a)
    pad = np.zeros(n + offset_words, dtype=np.uint32); pad[offset_words:] = a
    src = api.Buffer.from_array(ctx, pad) if n + offset_words else api.Buffer(ctx, 16)
    want_b = api.Buffer(ctx, max(pad.nbytes, 16)); got_b = api.Buffer(ctx, max(pad.nbytes, 16))
    off = 4 * offset_words
    api.check(lib.clrt_scan_inclusive_u32(ctx.handle, C.c_void_p(src.getCLMem() + off), C.c_void_p(want_b.getCLMem() + off), n))
    want = want_b.copyToHost(np.uint32, n + offset_words)[offset_words:]
    if in_place:
        got_b.upload(pad)
        api.check(lib.clrt_scan_inclusive_u32_noalloc(C.c_void_p(ctx.stream), C.c_void_p(got_b.getCLMem() + off), C.c_void_p(got_b.getCLMem() + off), n))
    else:
        api.check(lib.clrt_scan_inclusive_u32_noalloc(C.c_void_p(ctx.stream), C.c_void_p(src.getCLMem() + off), C.c_void_p(got_b.getCLMem() + off), n))
    ctx.finish()
    got = got_b.copyToHost(np.uint32, n + offset_words)[offset_words:]
    for b in (src, want_b, got_b):
        b.free()
    return want, got


def test_noalloc_scan_equals_existing_entry_point_two_calls_in_a_row(ctx):
    lib = _lib.load()
    # first call: 16K-element tiles, several waves, ragged last tile; second call right after it: smaller, other data, 8K-element tiles —
    # the descriptors the first call left in the shared static array must not leak into it
    a1 = values(16384 * 148 * 3 + 12345, seed=1)
    want1, got1 = scan_both(ctx, lib, a1)
    assert np.array_equal(got1, want1)
    assert got1[-1] == np.uint32(a1.sum(dtype=np.uint64) % (1 << 32))
    a2 = values(100_003, seed=2)
    want2, got2 = scan_both(ctx, lib, a2)
    assert np.array_equal(got2, want2)
    # and back to the big size with the first data: same answer as the first time
    want3, got3 = scan_both(ctx, lib, a1)
    assert np.array_equal(got3, want1) and np.array_equal(want3, want1)


@pytest.mark.parametrize("n", [1, 31, 8192, 8193, (1 << 21) - 1, 1 << 21, 3_000_001])
def test_noalloc_scan_sizes_in_place_and_unaligned(ctx, n):
    lib = _lib.load()
    a = values(n, seed=n)
    want, got = scan_both(ctx, lib, a, in_place=True)
    assert np.array_equal(got, want)
    want, got = scan_both(ctx, lib, a, offset_words=3)            # 4-byte aligned only: the look-back kernel
    assert np.array_equal(got, want)


def test_noalloc_scan_same_stream_back_to_back_without_host_sync(ctx):
    """Calls ordered on one stream share the descriptor array safely: five scans enqueued with no synchronisation in between."""
    lib = _lib.load()
    sizes = [2_500_000, 70_000, 2_500_000, 1 << 21, 9_000]
    ins = [values(n, seed=10 + i) for i, n in enumerate(sizes)]
    bufs = [api.Buffer.from_array(ctx, a) for a in ins]
    for b, n in zip(bufs, sizes):
        api.check(lib.clrt_scan_inclusive_u32_noalloc(C.c_void_p(ctx.stream), C.c_void_p(b.getCLMem()), C.c_void_p(b.getCLMem()), n))
    ctx.finish()
    for b, a, n in zip(bufs, ins, sizes):
        assert np.array_equal(b.copyToHost(np.uint32, n), np.cumsum(a, dtype=np.uint64).astype(np.uint32))
        b.free()


def test_noalloc_scan_rejects_sizes_beyond_the_static_array(ctx):
    lib = _lib.load()
    b = api.Buffer(ctx, 64)
    with pytest.raises(api.ClrtError, match="CLRT_SCAN_NOALLOC_MAX_ELEMENTS"):
        api.check(lib.clrt_scan_inclusive_u32_noalloc(C.c_void_p(ctx.stream), C.c_void_p(b.getCLMem()), C.c_void_p(b.getCLMem()), MAX_ELEMENTS + 1))
    with pytest.raises(api.ClrtError, match="NULL buffer"):
        api.check(lib.clrt_scan_inclusive_u32_noalloc(C.c_void_p(ctx.stream), None, C.c_void_p(b.getCLMem()), 16))
    assert lib.clrt_scan_inclusive_u32_noalloc(C.c_void_p(ctx.stream), None, None, 0) == 0          # n == 0: nothing to do
    b.free()


def test_noalloc_scan_at_the_stated_maximum(ctx):
    """n = CLRT_SCAN_NOALLOC_MAX_ELEMENTS uses every descriptor of the static array: scan of ones = 1, 2, 3, ..."""
    lib = _lib.load()
    n = MAX_ELEMENTS
    b = api.Buffer.from_array(ctx, np.ones(n, dtype=np.uint32))
    api.check(lib.clrt_scan_inclusive_u32_noalloc(C.c_void_p(ctx.stream), C.c_void_p(b.getCLMem()), C.c_void_p(b.getCLMem()), n))
    ctx.finish()
    got = b.copyToHost(np.uint32, n)
    b.free()
    assert got[0] == 1 and got[-1] == n
    assert (np.diff(got[:: 4099].astype(np.int64)) == 4099).all()
    assert np.array_equal(got[-70_000:], np.arange(n - 70_000 + 1, n + 1, dtype=np.uint32))
