"""GPU parity: the Sort and Prefix Sum building blocks through the C ABI vs the oracle (bit-exact)."""
import numpy as np
import pytest
from openclraytracer_b200 import api, scenes, layouts as L

pytestmark = pytest.mark.gpu


def make_pairs(n, seed, key_mask=0x3FFFFFFF, dup=False):
    i = np.arange(n, dtype=np.uint64)
    k = (scenes.splitmix64(i + np.uint64(seed) * np.uint64(1000003)) & np.uint64(key_mask)).astype(np.uint32)
    if dup:                                   # "mesh-like": ~50 % adjacent duplicates
        k = (k >> np.uint32(1)) << np.uint32(1)
        k[1::2] = k[0::2][: len(k[1::2])]
    p = np.zeros(n, dtype=L.UINT2)
    p["x"], p["y"] = k, np.arange(n, dtype=np.uint32)
    return p


@pytest.mark.parametrize("n", [0, 1, 2, 31, 32, 33, 255, 1024, 1025, 4096, 50032, (1 << 19) + 7, 1 << 20, 3_000_001])
def test_radix_sort_matches_stable_oracle(ctx, O, n):
    p = make_pairs(n, seed=n, key_mask=0xFFFFFFFF if n % 2 else 0x3FFFFFFF, dup=(n % 3 == 0))
    buf = api.Buffer(ctx, max(p.nbytes, 16))
    if n:
        buf.upload(p)
    api.BitonicSort(ctx).sort(buf.getCLMem(), n)
    got = buf.copyToHost(L.UINT2, n)
    assert np.array_equal(got, O.sort_pairs_stable(p))
    buf.free()


@pytest.mark.parametrize("n", [1, 2, 1001, 70_000])
def test_radix_sort_of_a_slice_at_an_odd_pair_offset(ctx, O, n):
    """The interface asks for 8-byte alignment (a uint2); a slice that starts at an odd pair is not 16-byte aligned."""
    p = make_pairs(n + 3, seed=n)
    buf = api.Buffer.from_array(ctx, p)
    api.BitonicSort(ctx).sort(buf.getCLMem() + 8, n)
    got = buf.copyToHost(L.UINT2, n + 3)
    assert np.array_equal(got[1:1 + n], O.sort_pairs_stable(p[1:1 + n]))
    assert np.array_equal(got[:1], p[:1]) and np.array_equal(got[1 + n:], p[1 + n:])
    buf.free()


def test_radix_sort_few_distinct_keys_and_extremes(ctx, O):
    n = 200_000
    p = make_pairs(n, 5)
    p["x"] = np.where(np.arange(n) % 3 == 0, 0xFFFFFFFF, np.where(np.arange(n) % 3 == 1, 0, 0x80000000)).astype(np.uint32)
    buf = api.Buffer.from_array(ctx, p)
    api.BitonicSort(ctx).sort(buf.getCLMem(), n)
    assert np.array_equal(buf.copyToHost(L.UINT2, n), O.sort_pairs_stable(p))
    buf.free()


@pytest.mark.parametrize("k", [1, 5, 10, 14])
def test_bitonic_compat_reproduces_reference_network(ctx, O, k):
    """CLRT_SORT_BITONIC_COMPAT == the reference's unstable network, tie order included."""
    n = 1 << k
    p = make_pairs(n, seed=k, key_mask=0xFF)          # heavy duplication
    buf = api.Buffer.from_array(ctx, p)
    api.BitonicSort(ctx, mode=api.SORT_BITONIC_COMPAT).sort(buf.getCLMem(), n)
    assert np.array_equal(buf.copyToHost(L.UINT2, n), O.sort_pairs_bitonic(p))
    buf.free()


def test_bitonic_compat_rejects_non_pow2(ctx):
    buf = api.Buffer(ctx, 800)
    with pytest.raises(api.ClrtError, match="power of two"):
        api.BitonicSort(ctx, mode=api.SORT_BITONIC_COMPAT).sort(buf.getCLMem(), 100)
    buf.free()


# sizes around the tile sizes of the staged kernel (8192 below 2^21 elements, 16384 from there), one tile, one wave of 148 CTAs, a few
# waves with a ragged last tile, and the switch of tile size
@pytest.mark.parametrize("n", [0, 1, 2, 31, 4095, 4096, 4097, 8191, 8192, 8193, 16385, 65536, 1_000_003, 8192 * 296 + 17, (1 << 21) - 1, 1 << 21,
                               (1 << 21) + 1, 16384 * 148 + 1, 16384 * 148 * 3 + 5, 1 << 24])
def test_scan_matches_oracle(ctx, O, n):
    a = (scenes.splitmix64(np.arange(n, dtype=np.uint64)) & np.uint64(7)).astype(np.uint32)
    bi = api.Buffer(ctx, max(a.nbytes, 16)); bo = api.Buffer(ctx, max(a.nbytes, 16))
    if n:
        bi.upload(a)
    ps = api.PrefixSum(ctx)
    ps.computePrefixSum(bi.getCLMem(), bo.getCLMem(), n)
    want = O.scan_inclusive(a)
    assert np.array_equal(bo.copyToHost(np.uint32, n), want)
    ps.computePrefixSum(bi.getCLMem(), bi.getCLMem(), n)           # in place
    assert np.array_equal(bi.copyToHost(np.uint32, n), want)
    if n:
        assert want[-1] == a.sum(dtype=np.uint64) % (1 << 32)
    bi.free(); bo.free()


@pytest.mark.parametrize("n", [5, 8192 * 3 + 1, 3_000_001])
def test_scan_buffers_aligned_to_4_bytes_only(ctx, O, n):
    """Odd offsets take the register-prefetch kernel with the classic look-back (scan.cu kernel 1); in place and out of place."""
    a = (scenes.splitmix64(np.arange(n, dtype=np.uint64)) & np.uint64(15)).astype(np.uint32)
    want = O.scan_inclusive(a)
    bi = api.Buffer(ctx, a.nbytes + 32); bo = api.Buffer(ctx, a.nbytes + 32)
    pad = np.zeros(n + 3, dtype=np.uint32); pad[3:] = a
    bi.upload(pad)
    ps = api.PrefixSum(ctx)
    ps.computePrefixSum(bi.getCLMem() + 12, bo.getCLMem() + 4, n)
    assert np.array_equal(bo.copyToHost(np.uint32, n + 1)[1:], want)
    ps.computePrefixSum(bi.getCLMem() + 12, bi.getCLMem() + 12, n)
    assert np.array_equal(bi.copyToHost(np.uint32, n + 3)[3:], want)
    bi.free(); bo.free()


def test_scan_repeated_calls_reuse_descriptors(ctx, O):
    """Descriptors are never reset: a call is told from the previous ones by its generation number.  Alternating sizes and tile sizes
    on one context must not see stale aggregates."""
    ps = api.PrefixSum(ctx)
    for rep, n in enumerate([3_000_000, 100_000, 3_000_000, 2_500_000, 16384 * 148 * 2, 3_000_000]):
        a = (scenes.splitmix64(np.arange(n, dtype=np.uint64) + np.uint64(rep)) & np.uint64(7)).astype(np.uint32)
        b = api.Buffer.from_array(ctx, a)
        ps.computePrefixSum(b.getCLMem(), b.getCLMem(), n)
        assert np.array_equal(b.copyToHost(np.uint32, n), O.scan_inclusive(a)), (rep, n)
        b.free()


def test_scan_full_size_property(ctx):
    """BASELINE config 5's largest size, 2^28 elements: scan of ones = 1, 2, 3, ... (size-independent property)."""
    n = 1 << 28
    b = api.Buffer.from_array(ctx, np.ones(n, dtype=np.uint32))
    api.PrefixSum(ctx).computePrefixSum(b.getCLMem(), b.getCLMem(), n)
    got = b.copyToHost(np.uint32, n)
    b.free()
    assert got[0] == 1 and got[-1] == n
    assert (np.diff(got[:: 4099].astype(np.int64)) == 4099).all()
    assert np.array_equal(got[-70_000:], np.arange(n - 70_000 + 1, n + 1, dtype=np.uint32))
    assert np.array_equal(got[: 70_000], np.arange(1, 70_001, dtype=np.uint32))
    mid = 148 * 16384 * 37                                              # a wave boundary
    assert np.array_equal(got[mid - 20_000: mid + 20_000], np.arange(mid - 20_000 + 1, mid + 20_000 + 1, dtype=np.uint32))


def test_scan_wraps_like_uint32(ctx, O):
    a = np.full(100_000, 0xFFFF0000, dtype=np.uint32)
    b = api.Buffer.from_array(ctx, a)
    api.PrefixSum(ctx).computePrefixSum(b.getCLMem(), b.getCLMem(), len(a))
    assert np.array_equal(b.copyToHost(np.uint32, len(a)), O.scan_inclusive(a))
    b.free()


def test_sort_full_size_properties(ctx):
    """BASELINE config 5 size (2^26 here to bound memory/time): sortedness, stability, permutation — checked with numpy."""
    n = 1 << 26
    p = make_pairs(n, seed=11)
    buf = api.Buffer.from_array(ctx, p)
    api.BitonicSort(ctx).sort(buf.getCLMem(), n)
    got = buf.copyToHost(L.UINT2, n)
    buf.free()
    k, v = got["x"], got["y"]
    assert (k[1:] >= k[:-1]).all()
    assert (v[1:][k[1:] == k[:-1]] > v[:-1][k[1:] == k[:-1]]).all()       # stable: equal keys keep input order
    assert np.array_equal(p["x"][v], k)                                    # (key,value) pairs stay together
    seen = np.zeros(n, dtype=np.uint8); seen[v] = 1
    assert seen.all()                                                      # permutation


def test_buffer_semantics(ctx):
    """CLBuffer: grow-only resize, fill patterns, offset copies (R/src/OpenCLUtils/CLBuffer.cpp:40-130)."""
    b = api.Buffer(ctx, 1024)
    assert b.getSize() == 1024 and b.getActualSize() >= 1024
    b.fill(np.array([0xFFFFFFFF, 0xFFFFFFFF], dtype=np.uint32))
    assert (b.copyToHost(np.uint32) == 0xFFFFFFFF).all()
    b.fill(np.array([1, 2, 3, 4], dtype=np.uint32))
    assert np.array_equal(b.copyToHost(np.uint32, 8), [1, 2, 3, 4, 1, 2, 3, 4])
    b.resize(512)
    assert b.getSize() == 512 and b.getActualSize() >= 1024
    b.resize(4096)
    assert b.getSize() == 4096 and b.getActualSize() >= 4096
    b.upload(np.arange(16, dtype=np.uint32), offset=64)
    assert np.array_equal(b.copyToHost(np.uint32, 16, offset=64), np.arange(16))
    with pytest.raises(api.ClrtError):
        b.upload(np.zeros(8192, dtype=np.uint8))
    assert ctx.getMaxComputeUnits() == 148
    b.free()
