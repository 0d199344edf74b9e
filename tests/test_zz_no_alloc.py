"""kh_partition_fast_noalloc_dev: the hash partition with its scratch in a static device array (no allocation inside the call).
Collected last on purpose: a failure here hides no other test.

The entry point is called through the C ABI (ctypes) and checked against kh_partition_fast_dev, the existing entry point with the
same kernels: identical per-destination counts, and -- the order inside a destination being unspecified for both -- identical
sorted keys per destination (bit-exact, as every integer test here), plus a valid origin permutation.  Two calls in a row share
the static array (it keeps the previous call's contents): the second call uses other keys and FEWER destinations."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _partition(fn, dkeys, W, n, p, hash_kind, seed=9876543, want_perm=True):
    import torch
    out = torch.empty_like(dkeys)
    perm = torch.empty(n, dtype=torch.int32, device=dkeys.device) if want_perm else None
    counts = np.zeros(p, dtype=np.uint64)
    torch.cuda.synchronize()
    rc = fn(0, hash_kind, seed, W, p, dkeys.data_ptr(), n, out.data_ptr(), perm.data_ptr() if want_perm else None, counts.ctypes.data)
    return rc, out, perm, counts


def _segments(out, counts, W):
    a = out.cpu().numpy().view(np.uint64).reshape(-1, W)
    segs, off = [], 0
    for c in counts:
        s = a[off:off + int(c)]
        segs.append(s[np.lexsort(s.T[::-1])])
        off += int(c)
    return segs


@pytest.mark.parametrize("W,hash_name", [(1, "MURMUR64avx"), (2, "MURMUR64"), (1, "FARM")])
def test_partition_noalloc_matches_existing_entry_point_twice(kh, W, hash_name):
    import torch
    from kmerhash_b200._lib import lib
    from kmerhash_b200.hashmap import HASH_KINDS
    hk = HASH_KINDS[hash_name]
    rng = np.random.default_rng(4242 + W)
    for call, (n, p) in enumerate([(1_500_000, 37), (700_001, 8)]):           # second call: other keys, fewer destinations
        keys = rng.integers(0, 2**62, size=(n, W), dtype=np.uint64)
        keys[::5] = keys[0]                                                   # a heavy key
        flat = np.ascontiguousarray(keys if W == 2 else keys[:, 0])
        dk = torch.from_numpy(flat.view(np.int64)).cuda()
        rc0, out0, perm0, cnt0 = _partition(lib.kh_partition_fast_dev, dk, W, n, p, hk)
        rc1, out1, perm1, cnt1 = _partition(lib.kh_partition_fast_noalloc_dev, dk, W, n, p, hk)
        assert rc0 == 0 and rc1 == 0, lib.kh_last_error()
        assert int(cnt1.sum()) == n and np.array_equal(cnt0, cnt1)
        for a, b in zip(_segments(out0, cnt0, W), _segments(out1, cnt1, W)):
            assert np.array_equal(a, b)
        # the permutation names the origin of every permuted key and is a bijection
        pm = perm1.cpu().numpy().view(np.uint32).astype(np.int64)
        assert np.array_equal(np.sort(pm), np.arange(n))
        assert np.array_equal(keys[pm], out1.cpu().numpy().view(np.uint64).reshape(-1, W))
        # and without the permutation output
        rc2, out2, _, cnt2 = _partition(lib.kh_partition_fast_noalloc_dev, dk, W, n, p, hk, want_perm=False)
        assert rc2 == 0 and np.array_equal(cnt2, cnt0)
        for a, b in zip(_segments(out0, cnt0, W), _segments(out2, cnt2, W)):
            assert np.array_equal(a, b)


def test_partition_noalloc_limits(kh):
    import torch
    from kmerhash_b200._lib import lib
    from kmerhash_b200.hashmap import HASH_KINDS
    hk = HASH_KINDS["MURMUR64avx"]
    dk = torch.arange(1, 5001, dtype=torch.int64, device="cuda")
    rc, _, _, _ = _partition(lib.kh_partition_fast_noalloc_dev, dk, 1, 5000, 1025, hk)     # beyond the static scratch
    assert rc == -1 and b"static scratch" in lib.kh_last_error()
    rc, out, _, cnt = _partition(lib.kh_partition_fast_noalloc_dev, dk, 1, 5000, 1024, hk)   # the maximum itself
    assert rc == 0 and int(cnt.sum()) == 5000
    rc0, out0, _, cnt0 = _partition(lib.kh_partition_fast_dev, dk, 1, 5000, 1024, hk)
    assert rc0 == 0 and np.array_equal(cnt, cnt0)
    rc, _, _, cnt = _partition(lib.kh_partition_fast_noalloc_dev, dk, 1, 0, 4, hk)            # empty batch
    assert rc == 0 and int(cnt.sum()) == 0
