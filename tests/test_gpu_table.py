"""GPU tier: the table through the C ABI -- differential tests in the style of the reference's own
(test/unit/test_hashmap_robinhood_offsets.cpp:96-171,261-336,401-451: compare sorted contents and per-key
answers with a gold map), against the golden vectors of the compiled reference and against the oracle."""
import numpy as np
import pytest

from conftest import sort_pairs

pytestmark = pytest.mark.gpu

GEOMS_W1 = [(64, 1), (32, 1), (32, 2), (32, 4), (16, 2), (16, 4), (16, 8)]
GEOMS_W2 = [(64, 1), (32, 1), (32, 2), (32, 4)]


def all_geoms():
    return [(1, lb, g) for lb, g in GEOMS_W1] + [(2, lb, g) for lb, g in GEOMS_W2]


def check_against(m, om):
    gk, gv = sort_pairs(*m.to_vector())
    ok, ov = sort_pairs(*om.to_vector())
    assert m.size() == om.size()
    assert np.array_equal(gk, ok)
    assert np.array_equal(gv, ov)


@pytest.mark.parametrize("W,lb,g", all_geoms())
@pytest.mark.parametrize("red", ["count", "first", "last"])
def test_golden_benchmark_sequence(kh, golden, W, lb, g, red):
    """insert -> find -> count -> erase -> count2 (benchmark/BenchmarkHashTables.cpp:1191-1364) on the
    reference's own generate_input data; every answer compared with what the compiled reference returned."""
    tag = "k31" if W == 1 else "k63"
    keys, vals = golden["gen_keys_" + tag], golden["gen_vals_" + tag]
    m = kh.HashMap(W, "MURMUR64avx", red.upper())
    m.set_geometry(lb, g)
    if red == "count":
        new = m.insert(keys, default=1)
    else:
        new = m.insert(keys, vals)
    tk, tv = sort_pairs(*m.to_vector())
    assert new == len(tk) == m.size()
    assert np.array_equal(tk, golden["table_%s_keys_%s" % (red, tag)])
    assert np.array_equal(tv, golden["table_%s_vals_%s" % (red, tag)])
    q = np.ascontiguousarray(keys[:600])
    miss = np.ascontiguousarray(keys[:50] ^ np.uint64(5))
    assert np.array_equal(m.find(q, 0xFFFFFFFF), golden["find_%s_%s" % (red, tag)])
    assert np.array_equal(m.count(q), golden["count_%s_%s" % (red, tag)])
    assert np.array_equal(m.count(miss), golden["count_miss_%s_%s" % (red, tag)])
    assert m.erase(q) == int(golden["erased_%s_%s" % (red, tag)][0])
    assert np.array_equal(m.count(q), golden["count2_%s_%s" % (red, tag)])
    assert m.size() == int(golden["size_after_%s_%s" % (red, tag)][0])


@pytest.mark.parametrize("hash_kind", ["IDEN", "MURMUR64", "MURMUR64avx", "MURMUR32", "FARM", "FARM32", "CLHASH", "CRC32C"])
@pytest.mark.parametrize("W", [1, 2])
def test_every_store_hash(kh, oracle_mod, hash_kind, W):
    rng = np.random.default_rng(11)
    base = rng.integers(0, 2**62, size=(4000, W), dtype=np.uint64)
    keys = base[rng.integers(0, 4000, size=30000)]
    kk = keys if W == 2 else keys[:, 0].copy()
    m = kh.HashMap(W, hash_kind, "COUNT")
    om = oracle_mod.port.Map(W, "COUNT")
    assert m.insert(kk, default=1) == om.insert(kk, default=1)
    check_against(m, om)


@pytest.mark.parametrize("W,lb,g", all_geoms())
def test_growth_from_128_and_load_factors(kh, oracle_mod, W, lb, g):
    """insert_no_estimate: starts at capacity 128 and doubles (config 2: load-factor sweep incl. rehash)."""
    rng = np.random.default_rng(5)
    keys = rng.integers(0, 2**62, size=(60000, W), dtype=np.uint64)
    keys = np.concatenate([keys, keys[:20000]])
    kk = keys if W == 2 else keys[:, 0].copy()
    for max_load in (0.5, 0.6, 0.7, 0.8, 0.9):
        m = kh.HashMap(W, "MURMUR64", "COUNT", "LINEARPROBE")
        m.set_geometry(lb, g)
        m.set_load_factors(0.35 if max_load > 0.35 else 0.1, max_load)
        assert m.capacity() == 128
        om = oracle_mod.port.Map(W, "COUNT")
        assert m.insert_no_estimate(kk) == om.insert(kk)
        assert m.size() <= max_load * m.capacity() + 1
        assert m.capacity() >= 128 and (m.capacity() & (m.capacity() - 1)) == 0
        check_against(m, om)
        q = kk[::7]
        assert np.array_equal(m.find(q, 0), om.find(q, 0))


@pytest.mark.parametrize("W", [1, 2])
@pytest.mark.parametrize("red", ["COUNT", "FIRST", "LAST"])
def test_multi_batch_and_values(kh, oracle_mod, W, red):
    """several batches with explicit values; FIRST keeps the first value ever seen, LAST the last, COUNT adds"""
    rng = np.random.default_rng(17)
    pool = rng.integers(0, 2**62, size=(3000, W), dtype=np.uint64)
    m = kh.HashMap(W, "MURMUR64avx", red)
    om = oracle_mod.port.Map(W, red)
    for b in range(4):
        keys = pool[rng.integers(0, 3000, size=20000)]
        kk = keys if W == 2 else keys[:, 0].copy()
        vals = rng.integers(1, 1000, size=20000).astype(np.uint32)
        assert m.insert(kk, vals, estimate=bool(b & 1)) == om.insert(kk, vals)
        check_against(m, om)


@pytest.mark.parametrize("W,lb,g", all_geoms())
def test_erase_is_tombstone_free_and_chains_survive(kh, oracle_mod, W, lb, g):
    """fill close to max load so probe chains cross buckets, erase half, then every survivor must still be
    found, erased keys must be absent, and re-inserting reuses the freed slots without duplicates."""
    rng = np.random.default_rng(23)
    n = 50000
    keys = rng.integers(0, 2**62, size=(n, W), dtype=np.uint64)
    kk = keys if W == 2 else keys[:, 0].copy()
    m = kh.HashMap(W, "MURMUR64avx", "COUNT", max_load=0.9, min_load=0.01)
    m.set_geometry(lb, g)
    m.reserve(int(n * 0.99))                   # 2^16 slots -> load ~0.76
    om = oracle_mod.port.Map(W, "COUNT")
    m.insert_no_estimate(kk); om.insert(kk)
    cap = m.capacity()
    for rnd in range(3):
        victims = kk[rng.permutation(n)[: n // 2]]
        victims = np.concatenate([victims, victims[:1000]])       # duplicates inside the erase batch
        assert m.erase(victims) == om.erase(victims)
        assert m.capacity() == cap
        assert np.array_equal(m.count(kk), om.count(kk))
        assert np.array_equal(m.find(kk, 77), om.find(kk, 77))
        check_against(m, om)
        assert m.insert_no_estimate(kk) == om.insert(kk)
        check_against(m, om)
    # a rebuild (rehash) clears the sticky bits and keeps the contents
    m.rehash(cap * 2)
    check_against(m, om)


def test_shrink_after_erase(kh, oracle_mod):
    rng = np.random.default_rng(29)
    keys = rng.integers(0, 2**62, size=100000, dtype=np.uint64)
    m = kh.HashMap(1, "MURMUR64avx", "COUNT", min_load=0.35, max_load=0.8)
    om = oracle_mod.port.Map(1, "COUNT")
    m.insert(keys); om.insert(keys)
    cap = m.capacity()
    assert m.erase(keys[:90000]) == om.erase(keys[:90000])
    assert m.capacity() < cap                   # reserve(lsize) after the load fell below min_load
    check_against(m, om)


@pytest.mark.parametrize("W", [1, 2])
def test_find_existing_update_clear(kh, oracle_mod, W):
    rng = np.random.default_rng(31)
    keys = rng.integers(0, 2**62, size=(20000, W), dtype=np.uint64)
    kk = keys if W == 2 else keys[:, 0].copy()
    m = kh.HashMap(W, "FARM", "COUNT")
    om = oracle_mod.port.Map(W, "COUNT")
    m.insert(kk[:10000]); om.insert(kk[:10000])
    q = kk[rng.permutation(20000)[:8000]]
    gk, gv = m.find_existing(q)
    ek, ev = om.find_existing(q)
    assert np.array_equal(gk, ek) and np.array_equal(gv, ev)          # hits only, in query order
    upd = np.unique(q, axis=0)
    newv = rng.integers(5, 500, size=len(upd)).astype(np.uint32)
    n_upd = m.update(upd, newv)
    assert n_upd == int(om.count(upd).sum())
    got = m.find(upd, 0)
    present = om.count(upd).astype(bool)
    assert np.array_equal(got[present], newv[present]) and not got[~present].any()
    m.clear()
    assert m.size() == 0 and not m.count(kk).any()
    assert m.insert(kk[:10]) == 10


def test_empty_batches_and_edge_keys(kh, oracle_mod):
    m = kh.HashMap(1, "MURMUR64avx", "COUNT")
    e = np.zeros(0, dtype=np.uint64)
    assert m.insert(e) == 0 and m.erase(e) == 0 and m.size() == 0
    assert m.count(e).shape == (0,) and m.find(e).shape == (0,)
    assert m.to_vector()[0].shape == (0,)
    # key 0 (poly-A), the largest 31-mer, and the all-ones word (not a 31-mer but a legal uint64 key,
    # BenchmarkDistHashTables uses raw uint64 keys) all behave like any other key
    edge = np.array([0, (1 << 62) - 1, 0xFFFFFFFFFFFFFFFF, 0xFFFFFFFFFFFFFFFE, 0, 0xFFFFFFFFFFFFFFFF], dtype=np.uint64)
    assert m.insert(edge) == 4
    assert m.find(edge, 9).tolist() == [2, 1, 2, 1, 2, 2]
    k, v = sort_pairs(*m.to_vector())
    assert k.tolist() == [0, (1 << 62) - 1, 0xFFFFFFFFFFFFFFFE, 0xFFFFFFFFFFFFFFFF] and v.tolist() == [2, 1, 1, 2]
    assert m.erase(edge[2:3]) == 1 and m.size() == 3
    assert m.count(edge).tolist() == [1, 1, 0, 1, 1, 0]
    m2 = kh.HashMap(2, "MURMUR64avx", "COUNT")
    edge2 = np.array([[0, 0], [0xFFFFFFFFFFFFFFFF, 0xFFFFFFFFFFFFFFFF], [0xFFFFFFFFFFFFFFFF, 0], [0, 0xFFFFFFFFFFFFFFFF],
                      [0xFFFFFFFFFFFFFFFF, 0xFFFFFFFFFFFFFFFF]], dtype=np.uint64)
    assert m2.insert(edge2) == 4
    assert m2.find(edge2, 0).tolist() == [1, 2, 1, 1, 2]


def test_bad_arguments(kh):
    with pytest.raises(ValueError):
        kh.HashMap(3)
    m = kh.HashMap(1)
    with pytest.raises(ValueError):
        m.set_load_factors(0.9, 0.5)
    with pytest.raises(ValueError):
        m.insert(np.zeros((4, 2), dtype=np.uint64))


@pytest.mark.parametrize("W", [1, 2])
def test_hll_estimate(kh, oracle_mod, golden, W):
    tag = "k31" if W == 1 else "k63"
    keys = golden["gen_keys_" + tag]
    m = kh.HashMap(W, "MURMUR64avx", "COUNT")
    est, regs = m.estimate_distinct(keys, return_registers=True)
    pest, pregs = oracle_mod.port.hll_estimate(oracle_mod.port.hash("MURMUR64avx", keys, 43), 64)
    assert np.array_equal(regs, pregs)                     # registers bit-identical
    assert est == pytest.approx(float(golden["hll_est_" + tag][0]), rel=1e-12)
    m32 = kh.HashMap(W, "MURMUR32", "COUNT")
    est32, regs32 = m32.estimate_distinct(keys, return_registers=True)
    _, pregs32 = oracle_mod.port.hll_estimate(oracle_mod.port.hash("MURMUR32", keys, 43), 32)
    assert np.array_equal(regs32, pregs32)


def test_device_pointer_path(kh, oracle_mod):
    import torch
    rng = np.random.default_rng(41)
    keys = rng.integers(0, 2**62, size=300000, dtype=np.uint64)
    keys = np.concatenate([keys, keys[:100000]])
    d = torch.from_numpy(keys.view(np.int64)).cuda()
    m = kh.HashMap(1, "MURMUR64avx", "COUNT")
    om = oracle_mod.port.Map(1, "COUNT")
    assert m.insert(d) == om.insert(keys)
    assert np.array_equal(m.find(d, 0).cpu().numpy().view(np.uint32), om.find(keys, 0))
    assert np.array_equal(m.count(d).cpu().numpy(), om.count(keys))
    assert m.erase(d[:1000]) == om.erase(keys[:1000])
    check_against(m, om)


def test_large_batch_properties(kh):
    """size-independent properties at a scale the oracle is too slow for: 20M keys with 5x repeats"""
    rng = np.random.default_rng(43)
    uniq = rng.integers(0, 2**62, size=4_000_000, dtype=np.uint64)
    uniq = np.unique(uniq)
    reps = rng.integers(1, 10, size=len(uniq))
    keys = np.repeat(uniq, reps)
    rng.shuffle(keys)
    m = kh.HashMap(1, "MURMUR64avx", "COUNT")
    assert m.insert(keys) == len(uniq) == m.size()
    tk, tv = sort_pairs(*m.to_vector())
    assert np.array_equal(tk, uniq)                         # identical key set
    assert int(tv.sum()) == len(keys)                       # counts add up to the input size
    assert np.array_equal(tv, reps)                         # and equal the true multiplicities
    assert m.count(keys[:1_000_000]).all()
    assert m.erase(keys[:1_000_000]) == len(np.unique(keys[:1_000_000]))
    assert not m.count(keys[:1_000_000]).any()
    assert m.erase(keys[:1_000_000]) == 0                   # idempotent


@pytest.mark.parametrize("n_gpus", [1, 2])
def test_multi_gpu_map(kh, oracle_mod, n_gpus):
    if kh.device_count() < n_gpus:
        pytest.skip("needs %d GPUs" % n_gpus)
    rng = np.random.default_rng(47)
    raw = rng.integers(0, 2**62, size=60000, dtype=np.uint64)
    raw = np.concatenate([raw, raw[:30000]])
    canon = oracle_mod.port.canonical(31, raw)
    m = kh.MultiGpuMap(n_gpus, 1, "MURMUR64avx", "MURMUR64avx", "COUNT", canonical_k=31)
    om = oracle_mod.port.Map(1, "COUNT")
    assert m.insert(raw) == om.insert(canon) == m.size()
    ranks = oracle_mod.port.rank("MURMUR64avx", canon, n_gpus)
    for r in range(n_gpus):      # per-rank partition contents identical to the reference's rule
        gk, gv = sort_pairs(*m.to_vector_rank(r))
        mine = np.unique(canon[ranks == r])
        assert np.array_equal(gk, mine)
        assert np.array_equal(gv, om.find(mine, 0))
    q = raw[rng.permutation(len(raw))[:20000]]
    q = np.concatenate([q, q[:100] ^ np.uint64(3)])
    cq = oracle_mod.port.canonical(31, q)
    assert np.array_equal(m.count(q), om.count(cq))         # answers in INPUT order
    assert np.array_equal(m.find(q, 0), om.find(cq, 0))
    assert m.erase(q) == om.erase(cq)
    assert m.size() == om.size()


@pytest.mark.parametrize("W", [1, 2])
@pytest.mark.parametrize("hash_kind", ["MURMUR64avx", "MURMUR32", "IDEN"])
@pytest.mark.parametrize("window_bytes", [1 << 12, 1 << 16])
def test_windowed_bulk_insert_matches_oracle(kh, oracle_mod, W, hash_kind, window_bytes):
    """L2-windowed insert (batch sorted by table region first) forced on at test sizes: same contents, same answers,
    also for 32-bit and identity store hashes (top bits of the spread hash pick the region)."""
    rng = np.random.default_rng(17)
    base = rng.integers(0, 2**62, size=(40000, W), dtype=np.uint64)
    keys = base[rng.integers(0, 40000, size=300000)]
    kk = np.ascontiguousarray(keys if W == 2 else keys[:, 0])
    m = kh.HashMap(W, hash_kind, "COUNT")
    m.set_insert_window(True, window_bytes, 1)
    om = oracle_mod.port.Map(W, "COUNT")
    assert m.insert(kk, default=1) == om.insert(kk, default=1)          # HLL presize + windowed single piece
    check_against(m, om)
    more = np.ascontiguousarray(np.concatenate([kk[:50000], kk[:1000] ^ np.uint64(3)]))
    assert m.insert_no_estimate(more, default=2) == om.insert(more, default=2)   # second batch into a live table
    check_against(m, om)
    q = np.ascontiguousarray(kk[:5000])
    assert np.array_equal(m.find(q, 0), om.find(q, 0))
    assert m.erase(q) == om.erase(q)
    check_against(m, om)


def test_windowed_and_plain_insert_agree_at_scale(kh):
    """20 M keys (large enough for the default window thresholds): identical table either way, and the
    size-independent invariant sum(counts) == n."""
    rng = np.random.default_rng(3)
    uniq = rng.integers(0, 2**62, size=4_000_000, dtype=np.uint64)
    keys = np.ascontiguousarray(uniq[rng.integers(0, len(uniq), size=20_000_000)])
    out = []
    for on in (True, False):
        m = kh.HashMap(1, "MURMUR64avx", "COUNT", min_load=0.35, max_load=0.8)
        m.set_insert_window(on, 4 << 20, 1 << 20)
        new = m.insert(keys, default=1)
        k, v = sort_pairs(*m.to_vector())
        assert new == len(k) and int(v.astype(np.uint64).sum()) == len(keys)
        out.append((k, v))
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])
    assert np.array_equal(out[0][0], np.unique(keys))


@pytest.mark.parametrize("p", [1, 2, 3, 8])
@pytest.mark.parametrize("W", [1, 2])
def test_planned_exchange_halves(kh, oracle_mod, p, W):
    """khx_plan_scatter_dev / khx_insert_recv_dev with the all-to-all emulated on one GPU: every rank's table must
    hold exactly the reference's rank contents { (k, c) in the global count map : ref_rank(k, p) == r } and the
    metadata must agree with the oracle (counts per destination; HLL registers of the received keys)."""
    import torch
    o = oracle_mod.port
    rng = np.random.default_rng(40 + p + W)
    base = rng.integers(0, 2**62, size=(30000, W), dtype=np.uint64)
    batches = [np.ascontiguousarray(base[rng.integers(0, 30000, size=120000 + 1000 * g)]) for g in range(p)]
    batches = [b if W == 2 else np.ascontiguousarray(b[:, 0]) for b in batches]
    tables = [kh.HashMap(W, "MURMUR64avx", "COUNT") for _ in range(p)]
    for t in tables:
        t.set_insert_window(True, 1 << 14, 1)          # exercise the windowed path with the shipped histogram
    sent = []
    for g in range(p):
        d = torch.from_numpy(batches[g].view(np.int64)).cuda()
        out, meta, counts = tables[g].plan_scatter(d, p)
        ranks = o.rank("MURMUR64avx", batches[g], p)
        assert counts == np.bincount(ranks, minlength=p).tolist()
        assert meta[:, 0].tolist() == counts
        assert meta[:, 1:1025].sum(dim=1).tolist() == counts          # region histogram rows sum to the counts
        sent.append((out, meta, counts))
    allk = np.concatenate(batches)
    g_map = o.Map(W, "COUNT"); g_map.insert(allk, default=1)
    gk, gv = g_map.to_vector()
    owner = o.rank("MURMUR64avx", gk, p)
    for r in range(p):
        segs, metas = [], []
        for g in range(p):
            out, meta, counts = sent[g]
            start = sum(counts[:r])
            segs.append(out[start:start + counts[r]])
            metas.append(meta[r])
        recv = torch.cat(segs).contiguous()
        recv_meta = torch.stack(metas).contiguous()
        new = tables[r].insert_recv(recv, recv_meta, 1)
        ek, ev = sort_pairs(gk[owner == r], gv[owner == r])
        tk, tv = sort_pairs(*tables[r].to_vector())
        assert new == len(ek)
        assert np.array_equal(tk, ek) and np.array_equal(tv, ev)
        # shipped registers == registers of an HLL pass over the received keys
        hk = recv.cpu().numpy().view(np.uint64)
        _, regs = o.hll_estimate(o.hash("MURMUR64avx", hk if W == 2 else hk.reshape(-1), 43))
        merged = recv_meta[:, 1025:].max(dim=0).values.cpu().numpy().astype(np.uint8)
        assert np.array_equal(merged, np.asarray(regs, dtype=np.uint8))


@pytest.mark.parametrize("p", [2, 3, 8])
@pytest.mark.parametrize("W", [1, 2])
def test_fused_scatter_into_peer_buffers(kh, oracle_mod, p, W):
    """khx_plan_dev + khx_scatter_peers_dev: the rank scatter stores straight into the owners' receive buffers.
    p ranks emulated on one GPU (khx_peer_set maps plain device pointers); afterwards every receive buffer must
    hold exactly the keys the reference's rule assigns to that rank, segment by segment, and the receiver half
    must build the reference's per-rank table from it."""
    import torch
    from kmerhash_b200._lib import lib
    o = oracle_mod.port
    rng = np.random.default_rng(70 + p + W)
    base = rng.integers(0, 2**62, size=(20000, W), dtype=np.uint64)
    batches = [np.ascontiguousarray(base[rng.integers(0, 20000, size=90000 + 777 * g)]) for g in range(p)]
    batches = [b if W == 2 else np.ascontiguousarray(b[:, 0]) for b in batches]
    tables = [kh.HashMap(W, "MURMUR64avx", "COUNT") for _ in range(p)]
    dkeys = [torch.from_numpy(b.view(np.int64)).cuda() for b in batches]
    metas = [tables[g].plan(dkeys[g], p) for g in range(p)]
    allmeta = torch.stack(metas)                                   # [g][r][:]
    Cm = allmeta[:, :, 0].cpu().numpy().astype(np.uint64)
    for g in range(p):
        assert Cm[g].tolist() == np.bincount(o.rank("MURMUR64avx", batches[g], p), minlength=p).tolist()
    cap = int(Cm.sum(axis=0).max()) + 16
    peers = [lib.khx_peer_create(0, r, p, cap, W) for r in range(p)]
    try:
        assert all(peers)
        for r in range(p):
            for q in range(p):
                if q != r:
                    assert lib.khx_peer_set(peers[r], q, lib.khx_peer_recv_ptr(peers[q])) == 0
        for g in range(p):
            tables[g].scatter_peers(dkeys[g], peers[g], Cm[:g, :].sum(axis=0), Cm[g, :])
        allk = np.concatenate(batches)
        gm = o.Map(W, "COUNT"); gm.insert(allk, default=1)
        gk, gv = gm.to_vector()
        owner = o.rank("MURMUR64avx", gk, p)
        for r in range(p):
            n_r = int(Cm[:, r].sum())
            buf = torch.empty((n_r,) if W == 1 else (n_r, W), dtype=torch.int64, device="cuda")
            import ctypes
            torch.cuda.synchronize()
            # copy the receive buffer out for inspection (cudaMemcpy through torch: wrap the raw pointer)
            lib_cudart = ctypes.CDLL("libcudart.so")
            assert lib_cudart.cudaMemcpy(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(lib.khx_peer_recv_ptr(peers[r])),
                                         ctypes.c_size_t(n_r * 8 * W), 3) == 0
            got = buf.cpu().numpy().view(np.uint64)
            off = 0
            for g in range(p):                                     # segment of source g: same multiset as the rule gives
                c = int(Cm[g, r])
                seg = got[off:off + c]
                src = batches[g][o.rank("MURMUR64avx", batches[g], p) == r]
                a = np.sort(seg, axis=0) if W == 1 else seg[np.lexsort(seg.T[::-1])]
                b = np.sort(src, axis=0) if W == 1 else src[np.lexsort(src.T[::-1])]
                assert np.array_equal(a, b)
                off += c
            new = tables[r].insert_recv_ptr(lib.khx_peer_recv_ptr(peers[r]), n_r, allmeta[:, r, :].contiguous(), 1)
            ek, ev = sort_pairs(gk[owner == r], gv[owner == r])
            tk, tv = sort_pairs(*tables[r].to_vector())
            assert new == len(ek) and np.array_equal(tk, ek) and np.array_equal(tv, ev)
    finally:
        for h in peers:
            if h:
                lib.khx_peer_destroy(h)


@pytest.mark.parametrize("p,W,n_per,tiled", [(2, 1, 90000, False), (3, 2, 60000, False), (8, 1, 70000, False), (8, 2, 30000, False),
                                             (2, 1, 3000000, True), (8, 1, 1500000, True)])
def test_single_pass_slab_exchange(kh, oracle_mod, p, W, n_per, tiled):
    """khx_plan_scatter_slabs_dev + khx_insert_recv_slabs_dev (p ranks emulated on one GPU): the one-pass sender must
    produce the metadata of khx_plan_dev bit for bit and fill slab (owner r, sender g) with exactly the keys the
    reference's rule sends from g to r; the receiver, reading the slabs in place, must build the reference's per-rank
    table.  Two batches on alternating halves; then an overflowing slab: nothing may be written past it and the
    metadata must still carry the true counts."""
    import ctypes
    import torch
    from kmerhash_b200._lib import lib
    o = oracle_mod.port
    cudart = ctypes.CDLL("libcudart.so")
    rng = np.random.default_rng(170 + p + W)
    n_base = max(20000, n_per // 4)
    base = rng.integers(0, 2**62, size=(n_base, W), dtype=np.uint64)
    tables = [kh.HashMap(W, "MURMUR64avx", "COUNT") for _ in range(p)]
    if tiled:
        for t in tables:
            t.set_insert_window(True, 1 << 14, 1)
    slab_cap = int(n_per / p * 1.5) + 1024
    cap = 2 * p * slab_cap + 64                                    # 64 guard keys behind the last slab
    peers = [lib.khx_peer_create(0, r, p, cap, W) for r in range(p)]

    def read_buf(r):
        buf = torch.empty((cap, W), dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        assert cudart.cudaMemcpy(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(lib.khx_peer_recv_ptr(peers[r])),
                                 ctypes.c_size_t(cap * 8 * W), 3) == 0
        return buf.cpu().numpy().view(np.uint64)

    def srt(a):
        return np.sort(a[:, 0]) if W == 1 else a[np.lexsort(a.T[::-1])]

    try:
        assert all(peers)
        for r in range(p):
            for q in range(p):
                if q != r:
                    assert lib.khx_peer_set(peers[r], q, lib.khx_peer_recv_ptr(peers[q])) == 0
        gm = o.Map(W, "COUNT")
        for parity in (0, 1):
            batches = [np.ascontiguousarray(base[rng.integers(0, n_base, size=n_per + 777 * g)]) for g in range(p)]
            flat = [b if W == 2 else np.ascontiguousarray(b[:, 0]) for b in batches]
            dkeys = [torch.from_numpy(b.view(np.int64)).cuda() for b in flat]
            metas = [tables[g].plan_scatter_slabs(dkeys[g], peers[g], p, slab_cap, parity) for g in range(p)]
            for g in range(p):
                assert torch.equal(metas[g], tables[g].plan(dkeys[g], p))
            allmeta = torch.stack(metas)
            Cm = allmeta[:, :, 0].cpu().numpy().astype(np.uint64)
            assert int(Cm.max()) <= slab_cap
            for r in range(p):
                got = read_buf(r)
                for g in range(p):
                    lo = (parity * p + g) * slab_cap
                    src = batches[g][o.rank("MURMUR64avx", flat[g], p) == r]
                    assert int(Cm[g, r]) == len(src)
                    assert np.array_equal(srt(got[lo:lo + len(src)]), srt(src))
            gm.insert(np.concatenate(flat), default=1)
            gk, gv = gm.to_vector()
            owner = o.rank("MURMUR64avx", gk, p)
            for r in range(p):
                tables[r].insert_recv_slabs(peers[r], slab_cap, parity, Cm[:, r], allmeta[:, r, :].contiguous(), 1)
                ek, ev = sort_pairs(gk[owner == r], gv[owner == r])
                tk, tv = sort_pairs(*tables[r].to_vector())
                assert np.array_equal(tk, ek) and np.array_equal(tv, ev)
        # overflow: slabs far too small.  Guard keys behind each slab region must survive; counts stay true.
        small = 1000
        before = [read_buf(r) for r in range(p)]
        meta = tables[0].plan_scatter_slabs(dkeys[0], peers[0], p, small, 0)
        assert torch.equal(meta[:, 0], tables[0].plan(dkeys[0], p)[:, 0]) and int(meta[:, 0].max()) > small
        for r in range(p):
            after = read_buf(r)
            changed = np.nonzero((after != before[r]).any(axis=1))[0]
            assert changed.size == 0 or (changed.min() >= 0 and changed.max() < small)   # sender 0, half 0: slab [0, small)
    finally:
        for h in peers:
            if h:
                lib.khx_peer_destroy(h)



@pytest.mark.parametrize("p,W,n_per,pieces", [(2, 1, 90000, 1), (8, 1, 70000, 3), (3, 2, 60000, 2), (8, 1, 1500000, 4)])
def test_slab_exchange_without_metadata(kh, oracle_mod, p, W, n_per, pieces):
    """khx_scatter_slabs_dev (sender without metadata; the batch handed over in `pieces` chunks with accumulating counts, as
    behind a PCIe copy) + khx_insert_recv_slabs_dev(d_recv_meta = NULL): slab (owner r, sender g) must hold exactly the keys
    the reference's rule sends from g to r, the counts must be the true ones also when a slab overflows, and every
    receiver -- whose level-1 pass estimates from the received keys themselves -- must build the reference's per-rank table
    with the reference's capacity."""
    import ctypes
    import torch
    from kmerhash_b200._lib import lib
    o = oracle_mod.port
    cudart = ctypes.CDLL("libcudart.so")
    rng = np.random.default_rng(910 + p + W)
    n_base = max(20000, n_per // 4)
    base = rng.integers(0, 2**62, size=(n_base, W), dtype=np.uint64)
    tables = [kh.HashMap(W, "MURMUR64avx", "COUNT") for _ in range(p)]
    slab_cap = int(n_per / p * 1.5) + 1024
    cap = 2 * p * slab_cap + 64
    peers = [lib.khx_peer_create(0, r, p, cap, W) for r in range(p)]

    def read_buf(r):
        buf = torch.empty((cap, W), dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        assert cudart.cudaMemcpy(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(lib.khx_peer_recv_ptr(peers[r])),
                                 ctypes.c_size_t(cap * 8 * W), 3) == 0
        return buf.cpu().numpy().view(np.uint64)

    def srt(a):
        return np.sort(a[:, 0]) if W == 1 else a[np.lexsort(a.T[::-1])]

    try:
        assert all(peers)
        for r in range(p):
            for q in range(p):
                if q != r:
                    assert lib.khx_peer_set(peers[r], q, lib.khx_peer_recv_ptr(peers[q])) == 0
        gm = o.Map(W, "COUNT")
        ref = getattr(oracle_mod, "ref", None)              # the compiled reference (capacity rule incl. its cumulative HLL)
        refs = [ref.Map("OFFSETS_COUNT", W, "MURMUR64avx", 0.2, 0.8) for _ in range(p)] if ref is not None else None
        for parity in (0, 1):
            batches = [np.ascontiguousarray(base[rng.integers(0, n_base, size=n_per + 777 * g)]) for g in range(p)]
            flat = [b if W == 2 else np.ascontiguousarray(b[:, 0]) for b in batches]
            dkeys = [torch.from_numpy(b.view(np.int64)).cuda() for b in flat]
            counts = [torch.zeros(p, dtype=torch.int64, device="cuda") for _ in range(p)]
            for g in range(p):
                n = dkeys[g].shape[0]
                for i in range(pieces):
                    tables[g].scatter_slabs(dkeys[g][n * i // pieces:n * (i + 1) // pieces].contiguous(), peers[g], slab_cap, parity,
                                            counts[g], reset=(i == 0))
            for g in range(p):
                assert lib.kh_count_dev(tables[g]._h, None, 0, None) == 0          # synchronises table g's stream (the sender is asynchronous)
            Cm = torch.stack(counts).cpu().numpy().astype(np.uint64)
            assert int(Cm.max()) <= slab_cap
            for r in range(p):
                got = read_buf(r)
                for g in range(p):
                    lo = (parity * p + g) * slab_cap
                    src = batches[g][o.rank("MURMUR64avx", flat[g], p) == r]
                    assert int(Cm[g, r]) == len(src)
                    assert np.array_equal(srt(got[lo:lo + len(src)]), srt(src))
            gm.insert(np.concatenate(flat), default=1)
            gk, gv = gm.to_vector()
            owner = o.rank("MURMUR64avx", gk, p)
            allk = np.concatenate(flat)
            allowner = o.rank("MURMUR64avx", allk, p)
            for r in range(p):
                tables[r].insert_recv_slabs(peers[r], slab_cap, parity, Cm[:, r], None, 1, no_meta=True)
                ek, ev = sort_pairs(gk[owner == r], gv[owner == r])
                tk, tv = sort_pairs(*tables[r].to_vector())
                assert np.array_equal(tk, ek) and np.array_equal(tv, ev)
                if refs is not None:
                    refs[r].insert(np.ascontiguousarray(allk[allowner == r]), default=1)
                    assert tables[r].capacity() == refs[r].capacity()
        # overflow: the slabs far too small.  Nothing is written past a slab; the counts stay the true ones.
        small = 1000
        before = [read_buf(r) for r in range(p)]
        cnt = torch.zeros(p, dtype=torch.int64, device="cuda")
        tables[0].scatter_slabs(dkeys[0], peers[0], small, 0, cnt)
        assert lib.kh_count_dev(tables[0]._h, None, 0, None) == 0
        exp = np.bincount(o.rank("MURMUR64avx", flat[0], p), minlength=p)
        assert np.array_equal(cnt.cpu().numpy(), exp) and int(exp.max()) > small
        for r in range(p):
            after = read_buf(r)
            changed = np.nonzero((after != before[r]).any(axis=1))[0]
            assert changed.size == 0 or (changed.min() >= 0 and changed.max() < small)
    finally:
        for h in peers:
            if h:
                lib.khx_peer_destroy(h)


@pytest.mark.parametrize("n,distinct,skew", [(3_000_000, 40_000, 0.0), (3_000_000, 1_500_000, 0.0), (6_000_000, 5_000_000, 0.0),
                                             (3_000_000, 1_000_000, 0.5), (2_000_000, 1_900_000, 0.97)])
def test_fused_one_read_bulk_insert(kh, oracle_mod, n, distinct, skew):
    """insert() of a large COUNT batch takes the one-read route (region scatter in SLAB mode computing the HLL registers,
    then the tile insert on 32-bit hash tags): contents must equal the oracle's, the capacity must equal the one the
    classic estimate -> reserve -> insert route reaches (same HLL registers => same reserve), for tables with fewer and
    more tiles than regions, for a second batch into the filled table, and for skewed batches (region overflow list;
    at 97 % one key the whole batch falls back to the two-pass route)."""
    o = oracle_mod.port
    rng = np.random.default_rng(int(n + distinct + 100 * skew))
    base = rng.integers(0, 2**62, size=distinct, dtype=np.uint64)
    base[0] = np.uint64(0xFFFFFFFFFFFFFFFF)                              # the side-slot key rides along
    m = kh.HashMap(1, "MURMUR64avx", "COUNT")
    m.set_insert_window(True, 32 << 20, 1 << 20)
    c = kh.HashMap(1, "MURMUR64avx", "COUNT")
    c.set_insert_window(False, 32 << 20, 1 << 20)
    gm = o.Map(1, "COUNT")
    for batch in range(2):
        keys = base[rng.integers(0, distinct, size=n)]
        if skew:
            keys[rng.random(n) < skew] = base[1]
        keys = np.ascontiguousarray(keys)
        import torch
        d = torch.from_numpy(keys.view(np.int64)).cuda()
        new_m = m.insert(d, default=1)
        new_c = c.insert(d, default=1)
        before = gm.size()
        gm.insert(keys, default=1)
        assert new_m == new_c == gm.size() - before
        assert m.capacity() == c.capacity() and m.size() == gm.size()
        ek, ev = sort_pairs(*gm.to_vector())
        tk, tv = sort_pairs(*m.to_vector())
        assert np.array_equal(tk, ek) and np.array_equal(tv, ev)
    q = np.ascontiguousarray(np.concatenate([base[:1000], rng.integers(0, 2**62, size=1000, dtype=np.uint64)]))
    assert np.array_equal(m.find(q, 0), gm.find(q, 0))


@pytest.mark.parametrize("hash_kind", ["MURMUR32", "CRC32C", "FARM", "FARM32", "CLHASH", "MURMUR64", "IDEN"])
def test_fused_one_read_bulk_insert_hash_kinds(kh, oracle_mod, hash_kind):
    """the one-read route under every store hash (32-bit hashes take the 32-bit HLL branch): same contents as the
    oracle, same capacity as the classic estimate -> reserve -> insert route"""
    import torch
    o = oracle_mod.port
    rng = np.random.default_rng(len(hash_kind) + 5)
    base = rng.integers(0, 2**62, size=1_200_000, dtype=np.uint64)
    keys = np.ascontiguousarray(base[rng.integers(0, base.shape[0], size=2_500_000)])
    d = torch.from_numpy(keys.view(np.int64)).cuda()
    m = kh.HashMap(1, hash_kind, "COUNT")
    m.set_insert_window(True, 32 << 20, 1 << 20)
    c = kh.HashMap(1, hash_kind, "COUNT")
    c.set_insert_window(False, 32 << 20, 1 << 20)
    gm = o.Map(1, "COUNT")
    gm.insert(keys, default=1)
    assert m.insert(d, default=1) == c.insert(d, default=1) == gm.size()
    assert m.capacity() == c.capacity()
    ek, ev = sort_pairs(*gm.to_vector())
    tk, tv = sort_pairs(*m.to_vector())
    assert np.array_equal(tk, ek) and np.array_equal(tv, ev)


def test_peer_memory_barrier(kh):
    """khx_peer_sync_dev: p ranks emulated on one GPU, one host thread per rank (the flag kernels wait for each other, so
    they must be able to run concurrently).  Every rank must come back with the same p x p count matrix, row g = what
    rank g published; a second epoch without counts must pass as well."""
    import threading
    import torch
    from kmerhash_b200._lib import lib
    p = 4
    tables = [kh.HashMap(1, "MURMUR64avx", "COUNT") for _ in range(p)]
    peers = [lib.khx_peer_create(0, r, p, 1024, 1) for r in range(p)]
    try:
        assert all(peers)
        for r in range(p):
            for q in range(p):
                if q != r:
                    assert lib.khx_peer_set(peers[r], q, lib.khx_peer_recv_ptr(peers[q])) == 0
        rows = [torch.tensor([100 * g + c for c in range(p)], dtype=torch.int64, device="cuda") for g in range(p)]
        torch.cuda.synchronize()
        got, errs = [None] * p, []

        def run(r):
            try:
                got[r] = tables[r].peer_sync(peers[r], 1, rows[r], p)
                tables[r].peer_sync(peers[r], 2)
            except Exception as e:      # noqa: BLE001
                errs.append(e)

        th = [threading.Thread(target=run, args=(r,)) for r in range(p)]
        [t.start() for t in th]
        [t.join(60) for t in th]
        assert not errs, errs
        exp = np.array([[100 * g + c for c in range(p)] for g in range(p)], dtype=np.uint64)
        for r in range(p):
            assert np.array_equal(got[r], exp)
    finally:
        for h in peers:
            if h:
                lib.khx_peer_destroy(h)


def test_peer_barrier_timeout_poisons_every_rank(kh, monkeypatch):
    """a rank that gives up in the flag barrier must not leave its peers paired with the wrong epoch: it writes a poison
    value into every peer's copy of its flag, its own buffers become unusable, and a peer that reaches the barrier later
    (or any later barrier) fails too instead of passing with stale counts"""
    import torch
    from kmerhash_b200._lib import lib, KmerHashError
    monkeypatch.setenv("KH_PEER_TIMEOUT_MS", "150")
    p = 2
    tables = [kh.HashMap(1, "MURMUR64avx", "COUNT") for _ in range(p)]
    peers = [lib.khx_peer_create(0, r, p, 1024, 1) for r in range(p)]
    try:
        assert all(peers)
        for r in range(p):
            for q in range(p):
                if q != r:
                    assert lib.khx_peer_set(peers[r], q, lib.khx_peer_recv_ptr(peers[q])) == 0
        row = torch.tensor([5, 6], dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        with pytest.raises(KmerHashError):
            tables[0].peer_sync(peers[0], 1, row, p)          # rank 1 never shows up
        assert lib.khx_peer_poisoned(peers[0]) == 1
        with pytest.raises(KmerHashError):
            tables[0].peer_sync(peers[0], 2)                   # poisoned buffers refuse further barriers
        with pytest.raises(KmerHashError):
            tables[1].peer_sync(peers[1], 1, row, p)          # the late peer sees the poison, not a passed barrier
        assert lib.khx_peer_poisoned(peers[1]) == 1
    finally:
        for h in peers:
            if h:
                lib.khx_peer_destroy(h)


@pytest.mark.parametrize("p,W", [(2, 1), (3, 2), (8, 1), (8, 2)])
def test_query_slab_exchange(kh, oracle_mod, p, W):
    """khx_query_scatter_slabs_dev / khx_query_slabs_dev / khx_unpermute_slabs_dev with p ranks emulated on one GPU: every
    rank's table holds its partition of a count map; every rank asks a batch (present and absent keys); find / count
    answers must come back in INPUT order and equal the oracle's; erase must remove exactly the asked keys."""
    import torch
    from kmerhash_b200._lib import lib
    o = oracle_mod.port
    rng = np.random.default_rng(270 + p + W)
    base = rng.integers(0, 2**62, size=(40000, W), dtype=np.uint64)
    sq = (lambda a: a if W == 2 else np.ascontiguousarray(a[:, 0]))
    stored = sq(np.ascontiguousarray(base[rng.integers(0, 30000, size=200000)]))          # last 10000 base keys never stored
    gm = o.Map(W, "COUNT"); gm.insert(stored, default=1)
    owner = o.rank("MURMUR64avx", stored, p)
    tables = [kh.HashMap(W, "MURMUR64avx", "COUNT") for _ in range(p)]
    for r in range(p):
        tables[r].insert(stored[owner == r], default=1)
    n_q = 50000
    slab_cap = int(n_q / p * 1.5) + 512
    peers = [lib.khx_peer_create_q(0, r, p, p * slab_cap, W) for r in range(p)]
    try:
        assert all(peers)
        for r in range(p):
            for q in range(p):
                if q != r:
                    assert lib.khx_peer_set(peers[r], q, lib.khx_peer_recv_ptr(peers[q])) == 0
        for op in ("find", "count", "erase"):
            asks = [sq(np.ascontiguousarray(base[rng.integers(0, 40000, size=n_q + 111 * g)])) for g in range(p)]
            dq = [torch.from_numpy(a.view(np.int64)).cuda() for a in asks]
            perms = [torch.empty(p * slab_cap, dtype=torch.int32, device="cuda") for _ in range(p)]
            cnts = [torch.zeros(p, dtype=torch.int64, device="cuda") for _ in range(p)]
            for g in range(p):
                tables[g].query_scatter_slabs(dq[g], peers[g], slab_cap, perms[g], cnts[g])
            Cm = torch.stack(cnts).cpu().numpy().astype(np.uint64)                          # [g][r]
            for g in range(p):
                assert Cm[g].tolist() == np.bincount(o.rank("MURMUR64avx", asks[g], p), minlength=p).tolist()
            hits = [tables[r].query_slabs(peers[r], op, slab_cap, Cm[:, r], 77) for r in range(p)]
            if op == "erase":
                allq = np.concatenate(asks)
                assert sum(hits) == gm.erase(allq)
                gk, gv = gm.to_vector()
                own = o.rank("MURMUR64avx", gk, p)
                for r in range(p):
                    ek, ev = sort_pairs(gk[own == r], gv[own == r])
                    tk, tv = sort_pairs(*tables[r].to_vector())
                    assert np.array_equal(tk, ek) and np.array_equal(tv, ev)
                continue
            for g in range(p):
                out = torch.full((asks[g].shape[0],), 0x55, dtype=torch.int32 if op == "find" else torch.uint8, device="cuda")
                tables[g].unpermute_slabs(peers[g], slab_cap, perms[g], cnts[g], out)
                got = out.cpu().numpy()
                exp = gm.find(asks[g], 77) if op == "find" else gm.count(asks[g])
                assert np.array_equal(got.astype(exp.dtype), exp)
            exp_hits = sum(int(gm.count(a).sum()) for a in asks)
            assert sum(hits) == exp_hits
    finally:
        for h in peers:
            if h:
                lib.khx_peer_destroy(h)


@pytest.mark.parametrize("max_load", [0.8, 0.97])
def test_tile_insert_deferred_chains_and_side_key(kh, oracle_mod, max_load):
    """shared-memory tile insert (kh_tile.cuh): chains that leave their 8192-slot tile and the all-ones key are
    deferred to the global kernel; a second large batch goes into the live (non-fresh) table; after an erase
    (sticky bits) inserts fall back to the global path.  Contents and answers equal the oracle's throughout."""
    rng = np.random.default_rng(29)
    n_uniq = int(4 * 8192 * max_load * 0.95)                 # fills 4 tiles close to max_load
    uniq = rng.integers(0, 2**62, size=n_uniq, dtype=np.uint64)
    keys = uniq[rng.integers(0, n_uniq, size=6 * n_uniq)]
    keys = np.ascontiguousarray(np.concatenate([keys, uniq, np.full(7, 2**64 - 1, dtype=np.uint64)]))
    m = kh.HashMap(1, "MURMUR64avx", "COUNT", min_load=0.1, max_load=max_load)
    m.set_insert_window(True, 1 << 12, 1)
    om = oracle_mod.port.Map(1, "COUNT")
    m.reserve(n_uniq)
    assert m.capacity() == 4 * 8192
    assert m.insert_no_estimate(keys, default=1) == om.insert(keys, default=1)     # fresh table
    check_against(m, om)
    more = np.ascontiguousarray(np.concatenate([keys[:40000], rng.integers(0, 2**62, size=300, dtype=np.uint64)]))
    assert m.insert_no_estimate(more, default=3) == om.insert(more, default=3)     # live table
    check_against(m, om)
    q = np.ascontiguousarray(keys[:3000])
    assert m.erase(q) == om.erase(q)
    assert m.insert_no_estimate(more, default=1) == om.insert(more, default=1)     # sticky bits: global path
    check_against(m, om)
    assert np.array_equal(m.find(q, 0), om.find(q, 0))


@pytest.mark.gpu
@pytest.mark.parametrize("n,distinct,bulk", [(3_000_000, 700_000, True), (3_000_000, 2_500_000, True), (200_000, 90_000, False)])
def test_capacity_follows_compiled_reference_over_overlapping_batches(kh, oracle_mod, n, distinct, bulk):
    """insert() reserves from ONE cumulative HyperLogLog over everything inserted with an estimate, as the reference does
    (KH/robinhood_offset_hashmap_ptr.hpp:291,2520-2535): re-inserting overlapping batches (the k-mer counting case) must
    leave the capacity where the compiled reference leaves it, batch after batch -- bulk route and general route."""
    ref = getattr(oracle_mod, "ref", None)
    if ref is None:
        pytest.skip("oracle/_ref not built")
    import torch
    rng = np.random.default_rng(n + distinct)
    base = rng.integers(0, 2**62, size=distinct, dtype=np.uint64)
    m = kh.HashMap(1, "MURMUR64avx", "COUNT", min_load=0.2, max_load=0.8)
    m.set_insert_window(bulk, 32 << 20, 1 << 20)
    rm = ref.Map("OFFSETS_COUNT", 1, "MURMUR64avx", 0.2, 0.8)
    for batch in range(4):
        lo = 0 if batch != 2 else distinct // 2                        # batch 2 draws from the upper half only
        keys = np.ascontiguousarray(base[rng.integers(lo, distinct, size=n)])
        m.insert(torch.from_numpy(keys.view(np.int64)).cuda(), default=1)
        rm.insert(keys, default=1)
        assert m.size() == rm.size()
        assert m.capacity() == rm.capacity(), (batch, m.capacity(), rm.capacity())


@pytest.mark.parametrize("W,lp,skew", [(1, 6, 0.0), (1, 7, 0.0), (2, 6, 0.0), (2, 8, 0.0), (1, 6, 0.3), (1, 6, 0.9), (2, 7, 0.3)])
def test_bulk_insert_cluster_region_insert(kh, oracle_mod, monkeypatch, W, lp, skew):
    """tables with more tiles than level-1 regions, up to 16 tiles per region (forced here with 64..256 regions; reached without
    forcing above 2^25 slots): a thread-block cluster per region inserts into its tiles over distributed shared memory
    (region_insert_kernel: clusters of 2, 4 and 8 CTAs here), 8- and 16-byte keys, a second batch into the filled table, skewed
    batches (overflow list; at 90 % the general route takes over)"""
    import torch
    monkeypatch.setenv("KH_BULK_LP", str(lp))
    monkeypatch.setenv("KH_CLUSTER", "1")
    o = oracle_mod.port
    rng = np.random.default_rng(170 + W + lp)
    distinct, n = 1_200_000, 2_500_000
    base = rng.integers(0, 2**62, size=(distinct, W), dtype=np.uint64)
    m = kh.HashMap(W, "MURMUR64avx", "COUNT")
    m.set_insert_window(True, 32 << 20, 1 << 20)
    gm = o.Map(W, "COUNT")
    for batch in range(2):
        keys = base[rng.integers(0, distinct, size=n)]
        if skew:
            keys[rng.random(n) < skew] = base[7]
        keys = np.ascontiguousarray(keys if W == 2 else keys[:, 0])
        new = m.insert(torch.from_numpy(keys.view(np.int64)).cuda(), default=1)
        before = gm.size()
        gm.insert(keys, default=1)
        assert new == gm.size() - before and m.size() == gm.size()
        ek, ev = sort_pairs(*gm.to_vector())
        tk, tv = sort_pairs(*m.to_vector())
        assert np.array_equal(tk, ek) and np.array_equal(tv, ev)


@pytest.mark.parametrize("W,exact,skew", [(1, "0", 0.0), (1, "1", 0.0), (2, "0", 0.0), (2, "1", 0.0), (1, "0", 0.3), (1, "0", 0.9)])
def test_bulk_insert_level2_split(kh, oracle_mod, monkeypatch, W, exact, skew):
    """tables with more tiles than level-1 regions (forced here with 64 regions; reached without forcing above 2^25 slots):
    the level-2 split of every region into its tiles, in slab mode (default: no histogram pass, overflow list) and in exact
    mode (histogram -> scan -> scatter), 8- and 16-byte keys, a second batch into the filled table, and skewed batches whose
    hot key overflows its tile's slab (30 %) or the overflow list itself (90 %: the general route takes over)."""
    import torch
    monkeypatch.setenv("KH_BULK_LP", "6")
    monkeypatch.setenv("KH_L2_EXACT", exact)
    monkeypatch.setenv("KH_CLUSTER", "0")
    o = oracle_mod.port
    rng = np.random.default_rng(77 + W)
    distinct, n = 1_200_000, 2_500_000
    base = rng.integers(0, 2**62, size=(distinct, W), dtype=np.uint64)
    m = kh.HashMap(W, "MURMUR64avx", "COUNT")
    m.set_insert_window(True, 32 << 20, 1 << 20)
    gm = o.Map(W, "COUNT")
    for batch in range(2):
        keys = base[rng.integers(0, distinct, size=n)]
        if skew:
            keys[rng.random(n) < skew] = base[7]
        keys = np.ascontiguousarray(keys if W == 2 else keys[:, 0])
        new = m.insert(torch.from_numpy(keys.view(np.int64)).cuda(), default=1)
        before = gm.size()
        gm.insert(keys, default=1)
        assert new == gm.size() - before and m.size() == gm.size()
        assert skew >= 0.5 or m.capacity() >= 128 * 8192 // W        # more than 64 tiles: level 2 ran
        ek, ev = sort_pairs(*gm.to_vector())
        tk, tv = sort_pairs(*m.to_vector())
        assert np.array_equal(tk, ek) and np.array_equal(tv, ev)


@pytest.mark.parametrize("p,k,canonical", [(2, 31, True), (8, 31, True), (4, 21, False), (2, 63, True), (8, 64, False)])
def test_kmer_generation_fused_with_sender(kh, oracle_mod, p, k, canonical):
    """khx_kmers_scatter_slabs_dev (k-mer tile kernel in SCATTER mode; p ranks emulated on one GPU): slab (owner r, sender g)
    must hold exactly the k-mers of g's reads that the reference's rule sends to r -- as a multiset, order is free -- and the
    counts must be the true ones; reads with N, EOL characters and lower case, a misaligned base pointer"""
    import ctypes
    import torch
    from conftest import make_reads
    from kmerhash_b200._lib import lib
    o = oracle_mod.port
    W = 1 if k <= 32 else 2
    cudart = ctypes.CDLL("libcudart.so")
    rng = np.random.default_rng(600 + p + k)
    inputs = [make_reads(rng, 900, 60, 260, n_rate=0.002, eol=(g % 2 == 1), lower=True) for g in range(p)]
    n_max = max(len(b) for b, _ in inputs)
    tables = [kh.HashMap(W, "MURMUR64avx", "COUNT") for _ in range(p)]
    slab_cap = (int(n_max / p * 1.6) + 4096 + 63) // 64 * 64
    cap = 2 * p * slab_cap + 64
    peers = [lib.khx_peer_create(0, r, p, cap, W) for r in range(p)]

    def read_buf(r):
        buf = torch.empty((cap, W), dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        assert cudart.cudaMemcpy(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(lib.khx_peer_recv_ptr(peers[r])),
                                 ctypes.c_size_t(cap * 8 * W), 3) == 0
        return buf.cpu().numpy().view(np.uint64)

    def srt(a):
        a = a.reshape(-1, W)
        return np.sort(a[:, 0]) if W == 1 else a[np.lexsort(a.T[::-1])]

    try:
        assert all(peers)
        for r in range(p):
            for q in range(p):
                if q != r:
                    assert lib.khx_peer_set(peers[r], q, lib.khx_peer_recv_ptr(peers[q])) == 0
        parity = 1
        counts = []
        for g, (bases, offs) in enumerate(inputs):
            shift = 5 * g + 1                                              # misaligned base pointer, offsets not starting at 0
            db = torch.from_numpy(np.concatenate([np.full(shift, 65, dtype=np.uint8), bases])).cuda()
            do = torch.from_numpy((offs + np.uint64(shift)).astype(np.int64)).cuda()
            c = torch.zeros(p, dtype=torch.int64, device="cuda")
            assert tables[g].kmers_scatter_slabs(db, do, k, canonical, peers[g], slab_cap, parity, c)
            counts.append(c.cpu().numpy())
        for r in range(p):
            got = read_buf(r)
            for g, (bases, offs) in enumerate(inputs):
                km = o.kmers_from_reads(bases, offs, k, canonical)
                own = o.rank("MURMUR64avx", km, p)
                exp = km[own == r]
                assert int(counts[g][r]) == len(exp)
                lo = (parity * p + g) * slab_cap
                assert np.array_equal(srt(got[lo:lo + len(exp)]), srt(exp)), (g, r)
    finally:
        for h in peers:
            if h:
                lib.khx_peer_destroy(h)


@pytest.mark.parametrize("W", [1, 2])
@pytest.mark.parametrize("red", ["FIRST", "LAST"])
@pytest.mark.parametrize("with_vals,lp", [(True, None), (False, None), (True, 6)])
def test_ordered_bulk_insert(kh, oracle_mod, monkeypatch, W, red, with_vals, lp):
    """FIRST / LAST tables through the bulk route (l1_scatter_kernel<W, true> records with batch positions ->
    tile_insert_ord_kernel): the value a key keeps is the one of its first / last occurrence in batch order, over three
    batches (fresh table, live table, live table after growth), with and without per-key values, with the level-2 split forced
    (lp = 6), against the oracle.  The all-ones key (side slot) and repeated keys at both ends of the batch are included."""
    import torch
    if lp is not None:
        monkeypatch.setenv("KH_BULK_LP", str(lp))
    o = oracle_mod.port
    rng = np.random.default_rng(500 + W + (red == "LAST") * 10)
    distinct, n = 900_000, 2_000_000
    base = rng.integers(0, 2**62, size=(distinct, W), dtype=np.uint64)
    base[5] = 2**64 - 1
    m = kh.HashMap(W, "MURMUR64avx", red)
    m.set_insert_window(True, 32 << 20, 1 << 20)
    m.kernel_timing(True)
    gm = o.Map(W, red)
    for batch in range(3):
        hi = distinct if batch else distinct // 2
        keys = base[rng.integers(0, hi, size=n)]
        keys[:3] = base[5]; keys[-3:] = base[5]; keys[7] = base[11]; keys[-7] = base[11]
        keys = np.ascontiguousarray(keys if W == 2 else keys[:, 0])
        vals = rng.integers(1, 2**31, size=n).astype(np.uint32) if with_vals else None
        dk = torch.from_numpy(keys.view(np.int64)).cuda()
        dv = torch.from_numpy(vals.view(np.int32)).cuda() if with_vals else None
        before = gm.size()
        new = m.insert(dk, dv, default=batch + 2)
        gm.insert(keys, vals, default=batch + 2)
        assert new == gm.size() - before and m.size() == gm.size()
        ek, ev = sort_pairs(*gm.to_vector())
        tk, tv = sort_pairs(*m.to_vector())
        assert np.array_equal(tk, ek) and np.array_equal(tv, ev)
    assert m.kernel_time("scatter")[1] >= 3                        # the bulk route ran for every batch (level-1 launches)


@pytest.mark.parametrize("red", ["FIRST", "LAST"])
def test_ordered_bulk_deferred_keys_keep_batch_order(kh, oracle_mod, red):
    """tiles filled close to max_load 0.97: chains leave their tile, the deferred records go to the general kernel in batch
    order with their values (sorted by the position the records carry); a second batch goes into the live table"""
    import torch
    rng = np.random.default_rng(31)
    n_uniq = int(4 * 8192 * 0.97 * 0.95)
    uniq = rng.integers(0, 2**62, size=n_uniq, dtype=np.uint64)
    keys = np.ascontiguousarray(np.concatenate([uniq[rng.integers(0, n_uniq, size=6 * n_uniq)], uniq]))
    m = kh.HashMap(1, "MURMUR64avx", red, min_load=0.1, max_load=0.97)
    m.set_insert_window(True, 1 << 12, 1)
    om = oracle_mod.port.Map(1, red)
    m.reserve(n_uniq)
    for batch in range(2):
        vals = rng.integers(1, 2**31, size=keys.shape[0]).astype(np.uint32)
        assert m.insert_no_estimate(keys, vals) == om.insert(keys, vals)
        check_against(m, om)
        keys = np.ascontiguousarray(np.concatenate([keys[:50000], rng.integers(0, 2**62, size=200, dtype=np.uint64)]))


@pytest.mark.parametrize("red,max_load", [("COUNT", 0.5), ("COUNT", 0.9), ("FIRST", 0.7), ("LAST", 0.8)])
def test_bulk_insert_no_estimate_grows_to_the_doubling_schedules_capacity(kh, oracle_mod, red, max_load):
    """insert_no_estimate from capacity 128 (config 2's doubling mode) on the bulk route: the table grows at once to the capacity
    the doubling schedule is bound to reach (lower bound of the batch's HyperLogLog estimate) instead of doubling 18 times; the
    final capacity is the schedule's -- the smallest 128 * 2^j whose max_load bound holds the size -- and the contents are
    the oracle's.  A second batch (mostly known keys) stays in place."""
    import torch
    rng = np.random.default_rng(901)
    distinct, n = 1_500_000, 4_000_000
    base = rng.integers(0, 2**62, size=distinct, dtype=np.uint64)
    m = kh.HashMap(1, "MURMUR64avx", red, min_load=0.2, max_load=max_load)
    m.set_insert_window(True, 32 << 20, 1 << 20)
    m.kernel_timing(True)
    om = oracle_mod.port.Map(1, red)
    assert m.capacity() == 128
    for batch in range(2):
        keys = np.ascontiguousarray(base[rng.integers(0, distinct if batch == 0 else distinct // 3, size=n)])
        vals = None if red == "COUNT" else rng.integers(1, 2**31, size=n).astype(np.uint32)
        dk = torch.from_numpy(keys.view(np.int64)).cuda()
        dv = None if vals is None else torch.from_numpy(vals.view(np.int32)).cuda()
        assert m.insert_no_estimate(dk, dv) == om.insert(keys, vals)
        assert m.size() == om.size()
        cap = 128
        while m.size() > max_load * cap:
            cap *= 2
        assert m.capacity() == cap
        ek, ev = sort_pairs(*om.to_vector())
        tk, tv = sort_pairs(*m.to_vector())
        assert np.array_equal(tk, ek) and np.array_equal(tv, ev)
    assert m.kernel_time("scatter")[1] >= 2 and m.kernel_time("rebuild")[1] <= 2
