"""GPU tier: batched hashes, k-mer generation, canonicalisation and rank partitioning through the C ABI,
bit-exact against the oracle and the committed golden vectors."""
import numpy as np
import pytest

from conftest import READ78, make_reads

pytestmark = pytest.mark.gpu

HASHES = ["IDEN", "MURMUR64", "MURMUR64avx", "MURMUR32", "MURMUR32avx", "MURMUR32sse", "FARM", "FARM32",
          "CLHASH", "CRC32C", "MURMUR_X86"]


@pytest.mark.parametrize("name", HASHES)
def test_hash_golden(kh, golden, name):
    for seed in (43, 37, 9876543):
        assert np.array_equal(kh.hash_batch(golden["hash_keys1"], name, seed), golden["hash_%s_%d_w1" % (name, seed)])
        assert np.array_equal(kh.hash_batch(golden["hash_keys2"], name, seed), golden["hash_%s_%d_w2" % (name, seed)])


@pytest.mark.parametrize("W", [1, 2])
def test_hash_random_vs_oracle(kh, oracle_mod, W):
    rng = np.random.default_rng(3 + W)
    for n in (0, 1, 31, 1000, 200003):          # empty, tiny, ragged tails
        keys = rng.integers(0, 2**64, size=(n, W), dtype=np.uint64)
        kk = keys if W == 2 else keys[:, 0].copy()
        for name in HASHES:
            got = kh.hash_batch(kk, name, 9876543)
            assert np.array_equal(got, oracle_mod.port.hash(name, kk, 9876543)), (name, n)


def test_seed_changes_hash(kh):
    """the reference's own hash tests check seed sensitivity (test/unit/test_kmer_hash.cpp:291-424)"""
    keys = np.arange(1, 2000, dtype=np.uint64)
    for name in ("MURMUR64", "MURMUR64avx", "MURMUR32", "FARM", "CLHASH", "CRC32C"):
        a, b = kh.hash_batch(keys, name, 43), kh.hash_batch(keys, name, 44)
        assert not np.array_equal(a, b), name
        assert len(np.unique(a)) == len(a), name     # no collisions among adjacent keys


@pytest.mark.parametrize("k", [5, 15, 21, 31, 33, 47, 63])
def test_kmer_golden(kh, golden, k):
    bases = np.frombuffer(READ78.encode(), dtype=np.uint8)
    offs = np.array([0, len(READ78)], dtype=np.uint64)
    assert np.array_equal(kh.kmers_from_reads(bases, offs, k, False), golden["kmer_windows_k%d" % k])
    assert np.array_equal(kh.kmers_from_reads(bases, offs, k, True), golden["kmer_canon_k%d" % k])
    w = golden["kmer_windows_k%d" % k]
    assert np.array_equal(kh.revcomp(w, k), golden["kmer_revcomp_k%d" % k])
    assert np.array_equal(kh.canonicalize(w, k), golden["kmer_canon_k%d" % k])


@pytest.mark.parametrize("k", [1, 2, 16, 31, 32, 33, 48, 63, 64])
@pytest.mark.parametrize("canonical", [False, True])
def test_kmers_from_reads_vs_oracle(kh, oracle_mod, k, canonical):
    rng = np.random.default_rng(100 + k)
    cases = [
        make_reads(rng, 300, 20, 260),                                  # ragged, some shorter than k
        make_reads(rng, 200, 100, 300, n_rate=0.02),                    # N splitting
        make_reads(rng, 200, 100, 300, n_rate=0.01, eol=True, lower=True),   # CR/LF inside reads, lower case
        make_reads(rng, 3, 5000, 9000, n_rate=0.001, eol=True),         # long sequences
        make_reads(rng, 50, 0, 3),                                      # empty / tiny reads
    ]
    for bases, offs in cases:
        exp = oracle_mod.port.kmers_from_reads(bases, offs, k, canonical)
        got = kh.kmers_from_reads(bases, offs, k, canonical)
        assert got.shape == exp.shape, (k, got.shape, exp.shape)
        assert np.array_equal(got, exp), k


def test_kmers_from_reads_other_bytes_map_to_A(kh, oracle_mod):
    read = b"ACGTRYKMACGTXXACGTACGT-ACGT.ACGT"
    offs = np.array([0, len(read)], dtype=np.uint64)
    for k in (3, 7):
        assert np.array_equal(kh.kmers_from_reads(read, offs, k), oracle_mod.port.kmers_from_reads(read, offs, k, False))


@pytest.mark.parametrize("k,canonical", [(31, True), (21, False), (32, True), (33, True), (63, True), (64, False), (1, False)])
def test_kmers_tile_kernel_large_inputs(kh, oracle_mod, k, canonical):
    """the streaming tile kernel (kh_kmergen.cuh) over inputs of many tiles: reads straddling tile borders, N runs, EOL
    characters, empty reads, a base pointer that is not 16-byte aligned and offsets that do not start at 0; device and
    host entry points; output identical to the oracle's, in the reference's order"""
    import torch
    rng = np.random.default_rng(1000 + k)
    cases = [make_reads(rng, 3000, 100, 151),                               # Illumina-like, 450 KB
             make_reads(rng, 400, 0, 2500, n_rate=0.003, eol=True, lower=True),
             make_reads(rng, 2, 150000, 200000, n_rate=0.0005, eol=True),   # chromosomes: one read over many tiles
             make_reads(rng, 20000, 0, 40)]                                 # many reads per tile, lots of them shorter than k
    for bases, offs in cases:
        exp = oracle_mod.port.kmers_from_reads(bases, offs, k, canonical)
        got = kh.kmers_from_reads(bases, offs, k, canonical)
        assert got.shape == exp.shape and np.array_equal(got, exp), k
        # device entry: junk in front (misaligned start), offsets[0] = shift
        for shift in (3, 16, 37):
            pad = rng.integers(65, 85, size=shift).astype(np.uint8)
            db = torch.from_numpy(np.concatenate([pad, bases])).cuda()
            do = torch.from_numpy((offs + np.uint64(shift)).astype(np.int64)).cuda()
            out = kh.kmers_from_reads_dev(db, do, k, canonical).cpu().numpy().view(np.uint64)
            assert out.shape == exp.shape and np.array_equal(out, exp), (k, shift)
            out2 = kh.kmers_from_reads_dev(db[shift:], torch.from_numpy(offs.astype(np.int64)).cuda(), k, canonical).cpu().numpy().view(np.uint64)
            assert np.array_equal(out2, exp), (k, shift)


def test_kmers_tile_kernel_eol_dense_falls_back(kh, oracle_mod):
    """more EOL characters inside a window than the tile kernel's context holds: the call must still be exact (it repeats
    through the warp-per-read kernel)"""
    rng = np.random.default_rng(5)
    seq = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=30000)]
    dense = np.full(12 * len(seq), 10, dtype=np.uint8)
    dense[0::12] = seq; dense[5::12] = 13                                     # every base followed by eleven EOL characters
    offs = np.array([0, len(dense) // 2, len(dense)], dtype=np.uint64)
    for k in (31, 63):
        exp = oracle_mod.port.kmers_from_reads(dense, offs, k, True)
        assert len(exp) > 0
        assert np.array_equal(kh.kmers_from_reads(dense, offs, k, True), exp)


@pytest.mark.parametrize("alphabet,ks", [("DNA5", (7, 21, 31, 42)), ("DNA16", (15, 17, 31, 32))])
def test_kmers_other_alphabets(kh, oracle_mod, alphabet, ks):
    """N3 of SURVEY.md 8f, the part that fits two key words: DNA5 (3 bits, the reference's DNA6 table) and DNA16 (4 bits, IUPAC)
    k-mers -- generation from reads (N-split, EOL skipped, many tiles, also the EOL-dense fallback), reverse complement (= bit
    reversal, KI/common/kmer.hpp:1476-1497) and canonical form against vectors of the COMPILED reference and against the
    oracle; the k-mers then count in a 16-byte-key table like any key"""
    import os
    al = {"DNA5": 5, "DNA16": 16}[alphabet]
    g = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "alphabet_vectors.npz")))
    o = oracle_mod.port
    rng = np.random.default_rng(al)
    sym = np.frombuffer(b"ACGTACGTACGTacgtMRWSYKVHDBU.-X?", dtype=np.uint8)
    for k in ks:
        w = g["windows_a%d_k%d" % (al, k)]
        assert np.array_equal(kh.canonicalize_alpha(w, k, alphabet, revcomp_only=True), g["revcomp_a%d_k%d" % (al, k)]), k
        assert np.array_equal(kh.canonicalize_alpha(w, k, alphabet), g["canonical_a%d_k%d" % (al, k)]), k
        cases = [make_reads(rng, 400, 0, 300, n_rate=0.004, eol=True, lower=True), make_reads(rng, 2, 30000, 50000, n_rate=0.0005, eol=True)]
        # IUPAC / gap symbols sprinkled in
        for bases, offs in cases:
            hit = rng.random(len(bases)) < 0.05
            keep = (bases != 10) & (bases != 13) & (bases != ord("N")) & (bases != ord("n"))
            bases[hit & keep] = sym[rng.integers(0, len(sym), size=int((hit & keep).sum()))]
            for canonical in (False, True):
                exp = o.kmers_from_reads_alpha(al, bases, offs, k, canonical)
                got = kh.kmers_from_reads(bases, offs, k, canonical, alphabet=alphabet)
                assert got.shape == exp.shape and np.array_equal(got, exp), (alphabet, k, canonical)
    # EOL-dense input -> serial fallback
    k = ks[2]
    seq = sym[rng.integers(0, len(sym), size=6000)]
    dense = np.full(16 * len(seq), 10, dtype=np.uint8)
    dense[0::16] = seq
    offs = np.array([0, len(dense)], dtype=np.uint64)
    exp = o.kmers_from_reads_alpha(al, dense, offs, k, True)
    assert len(exp) > 0 and np.array_equal(kh.kmers_from_reads(dense, offs, k, True, alphabet=alphabet), exp)
    # the table counts them like any 16-byte key
    bases, offs = make_reads(rng, 3000, 100, 150)
    km = kh.kmers_from_reads(bases, offs, 31, True, alphabet=alphabet)
    assert km.ndim == 2 and km.shape[1] == 2
    m = kh.HashMap(2, "MURMUR64avx", "COUNT")
    gm = o.Map(2, "COUNT")
    assert m.insert(km, default=1) == gm.insert(km, default=1)
    ek, ev = gm.to_vector()
    tk, tv = m.to_vector()
    a, b = np.lexsort(ek.T[::-1]), np.lexsort(tk.T[::-1])
    assert np.array_equal(ek[a], tk[b]) and np.array_equal(np.asarray(ev)[a], np.asarray(tv)[b])


def test_kmers_no_reads(kh):
    out = kh.kmers_from_reads(b"", np.array([0], dtype=np.uint64), 31)
    assert out.shape == (0,)


@pytest.mark.parametrize("p", [1, 2, 3, 8, 13])
@pytest.mark.parametrize("W", [1, 2])
def test_rank_golden_and_partition(kh, golden, oracle_mod, p, W):
    keys = golden["hash_keys%d" % W]
    if p > 1:
        for name in ("MURMUR64avx", "MURMUR64", "MURMUR32", "FARM", "CLHASH", "CRC32C", "IDEN"):
            assert np.array_equal(kh.rank_of(keys, p, name), golden["rank_%s_p%d_w%d" % (name, p, W)]), name
    rng = np.random.default_rng(p * 10 + W)
    for n in (0, 1, 255, 256, 257, 100000, 700001):
        k = rng.integers(0, 2**64, size=(n, W), dtype=np.uint64)
        kk = k if W == 2 else k[:, 0].copy()
        ranks = oracle_mod.port.rank("MURMUR64avx", kk, p)
        out, perm, counts = kh.partition(kk, p, "MURMUR64avx")
        assert counts == np.bincount(ranks, minlength=p).tolist()
        # stable: each destination segment holds its keys in input order (permute_by_bucketid semantics)
        exp_perm = np.argsort(ranks, kind="stable").astype(np.uint32)
        assert np.array_equal(perm, exp_perm), (n, p)
        assert np.array_equal(out.reshape(k.shape), k[exp_perm])


@pytest.mark.parametrize("p", [1, 2, 3, 8, 64, 1000])
@pytest.mark.parametrize("W", [1, 2])
def test_fast_partition_segments(kh, oracle_mod, p, W):
    """tile-sorting scatter: same per-destination counts and the same SET of keys per destination as the reference's
    rule; order inside a destination is free, perm maps every output position to its input index."""
    import torch
    rng = np.random.default_rng(p + W)
    for n in (0, 1, 2047, 2048, 2049, 300001):
        k = rng.integers(0, 2**64, size=(n, W), dtype=np.uint64)
        kk = k if W == 2 else k[:, 0].copy()
        ranks = oracle_mod.port.rank("MURMUR64avx", kk, p)
        t = torch.from_numpy(np.ascontiguousarray(kk).view(np.int64)).cuda()
        out, perm, counts = kh.partition_dev(t, p, "MURMUR64avx", want_perm=True, stable=False)
        assert counts == np.bincount(ranks, minlength=p).tolist()
        perm = perm.cpu().numpy().view(np.uint32)
        out = out.cpu().numpy().view(np.uint64).reshape(k.shape)
        assert np.array_equal(np.sort(perm), np.arange(n, dtype=np.uint32))
        assert np.array_equal(out, k[perm])
        assert np.array_equal(ranks[perm], np.repeat(np.arange(p), counts))


def test_unpermute(kh):
    import torch
    rng = np.random.default_rng(1)
    n = 100001
    perm = torch.from_numpy(rng.permutation(n).astype(np.int32)).cuda()
    for dt in (torch.uint8, torch.int32):
        v = torch.from_numpy(rng.integers(0, 200, size=n).astype(np.int32)).to(dt).cuda()
        out = kh.unpermute_dev(v, perm)
        exp = torch.empty_like(v)
        exp[perm.long()] = v
        assert torch.equal(out, exp)
