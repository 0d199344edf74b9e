"""CPU tier: the C-ABI library loads, exports every symbol include/kmerhash_b200.h declares, fails loudly
without a GPU (no CPU fallback), and the arithmetic the kernels inline (compiled for the host from the very
same headers, tests/host_check.cu) is bit-exact to the oracle and the golden vectors."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "kmerhash_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(kh[dx]?_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(kh):
    names = declared_functions()
    assert len(names) >= 50
    lib = C.CDLL(kh.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # and the python binding covers every one of them
    from kmerhash_b200._lib import SIGNATURES
    assert sorted(SIGNATURES) == names


def test_product_never_touches_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "kmerhash_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "libkmerhash_oracle" not in text and "libkmerhash_ref" not in text, f


def test_no_gpu_fails_loudly(kh):
    if kh.device_count() > 0:
        pytest.skip("a GPU is visible")
    with pytest.raises(kh.KmerHashError):
        kh.HashMap()
    with pytest.raises(kh.KmerHashError):
        kh.hash_batch(np.arange(8, dtype=np.uint64))
    with pytest.raises(kh.KmerHashError):
        kh.kmers_from_reads(b"ACGTACGT", np.array([0, 8], dtype=np.uint64), 3)
    with pytest.raises(kh.KmerHashError):
        kh.MultiGpuMap(1)
    assert "no CPU fallback" in kh.last_error() or "CUDA" in kh.last_error()


@pytest.fixture(scope="module")
def hostlib():
    path = os.path.join(ROOT, "tests", "libkh_host_check.so")
    if not os.path.exists(path):
        import __graft_entry__
        __graft_entry__.build_host_check()
    L = C.CDLL(path)
    u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
    L.khtest_hash.argtypes = [C.c_int, C.c_uint32, C.c_int, u64p, C.c_size_t, u64p]
    L.khtest_kmer.argtypes = [C.c_int, C.c_int, u64p, C.c_size_t, u64p]
    return L


def test_kernel_arithmetic_on_host_vs_golden(hostlib, golden, oracle_mod):
    k1, k2 = golden["hash_keys1"], golden["hash_keys2"]
    for name, kid in oracle_mod.HASH_KINDS.items():
        if name == "MURMUR32FINALIZER":
            continue
        for seed in (43, 37, 9876543):
            for W, keys in ((1, k1), (2, k2)):
                out = np.empty(keys.shape[0], dtype=np.uint64)
                assert hostlib.khtest_hash(kid, seed, W, np.ascontiguousarray(keys), keys.shape[0], out) == 0
                assert np.array_equal(out, golden["hash_%s_%d_w%d" % (name, seed, W)]), (name, seed, W)
    for k in (5, 15, 21, 31, 33, 47, 63):
        w = np.ascontiguousarray(golden["kmer_windows_k%d" % k])
        for mode, key in ((0, "kmer_canon_k%d"), (1, "kmer_revcomp_k%d")):
            out = np.empty_like(w)
            hostlib.khtest_kmer(k, mode, w, w.shape[0], out)
            assert np.array_equal(out, golden[key % k]), (k, mode)


def test_kernel_arithmetic_on_host_vs_oracle_random(hostlib, oracle_mod):
    rng = np.random.default_rng(99)
    port = oracle_mod.port
    for W in (1, 2):
        keys = rng.integers(0, 2**64, size=(20000, W), dtype=np.uint64)
        kk = keys if W == 2 else keys[:, 0].copy()
        for name, kid in oracle_mod.HASH_KINDS.items():
            if name == "MURMUR32FINALIZER":
                continue
            out = np.empty(20000, dtype=np.uint64)
            hostlib.khtest_hash(kid, 9876543, W, keys, 20000, out)
            assert np.array_equal(out, port.hash(name, kk, 9876543)), name
    for k in (1, 2, 16, 31, 32, 33, 48, 63, 64):
        W = (2 * k + 63) // 64
        km = rng.integers(0, 2**64, size=(5000, W), dtype=np.uint64)
        top = 2 * k - 64 * (W - 1)
        if top < 64:
            km[:, W - 1] &= np.uint64((1 << top) - 1)
        out = np.empty_like(km)
        hostlib.khtest_kmer(k, 0, km, 5000, out)
        exp = port.canonical(k, km if W == 2 else km[:, 0].copy())
        assert np.array_equal(out.reshape(exp.shape), exp), k


def test_alphabet_arithmetic_on_host_vs_golden(hostlib):
    """DNA5 / DNA16: the kernels' own helpers (kh_kmer.cuh: alpha_code, kmer_revcomp_alpha = bit reversal,
    kmer_canonical_alpha), compiled for the host, against the vectors the compiled reference produced
    (tests/golden/alphabet_vectors.npz, KI/common/alphabets.hpp:197-292, KI/common/kmer.hpp:1476-1497)"""
    u64p = np.ctypeslib.ndpointer(np.uint64, flags="C")
    hostlib.khtest_kmer_alpha.argtypes = [C.c_int, C.c_int, C.c_int, u64p, C.c_size_t, u64p]
    g = dict(np.load(os.path.join(ROOT, "tests", "golden", "alphabet_vectors.npz")))
    text = g["text"]
    for alpha, k in ((5, 7), (5, 21), (5, 31), (5, 42), (16, 15), (16, 17), (16, 31), (16, 32)):
        w = np.ascontiguousarray(g["windows_a%d_k%d" % (alpha, k)])
        W = 1 if w.ndim == 1 else w.shape[1]
        assert hostlib.khtest_kmer_words_alpha(alpha, k) == W
        for mode, name in ((1, "revcomp"), (0, "canonical")):
            out = np.empty_like(w)
            assert hostlib.khtest_kmer_alpha(alpha, k, mode, w, w.shape[0], out) == 0
            assert np.array_equal(out, g["%s_a%d_k%d" % (name, alpha, k)]), (alpha, k, name)
        # the first window, rebuilt from the code table: base i of the window sits B * (k - 1 - i) bits up
        B = 3 if alpha == 5 else 4
        v = 0
        for i in range(k):
            v = (v << B) | hostlib.khtest_alpha_code(alpha, int(text[i]))
        first = int(w[0]) if W == 1 else int(w[0, 0]) | (int(w[0, 1]) << 64)
        assert v == first, (alpha, k)


def test_murmur32fin_rejected(kh):
    # murmur3finalizer_avx32 is defined for keys <= 4 bytes only (murmurhash3finalizer_32_avx.hpp:527)
    with pytest.raises((ValueError, kh.KmerHashError)):
        kh.hash_batch(np.arange(4, dtype=np.uint64), "MURMUR32FINALIZER")


def _serial_kmers(path, k, oracle_mod):
    """serial parse of a FASTA / FASTQ file (one read per record) -> the oracle's k-mers, in order"""
    import numpy as np
    reads, cur, fastq = [], None, None
    lines = open(path, "rb").read().split(b"\n")
    i = 0
    while i < len(lines):
        ln = lines[i]
        if fastq is None and ln[:1] in (b"@", b">"):
            fastq = ln[:1] == b"@"
        if fastq:
            if ln[:1] == b"@" and i + 2 < len(lines) and lines[i + 2][:1] == b"+":
                reads.append(lines[i + 1].rstrip(b"\r")); i += 4; continue
        elif ln[:1] == b">":
            if cur is not None:
                reads.append(cur)
            cur = b""
        elif cur is not None:
            cur += ln + b"\n"
        i += 1
    if cur is not None:
        reads.append(cur)
    bases = np.frombuffer(b"".join(reads), dtype=np.uint8)
    offs = np.concatenate([[0], np.cumsum([len(r) for r in reads])]).astype(np.uint64)
    return oracle_mod.port.kmers_from_reads(bases, offs, k, True)


@pytest.mark.parametrize("kind", ["fasta", "fastq"])
def test_partitioned_file_reader_of_the_kmercounter_driver(tmp_path, oracle_mod, kind):
    """drivers/BenchmarkKmerCounter reads files the reference's way (byte ranges, record-boundary search, k - 1 overlap for
    FASTA sequences cut at a range border, KI/io/kmer_file_helper.hpp:407-645): whatever the number of ranges, the reads
    it hands to the k-mer kernel must yield exactly the k-mers of a serial parse, in the same order (checked with the
    oracle's generator on the host; no GPU involved)."""
    import subprocess
    import numpy as np
    exe = os.path.join(ROOT, "drivers", "BenchmarkKmerCounter")
    rng = np.random.default_rng(5)
    path = str(tmp_path / ("in." + kind))
    with open(path, "w") as f:
        if kind == "fasta":
            for r in range(7):
                n = int(rng.integers(1, 9000)) if r != 3 else 10
                seq = "".join("ACGTN"[c] for c in rng.choice(5, size=n, p=[0.2495, 0.2495, 0.2495, 0.2495, 0.002]))
                f.write(">chr%d some description that is a little longer than usual %s\n" % (r, "x" * int(rng.integers(0, 90))))
                w = int(rng.integers(40, 80))
                for i in range(0, n, w):
                    f.write(seq[i:i + w] + ("\r\n" if r == 5 else "\n"))
                if r == 2:
                    f.write("\n")
        else:
            for r in range(900):
                n = int(rng.integers(1, 160))
                seq = "".join("ACGTN"[c] for c in rng.choice(5, size=n, p=[0.2475, 0.2475, 0.2475, 0.2475, 0.01]))
                qual = "".join(chr(c) for c in rng.integers(33, 74, size=n))
                if r % 5 == 0:
                    qual = "@" + qual[1:]            # a quality line that looks like a header
                if r % 7 == 0:
                    qual = "+" + qual[1:]
                f.write("@read%d/1\n%s\n+\n%s\n" % (r, seq, qual))
    exp = _serial_kmers(path, 31, oracle_mod)
    assert len(exp) > 1000
    for threads in (1, 2, 3, 7, 16, 61):
        out = str(tmp_path / "reads.bin")
        r = subprocess.run([exe, "--reads-only", out, path], capture_output=True, text=True, timeout=120,
                           env=dict(os.environ, KH_READ_THREADS=str(threads)))
        assert r.returncode == 0, r.stdout + r.stderr
        raw = open(out, "rb").read()
        nr = int(np.frombuffer(raw[:8], dtype=np.uint64)[0])
        offs = np.frombuffer(raw[8:8 + 8 * (nr + 1)], dtype=np.uint64)
        bases = np.frombuffer(raw[8 + 8 * (nr + 1):], dtype=np.uint8)
        got = oracle_mod.port.kmers_from_reads(bases, offs, 31, True)
        assert got.shape == exp.shape and np.array_equal(got, exp), (kind, threads)
