"""Generates tests/golden/*.npz from the COMPILED REFERENCE (oracle/_ref/libkmerhash_ref.so, i.e. the
reference's own headers built from /root/reference by oracle/Makefile).  Run in the build container:
    python tests/golden/make_golden.py
The vectors travel with the repo; /root/reference and this script's dependency on it do not."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import oracle  # noqa: E402

assert oracle.ref is not None, "build oracle/_ref first (make -C oracle)"
ref = oracle.ref
HERE = os.path.dirname(os.path.abspath(__file__))
READ = "ACGTTGCAAGGCTTAACCGGTACGATCGATTACAGGCATCGATCGGATCCATGCAAGTCGATCGTAAGCTTGACTGAC"


def main():
    rng = np.random.default_rng(20261104)
    out = {}
    # --- hashes: 8- and 16-byte keys, store seeds and the dist seed
    k1 = rng.integers(0, 2**64, size=256, dtype=np.uint64)
    k1[:4] = [0, 1, 0x06c6c6c6c6c6c6c6, 0xFFFFFFFFFFFFFFFF]
    k2 = rng.integers(0, 2**64, size=(256, 2), dtype=np.uint64)
    k2[0] = [0x12936368d4e42d8d, 0x06f90a7c16b18d8f]
    out["hash_keys1"], out["hash_keys2"] = k1, k2
    for name in ["IDEN", "MURMUR64", "MURMUR64avx", "MURMUR32", "MURMUR32avx", "MURMUR32sse", "FARM", "FARM32",
                 "CLHASH", "CRC32C", "MURMUR_X86"]:
        for seed in (43, 37, 9876543):
            out["hash_%s_%d_w1" % (name, seed)] = ref.hash(name, k1, seed)
            out["hash_%s_%d_w2" % (name, seed)] = ref.hash(name, k2, seed)
            # batch (SIMD) mode of the reference must agree with its scalar mode
            assert np.array_equal(ref.hash(name, k1, seed, batch=True), out["hash_%s_%d_w1" % (name, seed)]), name
            assert np.array_equal(ref.hash(name, k2, seed, batch=True), out["hash_%s_%d_w2" % (name, seed)]), name
    # --- k-mers: windows, revcomp, canonical for several k
    for k in (5, 15, 21, 31, 33, 47, 63):
        w = ref.kmer_windows(k, READ)
        out["kmer_windows_k%d" % k] = w
        out["kmer_revcomp_k%d" % k] = ref.revcomp(k, w)
        out["kmer_canon_k%d" % k] = ref.canonical(k, w)
    # --- ranks
    for p in (2, 3, 8, 13):
        for name in ("MURMUR64avx", "MURMUR64", "MURMUR32", "FARM", "CLHASH", "CRC32C", "IDEN"):
            out["rank_%s_p%d_w1" % (name, p)] = ref.rank(name, k1, p)
            out["rank_%s_p%d_w2" % (name, p)] = ref.rank(name, k2, p)
    # --- generate_input + table contents (k=31 and k=63), all reducers, query / erase answers
    for w, tag in ((1, "k31"), (2, "k63")):
        keys, vals = ref.generate_input(w, 3000, 10, canonical=(w == 2))
        out["gen_keys_" + tag], out["gen_vals_" + tag] = keys, vals
        q = np.ascontiguousarray(keys[:600])
        miss = np.ascontiguousarray(keys[:50] ^ np.uint64(0x5))     # mostly absent keys
        for mk, red in (("OFFSETS_COUNT", "count"), ("OFFSETS_FIRST", "first"), ("OFFSETS_LAST", "last")):
            m = ref.Map(mk, w, "MURMUR64avx")
            if red == "count":
                m.insert(keys, default=1)
            else:
                m.insert(keys, vals)
            tk, tv = m.to_vector()
            order = np.lexsort(tk.T[::-1]) if w == 2 else np.argsort(tk, kind="stable")
            out["table_%s_keys_%s" % (red, tag)], out["table_%s_vals_%s" % (red, tag)] = tk[order], tv[order]
            out["find_%s_%s" % (red, tag)] = m.find(q, 0xFFFFFFFF)
            out["count_%s_%s" % (red, tag)] = m.count(q)
            out["count_miss_%s_%s" % (red, tag)] = m.count(miss)
            out["erased_%s_%s" % (red, tag)] = np.array([m.erase(q)], dtype=np.uint64)
            out["count2_%s_%s" % (red, tag)] = m.count(q)
            out["size_after_%s_%s" % (red, tag)] = np.array([m.size()], dtype=np.uint64)
        # HLL registers are not exposed by the reference; its estimate is
        m = ref.Map("OFFSETS_COUNT", w, "MURMUR64avx")
        out["hll_est_" + tag] = np.array([m.hll_estimate(keys)])
    out["rand_words_43"] = ref.rand_words(43, 133)
    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
