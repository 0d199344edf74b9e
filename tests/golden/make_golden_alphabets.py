"""Generates tests/golden/alphabet_vectors.npz from the COMPILED REFERENCE (oracle/_ref: Kmer<k, DNA5 / DNA16, uint64_t>):
sliding windows, reverse complement and canonical form for the (alphabet, k) pairs the harness instantiates.
    python tests/golden/make_golden_alphabets.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import oracle  # noqa: E402

assert oracle.ref is not None, "build oracle/_ref first (make -C oracle)"
HERE = os.path.dirname(os.path.abspath(__file__))
PAIRS = ((5, 7), (5, 21), (5, 31), (5, 42), (16, 15), (16, 17), (16, 31), (16, 32))


def main():
    rng = np.random.default_rng(20261105)
    sym = np.frombuffer(b"ACGTACGTACGTNacgtnMRWSYKVHDBU.-Xx?Zq", dtype=np.uint8)
    text = sym[rng.integers(0, len(sym), size=600)].tobytes()
    out = {"text": np.frombuffer(text, dtype=np.uint8)}
    for alpha, k in PAIRS:
        w = oracle.ref.kmer_windows_alpha(alpha, k, text)
        out["windows_a%d_k%d" % (alpha, k)] = w
        out["revcomp_a%d_k%d" % (alpha, k)] = oracle.ref.revcomp_alpha(alpha, k, w)
        out["canonical_a%d_k%d" % (alpha, k)] = oracle.ref.canonical_alpha(alpha, k, w)
    np.savez_compressed(os.path.join(HERE, "alphabet_vectors.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
