import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden", "reference_vectors.npz")
READ78 = "ACGTTGCAAGGCTTAACCGGTACGATCGATTACAGGCATCGATCGGATCCATGCAAGTCGATCGTAAGCTTGACTGAC"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    return dict(np.load(GOLDEN))


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle
    return oracle


@pytest.fixture(scope="session")
def kh():
    import kmerhash_b200
    return kmerhash_b200


def sort_pairs(keys, vals):
    """sorted (key, value) view used by every table-content comparison"""
    keys = np.asarray(keys, dtype=np.uint64)
    vals = np.asarray(vals)
    if keys.ndim == 2:
        order = np.lexsort(keys.T[::-1])
    else:
        order = np.argsort(keys, kind="stable")
    return keys[order], vals[order]


def make_reads(rng, n_reads, min_len, max_len, n_rate=0.0, eol=False, lower=False):
    """random reads concatenated + offsets; optional N's, embedded newlines and lower case"""
    alphabet = np.frombuffer(b"ACGT", dtype=np.uint8)
    pieces, offsets = [], [0]
    for _ in range(n_reads):
        L = int(rng.integers(min_len, max_len + 1))
        s = alphabet[rng.integers(0, 4, size=L)].copy()
        if lower:
            mask = rng.random(L) < 0.3
            s[mask] |= 0x20
        if n_rate > 0:
            mask = rng.random(L) < n_rate
            s[mask] = ord("N") if rng.random() < 0.5 else ord("n")
        if eol and L > 4:
            pos = rng.integers(1, L, size=max(1, L // 40))
            s = np.insert(s, pos, ord("\n"))
            if rng.random() < 0.5:
                s = np.insert(s, int(rng.integers(1, len(s))), ord("\r"))
        pieces.append(s)
        offsets.append(offsets[-1] + len(s))
    bases = np.concatenate(pieces) if pieces else np.zeros(0, dtype=np.uint8)
    return bases, np.array(offsets, dtype=np.uint64)
