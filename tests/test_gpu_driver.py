"""The reference-shaped C++ driver (drivers/BenchmarkHashTables over include/kmerhash_b200.hpp) against the oracle:
the driver generates the reference's input itself (srand(23) recipe, glibc rand + libstdc++ random_shuffle), the
oracle restates that generator, so table contents and phase results must agree exactly."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, sort_pairs

DRIVER = os.path.join(ROOT, "drivers", "BenchmarkHashTables")


def run_driver(tmp_path, *args):
    dump = str(tmp_path / "table.bin")
    r = subprocess.run([DRIVER, *args, "--dump", dump], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = np.fromfile(dump, dtype=np.uint8)
    rec, n = np.frombuffer(raw[:16].tobytes(), dtype=np.uint64)
    body = raw[16:].reshape(int(n), int(rec))
    kw = (int(rec) - 8) // 8
    keys = np.ascontiguousarray(body[:, : 8 * kw]).view(np.uint64).reshape(int(n), kw)
    vals = np.ascontiguousarray(body[:, 8 * kw: 8 * kw + 4]).view(np.uint32).reshape(int(n))
    nums = {k: int(v) for k, v in re.findall(r"(size after erase|present after erase|present|erased) (\d+)", r.stdout)}
    return (keys[:, 0].copy() if kw == 1 else keys), vals, nums, r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("k,mode,mapname", [
    (31, "COUNT", "robinhood_offset_overflow"),
    (31, "FIRST", "robinhood_offset_overflow"),     # the literal default: DiscardReducer, value = input index
    (63, "COUNT", "robinhood_offset_overflow"),
    (31, "FIRST", "linearprobe"),
    (31, "FIRST", "classic_robinhood"),
])
def test_driver_matches_oracle(tmp_path, oracle_mod, k, mode, mapname):
    n, qf = 300_000, 10
    args = ["-N", str(n), "-Q", str(qf), "-k", str(k), "-m", mapname]
    if mode == "COUNT":
        args.append("--count")
    keys, vals, nums, out = run_driver(tmp_path, *args)
    ik, iv = oracle_mod.port.generate_input(k, n)
    m = oracle_mod.port.Map(oracle_mod.port.words(k), mode)
    if mode == "COUNT":
        m.insert(ik, default=1)
    else:
        m.insert(ik, vals=iv)
    ek, ev = m.to_vector()
    a, b = sort_pairs(keys, vals), sort_pairs(ek, ev)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    q = ik[: n // qf]
    erased = m.erase(q)
    assert nums["erased"] == erased and nums["size after erase"] == m.size()
    if "present after erase" in nums:
        assert nums["present after erase"] == 0 and nums["present"] == len(q)
    for phase in ("insert", "count", "erase", "count2"):
        assert phase in out


def test_driver_is_built():
    """build() compiles the driver on the CPU box (no GPU needed to link against the C ABI)"""
    assert os.path.exists(DRIVER)
    r = subprocess.run([DRIVER, "--help"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "BenchmarkHashTables" in r.stdout
