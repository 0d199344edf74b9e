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


COUNTER = os.path.join(ROOT, "drivers", "BenchmarkKmerCounter")


def _write_fastq(path, reads):
    with open(path, "w") as f:
        for i, r in enumerate(reads):
            f.write("@r%d\n%s\n+\n%s\n" % (i, r, "I" * len(r)))


@pytest.mark.gpu
@pytest.mark.parametrize("gpus", [1, 2])
def test_kmercounter_driver_matches_oracle(tmp_path, oracle_mod, gpus):
    """BenchmarkKmerCounter on a FASTQ file (reads with N runs and short reads): the [Kmer][count] records it writes
    equal the oracle's canonical 31-mer count map (CountType = uint16_t, benchmark/BenchmarkKmerCounter.cpp:184)."""
    import kmerhash_b200 as kh
    if gpus > kh.device_count():
        pytest.skip("needs %d GPUs" % gpus)
    rng = np.random.default_rng(11)
    genome = "".join("ACGT"[c] for c in rng.integers(0, 4, size=20000))
    reads = []
    for i in range(3000):
        ln = int(rng.integers(20, 151))
        s = int(rng.integers(0, len(genome) - ln))
        r = genome[s:s + ln]
        if i % 7 == 0 and ln > 40:
            p = int(rng.integers(0, ln - 2))
            r = r[:p] + "NN" + r[p + 2:]
        reads.append(r)
    fq = str(tmp_path / "reads.fastq")
    out = str(tmp_path / "counts.bin")
    _write_fastq(fq, reads)
    r = subprocess.run([COUNTER, "-G", str(gpus), "-O", out, "--sample", "5000", fq], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = np.fromfile(out, dtype=np.uint8).reshape(-1, 10)
    keys = np.ascontiguousarray(raw[:, :8]).view(np.uint64).reshape(-1)
    cnts = np.ascontiguousarray(raw[:, 8:]).view(np.uint16).reshape(-1)

    bases = np.frombuffer("".join(reads).encode(), dtype=np.uint8)
    offsets = np.concatenate([[0], np.cumsum([len(x) for x in reads])]).astype(np.uint64)
    exp = oracle_mod.port.kmers_from_reads(bases, offsets, 31, True)
    m = oracle_mod.port.Map(1, "COUNT")
    m.insert(exp, default=1)
    ek, ev = m.to_vector()
    a, b = sort_pairs(keys, cnts.astype(np.uint32)), sort_pairs(ek, ev & 0xFFFF)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert "kmers %d distinct %d" % (len(exp), len(ek)) in r.stdout
    for phase in ("read", "parse", "insert", "count", "find", "write"):
        assert phase in r.stdout


def test_kmercounter_driver_is_built():
    assert os.path.exists(COUNTER)
    r = subprocess.run([COUNTER, "--help"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "BenchmarkKmerCounter" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("world,pipe,mode", [(2, "1", "sm"), (2, "4", "sm"), (2, "4", "ce")])
def test_distributed_map_on_real_gpus_under_torchrun(world, pipe, mode):
    """one process per GPU (torchrun, NCCL + CUDA-IPC mapped peer buffers): tools/dist_parity.py checks every rank's table
    contents against the oracle's per-rank rule through the slab exchange, its overflow fallback, slab growth, empty
    batches, queries in input order and erase.  Needs >= 2 visible GPUs (the emulated-p tests cover the kernels on one)."""
    import subprocess
    import socket
    import sys
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    # 5 M keys per rank: large enough for the bulk insert and for the pipelined exchange (run with it off, with the sender
    # kernels storing over NVLink, and with the copy engines moving the partitioned runs)
    env = dict(os.environ, KH_PARITY_SHORT="1", KH_PARITY_N0="5000000,150000", KH_PIPE_CHUNKS=pipe, KH_PIPE_MODE=mode)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                        "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tools", "dist_parity.py")],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count(" OK") >= world


@pytest.mark.gpu
@pytest.mark.parametrize("gpus", [1, 2])
@pytest.mark.parametrize("index", ["count", "first", "last"])
def test_dist_hash_tables_driver(kh, gpus, index):
    """drivers/BenchmarkDistHashTables (the reference's driver shape over khd_*): --check compares every count / find answer
    with a host gold map that applies the reducer in input order -- for FIRST / LAST on several GPUs that is the exchange
    order (sender slice, then position) the reference's owner ranks see; after erasing every key the map must be empty"""
    if kh.device_count() < gpus:
        pytest.skip("needs %d GPUs" % gpus)
    exe = os.path.join(ROOT, "drivers", "BenchmarkDistHashTables")
    r = subprocess.run([exe, "-C", "400000", "-R", "4", "-G", str(gpus), "-I", index, "--check", "--missing-frac", "0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 mismatches" in r.stdout
    for phase in ("init", "copy_insert", "insert", "count", "find", "erase"):
        assert phase in r.stdout
    r = subprocess.run([exe, "-C", "200000", "-R", "8", "-G", str(gpus), "-I", index, "--check", "--missing-frac", "0.25"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "0 mismatches" in r.stdout, r.stdout + r.stderr
