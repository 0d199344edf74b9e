#!/usr/bin/env python
"""Measurement of BASELINE.json configs[1..3] (the parity-case configs that are not bench.py's line) on the GPU box.
One JSON line per case; results are cross-checked through size-independent properties (distinct count against
numpy / the HLL-free exact count, every inserted key found, erased keys gone).

  lp        config 2: hashmap_linearprobe_doubling, --max_load sweep 0.5 .. 0.9 (min 0.35), both from capacity 128
            through doubling rehashes (insert_no_estimate) and presized (insert with estimate)
  k63       config 3: Kmer<63> (two 64-bit words) canonical keys, 100 M inserts / 10 M queries,
            StoreHash in {MURMUR64, MURMUR64avx, FARM, CLHASH}
  counter   config 4 shape: synthetic 150-bp reads (100 Mbp random genome, both strands, 1 % substitutions) ->
            canonical 31-mers (kh_kmers_from_reads) -> count map.  1 GPU: this process; N GPUs: under torchrun
            (reads split evenly by read index, DistributedMap insert = partition + exchange + local insert)

usage: python tools/bench_configs.py lp|k63|counter [--n N] [--reads R]
       python -m torch.distributed.run --nproc-per-node N ... tools/bench_configs.py counter
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import numpy as np
import torch

import kmerhash_b200 as kh
from kmerhash_b200 import ops


def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e


def ref_keys(n, k, canonical):
    """the reference's generate_input bytes (drivers/libkh_geninput.so, as bench.py)"""
    from bench import generate_workload
    return generate_workload(n, 1, seed=23, k=k, canonical=canonical)[0]


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def cpu_arm(keys, map_kind, hash_kind, min_load, max_load, n_sample=10_000_000):
    """the compiled reference (oracle/_ref) on the first n_sample keys, one thread, same map / hash / load factors"""
    import oracle
    if oracle.ref is None:
        return None
    ns = min(n_sample, len(keys))
    nq = max(1, ns // 10)
    try:
        secs, cnts = oracle.ref.baseline_run(np.ascontiguousarray(keys[:ns]), np.ascontiguousarray(keys[:nq]), map_kind, hash_kind,
                                             threads=1, min_load=min_load, max_load=max_load)
    except Exception as e:          # combination not instantiated in the harness
        return {"unavailable": str(e)}
    return {"kind": "reference", "cores": 1, "sample_keys": ns, "inserts_per_s": ns / (secs["estimate"] + secs["insert"]),
            "finds_per_s": nq / secs["find"], "counts_per_s": nq / secs["count"], "erases_per_s": nq / secs["erase"]}


def roofline(n, nq, ph, W):
    """phase-level, SURVEY.md 8d bytes: insert 72 / 80 B per key, find 44 / 52 B"""
    pk = peak_gbs()
    bi, bf = (72, 44) if W == 1 else (80, 52)
    gi, gf = bi * n / ph[0] / 1e6, bf * nq / ph[1] / 1e6
    return {"bound": "hbm", "peak": pk, "unit": "GB/s", "insert": {"achieved": gi, "frac": gi / pk, "bytes_per_key": bi},
            "find": {"achieved": gf, "frac": gf / pk, "bytes_per_key": bf}}


def gen_keys(n, words, repeats=10, seed=23):
    """generate_input's shape (benchmark/BenchmarkHashTables.cpp:181-227): 62 significant bits per word"""
    rng = np.random.default_rng(seed)
    n_uniq = int(n / ((repeats + 1) / 2.0)) + 1
    shape = (n_uniq,) if words == 1 else (n_uniq, words)
    uniq = rng.integers(0, 2**63, size=shape, dtype=np.uint64) & np.uint64(0x7FFFFFFF7FFFFFFF)
    if words == 1:
        uniq &= np.uint64((1 << 62) - 1)
    else:
        uniq[:, words - 1] &= np.uint64((1 << 62) - 1)     # k = 63: top word holds 62 bits
    reps = rng.integers(1, repeats + 1, size=n_uniq)
    keys = np.repeat(uniq, reps, axis=0)[:n]
    if len(keys) < n:
        keys = np.concatenate([keys, keys[: n - len(keys)]])
    keys = keys[rng.permutation(n)]
    return np.ascontiguousarray(keys)


def timed_phases(m, d_keys, d_q, estimate, steps=3, start_cap=128):
    out = []
    for _ in range(steps + 1):
        m.clear(); m.rehash(start_cap)
        e0 = ev(); new = m.insert(d_keys, estimate=estimate) if estimate else m.insert_no_estimate(d_keys)
        e1 = ev(); f = m.find(d_q, 0)
        e2 = ev(); c = m.count(d_q)
        e3 = ev(); er = m.erase(d_q)
        e4 = ev(); c2 = m.count(d_q)
        e5 = ev(); torch.cuda.synchronize()
        out.append([a.elapsed_time(b) for a, b in ((e0, e1), (e1, e2), (e2, e3), (e3, e4), (e4, e5))])
        chk = (new, int((f != 0).sum()), int(c.sum()), er, int(c2.sum()), m.capacity())
    return np.array(out[1:]).mean(axis=0), chk


def run_lp(args):
    n, nq = args.n, args.n // 10
    keys = ref_keys(n, 31, False)
    distinct = int(np.unique(keys).shape[0])
    d = torch.from_numpy(keys.view(np.int64)).cuda()
    q = d[:nq].contiguous()
    q_distinct = int(np.unique(keys[:nq]).shape[0])
    for kind, ctor in (("hashmap_linearprobe_doubling", kh.hashmap_linearprobe_doubling),
                       ("hashmap_robinhood_doubling", kh.hashmap_robinhood_doubling)):
        for max_load in (0.5, 0.6, 0.7, 0.8, 0.9):
            for mode in ("doubling", "presized"):
                m = ctor(1, "MURMUR64", min_load=0.35, max_load=max_load)
                ph, chk = timed_phases(m, d, q, estimate=(mode == "presized"))
                assert chk[0] == distinct and chk[1] == nq and chk[2] == nq and chk[3] == q_distinct and chk[4] == 0, chk
                cpu = cpu_arm(keys, "LINEARPROBE" if "linear" in kind else "ROBINHOOD", "MURMUR64", 0.35, max_load) if mode == "doubling" else None
                print(json.dumps({"config": 2, "map": kind, "max_load": max_load, "mode": mode, "n": n, "nq": nq,
                                  "roofline": roofline(n, nq, ph, 1), "cpu_baseline": cpu,
                                  "distinct": distinct, "capacity_after_insert": None, "final_capacity": chk[5],
                                  "inserts_per_s": n / ph[0] * 1e3, "finds_per_s": nq / ph[1] * 1e3,
                                  "counts_per_s": nq / ph[2] * 1e3, "erases_per_s": nq / ph[3] * 1e3,
                                  "phase_ms": [round(float(x), 3) for x in ph]}), flush=True)
                del m


def run_k63(args):
    n, nq = args.n, args.n // 10
    raw = ref_keys(n, 63, False)
    keys = ref_keys(n, 63, True)                    # generate_input(canonical = true), the reference's -c
    d = torch.from_numpy(raw.view(np.int64)).cuda()
    t0 = ev(); d = ops.canonicalize(d, 63); t1 = ev(); torch.cuda.synchronize()
    canon_ms = t0.elapsed_time(t1)
    assert np.array_equal(d.cpu().numpy().view(np.uint64), keys), "device canonicalisation differs from the reference generator's"
    # exact distinct count of the canonical keys on the device (sort of (hi, lo) pairs)
    u = torch.unique(d, dim=0).shape[0] if n <= 20_000_000 else None
    q = d[:nq].contiguous()
    for h in ("MURMUR64", "MURMUR64avx", "FARM", "CLHASH"):
        m = kh.hashmap_robinhood_offsets_count(2, h, min_load=0.35, max_load=0.8)
        ph, chk = timed_phases(m, d, q, estimate=True)
        assert chk[1] == nq and chk[2] == nq and chk[4] == 0 and (u is None or chk[0] == u), (chk, u)
        u = chk[0]                      # every hash variant must find the same number of distinct keys
        print(json.dumps({"config": 3, "k": 63, "hash": h, "n": n, "nq": nq, "distinct": chk[0],
                          "roofline": roofline(n, nq, ph, 2), "cpu_baseline": cpu_arm(keys, "OFFSETS_COUNT", h, 0.35, 0.8),
                          "canonicalize_keys_per_s": n / canon_ms * 1e3,
                          "inserts_per_s": n / ph[0] * 1e3, "finds_per_s": nq / ph[1] * 1e3,
                          "counts_per_s": nq / ph[2] * 1e3, "erases_per_s": nq / ph[3] * 1e3,
                          "phase_ms": [round(float(x), 3) for x in ph], "final_capacity": chk[5]}), flush=True)
        del m


def synth_reads(n_reads, read_len, genome_len, rank, seed=1, err=0.01, chunk=1 << 20):
    """ASCII reads on the device: uniform random genome (same on every rank), reads sampled uniformly from
    either strand, 1 % substitutions"""
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    genome = torch.randint(0, 4, (genome_len,), generator=g, device="cuda", dtype=torch.uint8)
    g2 = torch.Generator(device="cuda"); g2.manual_seed(1000 + rank)
    lut = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device="cuda")
    out = torch.empty(n_reads * read_len, dtype=torch.uint8, device="cuda")
    ar = torch.arange(read_len, device="cuda")
    for s in range(0, n_reads, chunk):
        m = min(chunk, n_reads - s)
        st = torch.randint(0, genome_len - read_len, (m,), generator=g2, device="cuda")
        codes = genome[st[:, None] + ar[None, :]]
        rc = torch.rand(m, generator=g2, device="cuda") < 0.5
        codes = torch.where(rc[:, None], 3 - codes.flip(1), codes)
        sub = torch.rand((m, read_len), generator=g2, device="cuda") < err
        shift = torch.randint(1, 4, (m, read_len), generator=g2, device="cuda", dtype=torch.uint8)
        codes = torch.where(sub, (codes + shift) & 3, codes)
        out[s * read_len:(s + m) * read_len] = lut[codes.long()].reshape(-1)
    offs = (torch.arange(n_reads + 1, device="cuda", dtype=torch.int64) * read_len)
    return out, offs


def run_counter(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    k, L = 31, 150
    n_reads = args.reads // world
    bases, offs = synth_reads(n_reads, L, args.genome, rank)
    n_kmers = n_reads * (L - k + 1)
    if world > 1:
        import torch.distributed as dist
        from kmerhash_b200.distributed import DistributedMap
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dm = DistributedMap(1, "MURMUR64avx", "MURMUR64avx", "COUNT", canonical_k=0, device=local, chunks=1)
        dm.e.table.set_load_factors(0.35, 0.8)
        table = dm.e.table
    else:
        table = kh.hashmap_robinhood_offsets_count(1, "MURMUR64avx", device=local, min_load=0.35, max_load=0.8)
    tg, ti = [], []
    for it in range(args.steps + 1):
        table.clear(); table.rehash(128)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        fused = world > 1 and os.environ.get("KH_COUNTER_FUSED", "1") not in ("", "0")
        if fused:                             # k-mers generated inside the sender kernel: no separate generation phase
            e0 = ev(); e1 = ev()
            dm.insert_reads(bases, offs, k, canonical=True); dist.barrier()
            e2 = ev(); torch.cuda.synchronize()
            kmers = None
        else:
            e0 = ev(); kmers = ops.kmers_from_reads_dev(bases, offs, k, canonical=True)
            e1 = ev()
            if world > 1:
                dm.insert(kmers); dist.barrier()
            else:
                table.insert(kmers, estimate=True)
            e2 = ev(); torch.cuda.synchronize()
            assert kmers.shape[0] == n_kmers
        if it:
            tg.append(e0.elapsed_time(e1)); ti.append(e1.elapsed_time(e2))
        if it < args.steps and kmers is not None:
            del kmers
    t = torch.tensor([float(np.mean(tg)), float(np.mean(ti))], device="cuda")
    size = torch.tensor([table.size()], device="cuda", dtype=torch.int64)
    # property: the sum of all counts equals the number of k-mers generated
    _, vals = table.to_vector()
    vsum = torch.tensor([int(np.asarray(vals, dtype=np.uint64).sum())], device="cuda", dtype=torch.int64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(size); dist.all_reduce(vsum)
    if rank == 0:
        tgen, tins = t.tolist()
        total = n_kmers * world
        assert int(vsum.item()) == total, (int(vsum.item()), total)
        print(json.dumps({"config": 4, "workload": "BenchmarkKmerCounter shape: %d reads x %d bp (genome %d bp, both strands, "
                          "1%% substitutions) -> canonical 31-mers -> count map" % (n_reads * world, L, args.genome),
                          "n_gpus": world, "kmers": total, "distinct": int(size.item()), "kmer_gen_fused_with_sender": bool(fused),
                          "kmer_gen_per_s": (total / tgen * 1e3 if tgen > 1e-3 else None), "inserts_per_s": total / tins * 1e3,
                          "kmers_counted_per_s_total": total / (tgen + tins) * 1e3,
                          "phase_ms": {"kmer_gen": tgen, "insert": tins},
                          "kmer_gen_GBps_algorithmic": (total * 9.25 / tgen / 1e6 if tgen > 1e-3 else None),
                          "sum_of_counts_equals_kmers": True}), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["lp", "k63", "counter"])
    ap.add_argument("--n", type=int, default=100_000_000)
    ap.add_argument("--reads", type=int, default=8_333_334)
    ap.add_argument("--genome", type=int, default=100_000_000)
    ap.add_argument("--steps", type=int, default=2)
    args = ap.parse_args()
    {"lp": run_lp, "k63": run_k63, "counter": run_counter}[args.what](args)


if __name__ == "__main__":
    main()
