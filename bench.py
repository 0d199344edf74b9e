#!/usr/bin/env python
"""bench.py -- BenchmarkHashTables-shaped measurement of the hot path (BASELINE.json metric:
k-mer inserts/s & finds/s, k=31, 100M keys, ~5x repeats, then 10M find/count/erase/count).

One step = the reference driver's phase sequence (benchmark/BenchmarkHashTables.cpp:1191-1364) on a
default-constructed count table:  insert (HLL estimate -> reserve -> batch insert, the reference's
insert()) -> find -> count -> erase -> count2.  The input is the reference's own: generate_input
(benchmark/BenchmarkHashTables.cpp:181-227; drivers/generate_input.hpp calls glibc rand() and libstdc++'s
random_shuffle exactly as the reference does), queries = the first n/10 input keys.
  value      inserts/s of the insert phase with the keys already resident in HBM (whole job, all GPUs)
  e2e        the same through the host API: pinned HOST buffers in, HOST results out, copies inside the timing
  roofline   the insert PHASE (every kernel an insert needs) on 72 B/key against the measured HBM
             peak; per kernel its OWN bytes with the ncu DRAM traffic beside them; and the second ceiling of
             SURVEY.md 8d: operations/s against the random 32-B-sector rates measured live at the table's footprint
  cpu_baseline  the reference's own CPU table (oracle/_ref, compiled from the reference headers) on a bounded
                sample of the same workload, 1 thread (BenchmarkHashTables is single-process)
  --impl reference   the reference's CPU implementation alone: value = the stock single-process path on the FULL
                config; the all-cores key-space-sharded run (BenchmarkDistHashTables without MPI) as an extra field
  --config 1 (default) | 5w | 5s    BASELINE.json configs[0] | configs[4] weak (1 B keys per GPU) | strong (8 B keys)
N > 1 (torchrun): one process per GPU, each rank brings its own keys (weak scaling); owner rank =
DistHash_9876543(key) & (p-1); exchange by peer stores over NVLink; local insert; queries answered in place.
Before anything is timed the N > 1 arm checks every rank's table contents, find/count answers and erase
against numpy (exact counts) and the owner rule restated here (murmur3 x86_128, seed 9876543).
"""
import argparse
import ctypes as C
import gc
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_KEYS_DEFAULT = 100_000_000
N_QUERY_DEFAULT = 10_000_000
REPEATS = 10
BYTES_PER_INSERT = 72     # 8 B key read + one 32-B sector read + one 32-B sector written back
BYTES_PER_FIND = 44       # 8 + 32 + 4
DIST_SEED = 9876543


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="1", choices=["1", "5w", "5s"])
    ap.add_argument("--n-keys", type=int, default=None)
    ap.add_argument("--n-query", type=int, default=None)
    ap.add_argument("--hash", default="MURMUR64avx")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--ref-full-steps-max", type=int, default=40,
                    help="reference arm: above this many steps the full-size 1-thread run would take too long; sample instead")
    a = ap.parse_args()
    if a.n_keys is None:
        a.n_keys = N_KEYS_DEFAULT if a.config == "1" else (1_000_000_000 if a.config == "5w" else 8_000_000_000 // max(a.gpus, 1))
    if a.n_query is None:
        a.n_query = a.n_keys // 10 if a.config == "1" else min(a.n_keys // 10, 100_000_000)
    return a


# ---------------------------------------------------------------------------------------------------
# input
# ---------------------------------------------------------------------------------------------------
_GEN = None


def _generator():
    global _GEN
    if _GEN is None:
        path = os.path.join(ROOT, "drivers", "libkh_geninput.so")
        if not os.path.exists(path):
            subprocess.run(["make", "-C", os.path.join(ROOT, "drivers"), "libkh_geninput.so"], check=True, capture_output=True)
        _GEN = C.CDLL(path)
        _GEN.khg_generate_input.restype = C.c_size_t
        _GEN.khg_generate_input.argtypes = [C.c_int, C.c_size_t, C.c_size_t, C.c_int, C.c_uint, C.c_void_p, C.c_void_p]
    return _GEN


def generate_workload(n, nq, seed=23, k=31, canonical=False):
    """the reference's generate_input (benchmark/BenchmarkHashTables.cpp:181-227) through drivers/libkh_geninput.so:
    srand(seed) (the reference: 23), 62-bit words from two rand() calls, rand() % 10 extra copies of each key,
    std::random_shuffle; queries = the first nq input keys (-Q, :1400)"""
    W = 1 if k <= 32 else 2
    keys = np.empty(n * W, dtype=np.uint64)
    got = _generator().khg_generate_input(k, n, REPEATS, int(canonical), seed, keys.ctypes.data, None)
    assert got == n, (got, n)
    if W > 1:
        keys = keys.reshape(n, W)
    return keys, np.ascontiguousarray(keys[:nq])


def device_keys_config5(n, rank, repeats=8):
    """BenchmarkDistHashTables shape (SURVEY.md 8d config 5 ii): uint64 keys, -R 8 (every distinct key 1..8 times),
    fixed seed 1000 + rank; generated on the device (host rand() would take minutes per billion keys)"""
    import torch
    g = torch.Generator(device="cuda"); g.manual_seed(1000 + rank)
    n_uniq = int(n / ((repeats + 1) / 2.0)) + 1
    uniq = torch.randint(0, 1 << 62, (n_uniq,), generator=g, device="cuda", dtype=torch.int64)
    reps = torch.randint(1, repeats + 1, (n_uniq,), generator=g, device="cuda", dtype=torch.int64)
    keys = torch.repeat_interleave(uniq, reps)[:n]
    del uniq, reps
    if keys.shape[0] < n:
        keys = torch.cat([keys, keys[: n - keys.shape[0]]])
    out = torch.empty_like(keys)
    step = 1 << 28                                      # shuffle in pieces: a full randperm(n) of int64 doubles the footprint
    perm = torch.randperm(n, generator=g, device="cuda") if n <= step else None
    if perm is not None:
        out = keys[perm]
    else:
        # two-level shuffle: random rotation of interleaved strides, then a permutation inside every piece
        stride = (n + step - 1) // step
        idx = torch.arange(n, device="cuda", dtype=torch.int64)
        idx = (idx % stride) * ((n + stride - 1) // stride) + idx // stride
        idx = torch.clamp(idx, max=n - 1)
        out = keys[idx]
        del idx
        for s in range(0, n, step):
            m = min(step, n - s)
            out[s:s + m] = out[s:s + m][torch.randperm(m, generator=g, device="cuda")]
    del keys
    return out


# ---------------------------------------------------------------------------------------------------
# the owner rule restated in numpy (parity pre-check of the N > 1 arm; independent of the library and of oracle/):
# fsc::hash::murmur3avx64 = MurmurHash3_x86_128 of the 8 key bytes, low 64 bits (KH/murmurhash3_64_avx.hpp:1553-1598,
# ext smhasher MurmurHash3.cpp), rank = hash & (p - 1) (KH/distributed_batched_robinhood_map.hpp:164-178,652)
# ---------------------------------------------------------------------------------------------------
def np_murmur3_x86_128_lo64(keys, seed):
    M = np.uint64(0xFFFFFFFF)
    k = keys.astype(np.uint64)

    def mul(a, b):
        return (a * np.uint64(b)) & M

    def rotl(x, r):
        return ((x << np.uint64(r)) | (x >> np.uint64(32 - r))) & M

    def fmix(h):
        h = h ^ (h >> np.uint64(16)); h = mul(h, 0x85ebca6b); h = h ^ (h >> np.uint64(13)); h = mul(h, 0xc2b2ae35)
        return h ^ (h >> np.uint64(16))

    c1, c2, c3, c4 = 0x239b961b, 0xab0e9789, 0x38b34ae5, 0xa1e38b93
    h1 = np.full(k.shape, seed, dtype=np.uint64); h2 = h1.copy(); h3 = h1.copy(); h4 = h1.copy()
    # 8 bytes: the tail switch of MurmurHash3_x86_128, cases 8..5 -> k2, 4..1 -> k1
    k1 = k & M
    k2 = k >> np.uint64(32)
    k2 = mul(k2, c2); k2 = rotl(k2, 16); k2 = mul(k2, c3); h2 = h2 ^ k2
    k1 = mul(k1, c1); k1 = rotl(k1, 15); k1 = mul(k1, c2); h1 = h1 ^ k1
    ln = np.uint64(8)
    h1 ^= ln; h2 ^= ln; h3 ^= ln; h4 ^= ln
    h1 = (h1 + h2) & M; h1 = (h1 + h3) & M; h1 = (h1 + h4) & M
    h2 = (h2 + h1) & M; h3 = (h3 + h1) & M; h4 = (h4 + h1) & M
    h1 = fmix(h1); h2 = fmix(h2); h3 = fmix(h3); h4 = fmix(h4)
    h1 = (h1 + h2) & M; h1 = (h1 + h3) & M; h1 = (h1 + h4) & M
    h2 = (h2 + h1) & M
    return h1 | (h2 << np.uint64(32))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.lines, self.proc, self.gpu = [], None, gpu_index
        self.nvml = None

    def _nvml_loop(self):
        nv, h = self.nvml
        masks = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self._stop:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                r = int(get_reasons(h))
                self.samples.append((float(sm), [k for k, m in masks.items() if r & m]))
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        # in-process NVML (same counters nvidia-smi prints, without a second process querying the GPU every few ms,
        # which stalls sub-millisecond phases); nvidia-smi -lms only when the NVML binding is missing
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.gpu)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.nvml, self.samples, self._stop = (nv, h), [], False
            self.th = threading.Thread(target=self._nvml_loop, daemon=True)
            self.th.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
            # nvidia-smi's start-up (NVML init) can stall the driver for tens of ms: let it finish before anything is timed
            t0 = time.time()
            while not self.lines and time.time() - t0 < 5.0:
                time.sleep(0.01)
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.nvml:
            self._stop = True
            self.th.join(1.0)
            if not self.samples:
                return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["no samples"]}
            sm = sorted(x[0] for x in self.samples)
            busy = sm[len(sm) // 2:]
            reasons = sorted({r for x in self.samples for r in x[1]})
            return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = sorted(sm)[len(sm) // 2:]          # upper half = samples under load
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ---------------------------------------------------------------------------------------------------
# CPU side: the reference's own implementation (oracle/_ref, else the C port).  Only these functions
# touch oracle/ -- as the timed baseline, never on the product path.
# ---------------------------------------------------------------------------------------------------
def run_reference_phases(keys, queries, threads, hash_kind):
    """replays estimate -> insert -> find -> count -> erase -> count2 on the reference's
    hashmap_robinhood_offsets_count; returns (seconds per phase, counts, kind)"""
    import oracle
    if oracle.ref is not None:
        secs, cnts = oracle.ref.baseline_run(keys, queries, "OFFSETS_COUNT", hash_kind, threads=threads,
                                             min_load=0.35, max_load=0.8)
        return secs, cnts, "reference"
    # the reference did not compile here: time the plain-C port (single thread)
    m = oracle.port.Map(1, "COUNT")
    secs = {"partition": 0.0, "estimate": 0.0}
    t = time.perf_counter(); m.insert(keys); secs["insert"] = time.perf_counter() - t
    t = time.perf_counter(); f = m.find(queries, 0); secs["find"] = time.perf_counter() - t
    t = time.perf_counter(); c = m.count(queries); secs["count"] = time.perf_counter() - t
    t = time.perf_counter(); e = m.erase(queries); secs["erase"] = time.perf_counter() - t
    t = time.perf_counter(); c2 = m.count(queries); secs["count2"] = time.perf_counter() - t
    return secs, {"distinct": m.size() + e, "found": int((f != 0).sum()), "counted": int(c.sum()), "erased": e,
                  "counted2": int(c2.sum())}, "port"


def cpu_baseline(keys, queries, hash_kind, n_sample=20_000_000):
    """bounded sample of the same workload on ONE host core (BenchmarkHashTables is single-process)"""
    ns = min(n_sample, len(keys))
    nq = max(1, ns // 10)
    secs, cnts, kind = run_reference_phases(np.ascontiguousarray(keys[:ns]), np.ascontiguousarray(queries[:nq]), 1, hash_kind)
    t_ins = secs["estimate"] + secs["insert"]
    return {"value": ns / t_ins, "unit": "inserts/s", "cores": 1, "kind": kind,
            "sample": "first %d of the %d keys, %d queries; hashmap_robinhood_offsets_count<Kmer31,u32,%s>, "
                      "HLL estimate + reserve + insert timed together" % (ns, len(keys), nq, hash_kind),
            "finds_per_s": nq / secs["find"], "counts_per_s": nq / secs["count"], "erases_per_s": nq / secs["erase"],
            "seconds": {k: round(v, 4) for k, v in secs.items()}, "distinct": cnts["distinct"]}


def reference_arm(args):
    """value = the stock single-process path (BenchmarkHashTables: one thread, hashmap_robinhood_offsets_count) on OUR arm's
    config; every step replays the full phase sequence.  The all-cores run (key space sharded by the reference's rank rule,
    one table per thread, MPI all-to-all excluded: no MPI on this box) is reported beside it from one bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    full = (args.warmup + args.steps) <= args.ref_full_steps_max and args.config == "1"
    ns = args.n_keys if full else min(args.n_keys, 25_000_000)
    nq = args.n_query if full else max(1, ns // 10)
    keys, queries = generate_workload(ns, nq, seed=23)
    ins, fnd, tot = [], [], []
    kind, cnts = "reference", None
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        secs, cnts, kind = run_reference_phases(keys, queries, 1, args.hash)
        t1 = time.perf_counter()
        if it >= args.warmup:
            ins.append(secs["estimate"] + secs["insert"])
            fnd.append(secs["find"])
            tot.append(t1 - t0)
    t_ins, t_find = float(np.mean(ins)), float(np.mean(fnd))
    value = ns / t_ins
    # all host cores, one bounded sample (not the headline: the partition is harness code, reported on its own)
    ms = min(ns, 25_000_000)
    s_all, _, _ = run_reference_phases(np.ascontiguousarray(keys[:ms]), np.ascontiguousarray(queries[:max(1, ms // 10)]), cores, args.hash)
    line = {
        "impl": "reference", "metric": "k-mer inserts/s (k=31, count table)", "value": value, "unit": "inserts/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(tot)),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "BenchmarkHashTables default: %d random 31-mer uint64 keys (~5.5x repeats, %d distinct) "
                               "insert, then %d find/count/erase/count; robinhood_offsets count map; 1 host thread"
                               % (ns, cnts["distinct"], nq),
                   "n_keys": ns, "n_query": nq, "hash": args.hash, "load_factors": [0.35, 0.8],
                   "input": "generate_input of the reference (srand(23), rand(), random_shuffle), the same bytes as our arm",
                   "full_config": bool(full),
                   "insert_phase": "HLL estimate + reserve + batch insert (the reference's insert())"},
        "finds_per_s": nq / t_find,
        "cpu_baseline": {"value": value, "unit": "inserts/s", "cores": 1, "kind": kind,
                         "sample": "%d keys / %d queries per step, single process like BenchmarkHashTables" % (ns, nq)},
        "all_cores_sharded": {"cores": cores, "n_keys": ms,
                              "inserts_per_s_excl_partition": ms / (s_all["estimate"] + s_all["insert"]),
                              "inserts_per_s_incl_partition": ms / (s_all["partition"] + s_all["estimate"] + s_all["insert"]),
                              "finds_per_s": max(1, ms // 10) / s_all["find"],
                              "seconds": {k: round(v, 4) for k, v in s_all.items()},
                              "note": "one reference table per thread, keys sharded by the reference's rank rule; the partition "
                                      "is harness code (counting pass + scatter), MPI all-to-all excluded: no MPI on this box"},
        "e2e": {"value": value, "unit": "inserts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------
def measured_peaks():
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(pk["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed ncu --set full capture
    (profiles/ncu_traffic.json names the capture it was read from)"""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


def pinned_u64(arr):
    import torch
    t = torch.empty(arr.shape, dtype=torch.int64).pin_memory()
    t.numpy().view(np.uint64)[...] = arr
    return t


def random_sector_ceiling(table_bytes, n_access=100_000_000):
    """SURVEY.md 8d's second ceiling, measured live: uniformly random 32-B sectors of a buffer of the table's size,
    read-only (find / count) and read + atomic (insert / erase); sectors per second"""
    import kmerhash_b200 as kh
    buf = 1 << int(np.ceil(np.log2(max(table_bytes, 1 << 20))))
    out = {"footprint_bytes": int(buf)}
    for mode in ("read", "rmw"):
        ms = kh.probe_random_access(buf, n_access, 32, mode, iters=3)
        out["sectors_per_s_" + mode] = n_access / (ms * 1e-3)
    return out


def ours_single(args):
    import torch
    import kmerhash_b200 as kh
    assert torch.cuda.is_available() and kh.device_count() > 0, "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(0)
    n, nq = args.n_keys, args.n_query
    if args.config == "1":
        keys, queries = generate_workload(n, nq, seed=23)
        d_keys = torch.from_numpy(keys.view(np.int64)).cuda()
        workload = "BenchmarkHashTables default (BASELINE.json configs[0])"
    else:
        keys = None
        d_keys = device_keys_config5(n, 0)
        workload = "BenchmarkDistHashTables shape (BASELINE.json configs[4], 1 GPU point): uint64 keys, every distinct key 1..8 times"
    d_q = d_keys[:nq].contiguous()
    m = kh.HashMap(1, args.hash, "COUNT", "ROBINHOOD_OFFSETS", device=0, min_load=0.35, max_load=0.8)
    KINDS = ("insert", "scatter", "level2", "hll", "rebuild", "lookup")

    def ev():
        e = torch.cuda.Event(enable_timing=True); e.record(); return e

    def step_resident():
        m.clear(); m.rehash(128)                      # default-constructed table, as the driver starts with
        m.kernel_timing(True)                         # resets the per-kind kernel clocks
        e0 = ev(); new = m.insert(d_keys, estimate=True)
        e1 = ev()
        k_ins = {k: m.kernel_time(k) for k in KINDS}  # kernels of the insert phase (outside the timed phases)
        cap_after_insert = m.capacity()
        m.kernel_timing(True)
        e1b = ev(); f = m.find(d_q, 0)
        e2 = ev(); c = m.count(d_q)
        e3 = ev(); er = m.erase(d_q)
        e4 = ev(); c2 = m.count(d_q)
        e5 = ev(); torch.cuda.synchronize()
        k_rest = {k: m.kernel_time(k) for k in KINDS}
        t = [a.elapsed_time(b) for a, b in ((e0, e1), (e1b, e2), (e2, e3), (e3, e4), (e4, e5))]
        return t, (new, int((f != 0).sum().item()), int(c.sum().item()), er, int(c2.sum().item())), k_ins, k_rest, cap_after_insert

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(0); sampler.start()
    kh.reset_launch_count()
    torch.cuda.synchronize()
    times, checks, cap = [], None, 0
    kin = {k: [0.0, 0] for k in KINDS}
    krest = {k: [0.0, 0] for k in KINDS}
    for _ in range(args.steps):
        t, checks, k_ins, k_rest, cap = step_resident()
        times.append(t)
        for k in KINDS:
            kin[k][0] += k_ins[k][0]; kin[k][1] += k_ins[k][1]
            krest[k][0] += k_rest[k][0]; krest[k][1] += k_rest[k][1]
    torch.cuda.synchronize()
    launches = kh.launch_count()
    m.kernel_timing(False)
    S = float(args.steps)
    times = np.array(times)
    ph = times.mean(axis=0)
    # sanity: all queries are present, erased ones are gone
    assert checks[1] == nq and checks[2] == nq and checks[4] == 0, checks
    distinct = checks[0]
    if keys is not None and n <= 200_000_000:
        assert distinct == int(np.unique(keys).shape[0]), "distinct keys differ from numpy's exact count"

    peak, peak_src = measured_peaks()
    table_bytes = cap * 16
    # per-kernel OWN algorithmic bytes: level 1 reads the key and writes one 16-byte record; level 2 reads
    # and writes a record; the tile insert reads a record and writes every slot of the (fresh) table once
    own = {"scatter": ("l1_scatter_kernel (hash + HLL registers + partition into table regions: 8 B read + 16 B record written per key)", 24.0 * n),
           "level2": ("l2_hist/l2_scatter_kernel (split of the regions into table tiles: 4 B tag + 16 B read, 16 B written per key)", 36.0 * n),
           "insert": ("tile_insert_kernel + insert_kernel for deferred keys (16 B record read per key + 16 B per table slot written)", 16.0 * n + 16.0 * cap),
           "hll": ("hll_update_kernel (stand-alone HyperLogLog pass, 8 B per key)", 8.0 * n),
           "rebuild": ("rebuild kernels inside the insert phase", 32.0 * cap)}
    ncu_name = {"scatter": "l1_scatter_kernel", "insert": "tile_insert_kernel", "level2": "l2_scatter_kernel", "rebuild": "tile_rebuild_kernel"}
    kernels, kernel_ms_sum, traffic = {}, 0.0, 0
    for k in ("scatter", "level2", "insert", "hll", "rebuild"):
        ms, cnt = kin[k][0] / S, kin[k][1] / S
        if cnt == 0:
            continue
        kernel_ms_sum += ms
        gbps = own[k][1] / (ms * 1e-3) / 1e9
        tr = ncu_traffic(ncu_name.get(k, ""))
        if tr:
            traffic += tr
        kernels[own[k][0]] = {"ms_per_step": ms, "launches_per_step": cnt, "own_algorithmic_GBps": gbps, "frac_of_peak": gbps / peak,
                              "ncu_dram_bytes_per_launch": tr}
    achieved = BYTES_PER_INSERT * n / (ph[0] * 1e-3) / 1e9          # the whole insert PHASE on the 72 B/key budget
    rs = random_sector_ceiling(table_bytes)
    ins_rate, find_rate = n / (ph[0] * 1e-3), nq / (ph[1] * 1e-3)
    line = {
        "metric": "k-mer inserts/s (k=31, count table)", "value": ins_rate, "unit": "inserts/s",
        "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(ph.sum()),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "%s: %d random 31-mer uint64 keys (%d distinct) insert, then %d find/count/erase/count; "
                               "robinhood_offsets count map; 1 B200" % (workload, n, distinct, nq),
                   "n_keys": n, "n_query": nq, "hash": args.hash, "load_factors": [0.35, 0.8], "config": args.config,
                   "input": ("generate_input of the reference (srand(23), glibc rand(), libstdc++ random_shuffle) via drivers/libkh_geninput.so"
                             if args.config == "1" else "torch generator on the device, seed 1000 + rank"),
                   "l2": "inputs (%.1f GB keys, %.2f GB table) larger than the 126 MB L2; table rebuilt from capacity 128 every step"
                         % (8e-9 * n, table_bytes * 1e-9),
                   "insert_phase": "HLL estimate + reserve + batch insert (the reference's insert())"},
        "finds_per_s": find_rate, "counts_per_s": nq / (ph[2] * 1e-3),
        "erases_per_s": nq / (ph[3] * 1e-3), "count2_per_s": nq / (ph[4] * 1e-3),
        "phase_ms": {"insert": float(ph[0]), "find": float(ph[1]), "count": float(ph[2]), "erase": float(ph[3]), "count2": float(ph[4])},
        "roofline": {"bound": "hbm",
                     "kernel": "insert phase = every kernel an insert needs (level-1 scatter with the HyperLogLog registers, level-2 split "
                               "when the table has more tiles than regions, tile insert, global insert kernel for deferred keys), "
                               "CUDA-event time of the phase incl. the estimate read-back and reserve between them",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic or None, "peak_source": peak_src,
                     "algorithmic_bytes_per_key": BYTES_PER_INSERT, "phase_ms": float(ph[0]),
                     "kernel_ms_per_step": kernel_ms_sum, "share_of_step": kernel_ms_sum / float(ph.sum()),
                     "kernels": kernels,
                     "find": {"achieved": BYTES_PER_FIND * nq / (ph[1] * 1e-3) / 1e9, "frac": BYTES_PER_FIND * nq / (ph[1] * 1e-3) / 1e9 / peak,
                              "algorithmic_bytes_per_key": BYTES_PER_FIND,
                              "lookup_kernels_ms_per_step": krest["lookup"][0] / S},
                     "post_erase_rebuild_ms_per_step": krest["rebuild"][0] / S,
                     # second ceiling: random 32-B sectors at the table's footprint, measured in this run
                     "random_sector": dict(rs, insert_frac=2.0 * ins_rate / rs["sectors_per_s_rmw"],
                                           find_frac=find_rate / rs["sectors_per_s_read"],
                                           note="insert = 2 sectors per key (read + write back), find = 1; a fraction above 1 means the "
                                                "path left the random-access regime (the bulk insert streams)")},
        "gpu_launches": int(launches),
    }

    if not args.no_e2e and keys is not None:
        hk, hq = pinned_u64(keys), pinned_u64(queries)
        hk_np, hq_np = hk.numpy().view(np.uint64), hq.numpy().view(np.uint64)
        out32 = torch.empty(nq, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
        out8 = torch.empty(nq, dtype=torch.uint8).pin_memory().numpy()

        def step_e2e():
            m.clear(); m.rehash(128)
            t0 = time.perf_counter(); m.insert(hk_np, estimate=True)
            t1 = time.perf_counter(); m.find(hq_np, 0, out=out32)
            t2 = time.perf_counter(); m.count(hq_np, out=out8)
            t3 = time.perf_counter(); m.erase(hq_np)
            t4 = time.perf_counter(); m.count(hq_np, out=out8)
            t5 = time.perf_counter()
            return [1e3 * (b - a) for a, b in ((t0, t1), (t1, t2), (t2, t3), (t3, t4), (t4, t5))]

        for _ in range(args.warmup):
            step_e2e()
        et = np.array([step_e2e() for _ in range(args.steps)]).mean(axis=0)
        assert int(out8.sum()) == 0 and int((out32 != 0).sum()) == nq
        line["e2e"] = {"value": n / (et[0] * 1e-3), "unit": "inserts/s",
                       "h2d_bytes_per_step": int(8 * n + 4 * 8 * nq), "d2h_bytes_per_step": int(4 * nq + 2 * nq),
                       "finds_per_s": nq / (et[1] * 1e-3), "ms_per_step": float(et.sum()),
                       "phase_ms": {k: float(v) for k, v in zip(("insert", "find", "count", "erase", "count2"), et)},
                       "api": "kh_insert_host / kh_find_host / kh_count_host / kh_erase_host on pinned host buffers"}
    line["clocks"] = sampler.stop()      # sampled over the resident AND the end-to-end timed regions
    if not args.no_cpu_baseline and keys is not None:
        line["cpu_baseline"] = cpu_baseline(keys, queries, args.hash)
    print(json.dumps(line), flush=True)


def multi_parity_check(args, rank, world, dev):
    """before anything is timed: a 6 M-key batch per rank (large enough for the pipelined exchange and the bulk insert) through
    insert / find / count / erase of a distributed map of its own;
    every rank's table contents must equal {(k, c) of the global count map : owner(k) == rank}, answers must come back in
    input order.  Expected values from numpy (exact counts) and the owner rule restated above.  Raises on mismatch."""
    import torch
    import torch.distributed as dist
    from kmerhash_b200.distributed import DistributedMap
    dm = DistributedMap(1, args.hash, args.hash, "COUNT", canonical_k=0, device=dev.index)
    dm.e.table.set_load_factors(0.35, 0.8)
    nb = 6_000_000
    rng = np.random.default_rng(777)                      # the SAME global pool on every rank
    pool = rng.integers(0, 1 << 62, size=4_000_000, dtype=np.uint64)
    mine = pool[np.random.default_rng(1000 + rank).integers(0, len(pool), size=nb)]
    batches = [pool[np.random.default_rng(1000 + r).integers(0, len(pool), size=nb)] for r in range(world)]
    allk = np.concatenate(batches)
    uk, uc = np.unique(allk, return_counts=True)
    owner = (np_murmur3_x86_128_lo64(uk, DIST_SEED) & np.uint64(world - 1)).astype(np.int64) if (world & (world - 1)) == 0 else \
        (np_murmur3_x86_128_lo64(uk, DIST_SEED) % np.uint64(world)).astype(np.int64)
    d = torch.from_numpy(mine.view(np.int64)).to(dev)
    dm.insert(d)
    dist.barrier()
    gk, gv = dm.to_vector()
    o = np.argsort(gk)
    sel = owner == rank
    ok = np.array_equal(gk[o], uk[sel]) and np.array_equal(gv[o].astype(np.int64), uc[sel].astype(np.int64))
    # queries: half present, half absent, answers in input order
    q = np.concatenate([mine[:50_000], rng.integers(1 << 62, (1 << 63) - 1, size=50_000, dtype=np.uint64)])
    dq = torch.from_numpy(q.view(np.int64)).to(dev)
    f = dm.find(dq, 0).cpu().numpy().astype(np.int64)
    c = dm.count(dq).cpu().numpy()
    exp = np.zeros(len(q), dtype=np.int64)
    pos = np.searchsorted(uk, q[:50_000])
    exp[:50_000] = uc[pos]
    ok = ok and np.array_equal(f, exp) and np.array_equal(c.astype(np.int64), (exp != 0).astype(np.int64))
    er = dm.erase(dq)
    tot = torch.tensor([er], device=dev, dtype=torch.int64); dist.all_reduce(tot)
    allq = np.unique(np.concatenate([b[:50_000] for b in batches]))
    ok = ok and int(tot.item()) == len(allq)
    c2 = dm.count(dq).cpu().numpy()
    ok = ok and int(c2.sum()) == 0
    flag = torch.tensor([1 if ok else 0], device=dev, dtype=torch.int64)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) != 1:
        raise SystemExit("bench.py: distributed parity check FAILED on rank %d (ok=%s)" % (rank, ok))
    dm.close()
    return {"keys_per_rank": nb, "distinct_global": int(len(uk)), "checked": "per-rank sorted (key,count) == numpy count map restricted "
            "to owner(k) == rank (murmur3 x86_128 seed 9876543 restated in numpy); find/count in input order; erase; count after erase"}


def ours_multi(args):
    """N > 1: one process per GPU under torchrun, exchange over peer memory, weak scaling (each rank brings n keys)."""
    import torch
    import torch.distributed as dist
    import kmerhash_b200 as kh
    from kmerhash_b200.distributed import DistributedMap
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", ""):
        os.environ["NCCL_DEBUG"] = "WARN"          # keep NCCL's version banner out of stdout (one JSON line)
    dist.init_process_group("nccl", device_id=dev)
    n, nq = args.n_keys, args.n_query
    if args.config == "1":
        keys, queries = generate_workload(n, nq, seed=23 + rank)
        d_keys = torch.from_numpy(keys.view(np.int64)).cuda()
    else:
        keys = None
        d_keys = device_keys_config5(n, rank)
    d_q = d_keys[:nq].contiguous()
    dm = DistributedMap(1, args.hash, args.hash, "COUNT", canonical_k=0, device=local)
    dm.e.table.set_load_factors(0.35, 0.8)
    parity = None if args.no_parity else multi_parity_check(args, rank, world, dev)

    f_host = torch.empty(nq, dtype=torch.int32).pin_memory()

    def barrier():
        dist.barrier(); torch.cuda.synchronize()

    def reset():
        dm.e.table.clear(); dm.e.table.rehash(128)

    import kmerhash_b200.distributed as kd
    stage_log = []                                                    # host-side stage times of every find step on rank 0

    def step(k_in, q_in, to_host):
        reset(); barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e2 = torch.cuda.Event(enable_timing=True)
        e0.record(); dm.insert(k_in); xb = dm.last_exchange_bytes
        barrier(); e1.record()
        kd.STAGE_LOG = []
        t0 = time.perf_counter(); f = dm.find(q_in, 0); t1 = time.perf_counter()
        if to_host:
            f_host.copy_(f); f = f_host                               # D2H into pinned memory (a fresh pageable tensor costs ~10 ms of page faults)
        barrier(); t2 = time.perf_counter(); e2.record(); torch.cuda.synchronize()
        stage_log.append(dict(kd.STAGE_LOG, find_call=round(1e3 * (t1 - t0), 3), barrier_after=round(1e3 * (t2 - t1), 3)))
        kd.STAGE_LOG = None
        return e0.elapsed_time(e1), e1.elapsed_time(e2), f, xb

    step_log = []                                                     # rank 0's per-step times (device and e2e runs)

    def run(k_in, q_in, to_host):
        f = None
        for _ in range(args.warmup):
            # the SAME object lifetimes as the timed loop: the previous step's result stays referenced while the next step
            # allocates its own, so the caching allocator reaches its steady state (two result blocks) during the warm-up.
            # (Round 1's "one slow find step per run" was exactly that: the first cudaMalloc of a second 40 MB block, inside
            # the second timed step -- 2-50 ms with IPC-mapped peer buffers present.)
            _, _, f, _ = step(k_in, q_in, to_host)
        gc.collect(); gc.disable()                                    # no collector pause inside a sub-millisecond phase
        ti, tf, xb = [], [], 0
        for _ in range(args.steps):
            a, b, f, xb = step(k_in, q_in, to_host)
            ti.append(a); tf.append(b)
        t = torch.tensor([float(np.mean(ti)), float(np.mean(tf)), float(np.median(ti)), float(np.median(tf)),
                          float(np.max(tf))], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)                      # max over ranks
        assert int((f != 0).sum()) == nq                              # every query key is present
        gc.enable()
        step_log.append({"insert_ms": [round(x, 3) for x in ti], "find_ms": [round(x, 3) for x in tf],
                         "find_stages_host_ms": stage_log[-args.steps:]})
        return t.tolist(), xb

    sampler = ClockSampler(local)
    if rank == 0 and os.environ.get("KH_NO_SAMPLER", "0") in ("", "0"):
        sampler.start()
        time.sleep(float(os.environ.get("KH_SAMPLER_LEAD_S", "0.5")))   # let the sampler's start-up settle before anything is timed
    kh.reset_launch_count()
    (ti, tf, ti_med, tf_med, tf_max), xbytes = run(d_keys, d_q, False)
    launches = torch.tensor([kh.launch_count()], device="cuda", dtype=torch.int64)
    dist.all_reduce(launches)
    clocks = sampler.stop() if (rank == 0 and (sampler.nvml or sampler.proc)) else None
    sizes = torch.tensor([dm.local_size()], device="cuda", dtype=torch.int64)
    dist.all_reduce(sizes)
    e2e = None
    if not args.no_e2e and keys is not None:
        hk, hq = pinned_u64(keys), pinned_u64(queries)
        (ei, ef, _ei_med, ef_med, _), _ = run(hk, hq, True)
        e2e = {"value": world * n / (ei * 1e-3), "unit": "inserts/s", "h2d_bytes_per_step": int(8 * (n + nq)),
               "d2h_bytes_per_step": int(4 * nq), "finds_per_s": world * nq / (ef * 1e-3), "finds_per_s_median": world * nq / (ef_med * 1e-3),
               "api": "DistributedMap.insert / .find on pinned host tensors (per rank)"}
    if rank == 0:
        peak, peak_src = measured_peaks()
        achieved = BYTES_PER_INSERT * n / (ti * 1e-3) / 1e9           # per GPU: its n keys over the whole insert phase (sender, barrier, receiver)
        line = {
            "metric": "k-mer inserts/s (k=31, count table)", "value": world * n / (ti * 1e-3), "unit": "inserts/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ti + tf,
            "higher_is_better": True, "scaling": "weak" if args.config != "5s" else "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "BenchmarkDistHashTables-shaped: %d random 31-mer uint64 keys per GPU, "
                                   "hash-partitioned (murmur3 x86_128, seed 9876543) over %d GPUs, exchange over NVLink into CUDA-IPC mapped peer buffers (one piece: the sender kernel stores into the owners' slabs; "
                                   "several pieces: partition into local staging + copy engines, overlapped with the receiver's level 1; NCCL as fallback), "
                                   "count map insert, then %d finds per GPU with reply" % (n, world, nq),
                       "n_keys_per_gpu": n, "n_query_per_gpu": nq, "hash": args.hash, "config": args.config,
                       "exchange_pieces": dm.pipe_chunks,
                       "exchange_schedule": ("one piece, SM stores" if dm.pipe_chunks <= 1 else
                                             "%d pieces, %s" % (dm.pipe_chunks, "copy engines" if os.environ.get("KH_PIPE_MODE", "ce") == "ce" else "SM stores")),
                       "input": ("generate_input of the reference with srand(23 + rank)" if args.config == "1" else "torch generator on the device, seed 1000 + rank, every distinct key 1..8 times"),
                       "l2": "inputs larger than L2; tables rebuilt from capacity 128 every step"},
            "finds_per_s": world * nq / (tf * 1e-3), "finds_per_s_median": world * nq / (tf_med * 1e-3),
            "phase_ms": {"insert": ti, "find": tf, "insert_median": ti_med, "find_median": tf_med, "find_max": tf_max},
            "find_step_max_over_median": tf_max / tf_med, "rank0_steps": step_log[:1],
            "distinct_keys_total": int(sizes.item()),
            "parity_check": parity,
            "exchange": {"bytes_sent_per_rank_insert": int(xbytes), "GB_per_s_per_rank_if_serial": xbytes / (ti * 1e-3) / 1e9,
                         "nvlink_peak_GB_per_s_per_direction": 900},
            "roofline": {"bound": "hbm", "kernel": "insert phase per GPU (owner partition / sender scatter, transfer over NVLink, flag barriers, receiver's level-1 scatter + tile insert)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_key": BYTES_PER_INSERT},
            "gpu_launches": int(launches.item()), "clocks": clocks,
        }
        if e2e:
            line["e2e"] = e2e
        print(json.dumps(line), flush=True)
    dist.barrier()
    if hasattr(dm, "close"):
        dm.close()
    dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        reference_arm(args)
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        ours_multi(args)
    else:
        ours_single(args)


if __name__ == "__main__":
    main()
