set -x
mkdir -p gpurun_out
for st in 0 4000 16000 64000; do
KH_L1_STAGGER_NS=$st timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/bench_st$st.json 2> gpurun_out/bench_st.err; python -c "
import json,sys
d=json.load(open('gpurun_out/bench_st$st.json')); print('stagger $st', round(d['value']/1e9,2), d['phase_ms']['insert'], [round(v['ms_per_step'],3) for v in d['roofline']['kernels'].values()])"
done
timeout 600 python -m pytest tests/test_gpu_hash_kmer.py -m gpu -q -x -k "hash" 2>&1 | tail -n 3
timeout 600 python tools/bench_configs.py k63 --steps 2 > gpurun_out/config3_k63.jsonl 2> gpurun_out/config3.err; python -c "
import json
for l in open('gpurun_out/config3_k63.jsonl'):
    d=json.loads(l); print(d['hash'], round(d['inserts_per_s']/1e9,2), round(d['finds_per_s']/1e9,2), d['phase_ms'])"
exit 0
