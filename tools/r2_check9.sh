set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_table.py -m gpu -x -q --timeout 300 -k "cluster_region or level2 or poisons or barrier" > gpurun_out/pytest_cluster.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_cluster.log
tail -n 15 gpurun_out/pytest_cluster.log
for lp in 12 11 10; do
KH_CLUSTER=1 KH_BULK_LP=$lp timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/bench_lp$lp.json 2> gpurun_out/bench_lp.err; python -c "
import json
d=json.load(open('gpurun_out/bench_lp$lp.json')); print('lp $lp', round(d['value']/1e9,2), d['phase_ms']['insert'], [round(v['ms_per_step'],3) for v in d['roofline']['kernels'].values()])"
done
exit 0
