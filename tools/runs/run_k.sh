python -m pytest tests/test_gpu_parity.py tests/test_gpu_channel_shard.py tests/test_gpu_garbler.py -x -q -m gpu 2>&1 | tail -3
DASHB_QUAD_MAX=100000000 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or live or full_size" 2>&1 | tail -3
for Q in 0 75000 150000 300000; do
DASHB_QUAD_MAX=$Q python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r2_bench7_$Q.json 2> gpurun_out/r2_bench7.err; tail -2 gpurun_out/r2_bench7.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench7_$Q.json')); print('quad_max $Q', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'sync', d['e2e_sync']['value'], d['roofline']['busy_ms_by_layer_type'])"
done
