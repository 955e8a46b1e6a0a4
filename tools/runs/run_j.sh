python -m pytest tests/test_gpu_p2p_shard.py tests/test_gpu_channel_shard.py -x -q -m gpu 2>&1 | tail -15
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench6.json 2> gpurun_out/r2_bench6.err; tail -3 gpurun_out/r2_bench6.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench6.json')); print(d['value'], 'e2e', d['e2e']['value']); print(json.dumps(d.get('c3_batch_sharded'), indent=1)); print(json.dumps(d.get('c4_channel_sharded'), indent=1))"
