python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "async or direct_ingest or MAXPOOL or PROJ" 2>&1 | tail -3
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench5.json 2> gpurun_out/r2_bench5.err; tail -3 gpurun_out/r2_bench5.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench5.json')); print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['outputs_equal_across_steps'], 'sync', d['e2e_sync']['value'], 'comp', d['e2e_compressed']['value'], 'vals', d['e2e_values']['value'])"
DASHB_DIRECT_INGEST=0 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench5s.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2_bench5s.json')); print('staged', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'sync', d['e2e_sync']['value'])"
