python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "replicas or golden_final" 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench8.json 2> gpurun_out/r2_bench8.err; tail -3 gpurun_out/r2_bench8.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench8.json')); print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value']); print(d.get('per_inference_tables'))"
