timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden or relu or RELU or replicas" > gpurun_out/u_pytest.log 2>&1; tail -3 gpurun_out/u_pytest.log
timeout 600 python bench.py --no-extras > gpurun_out/u_bench.json 2> gpurun_out/u_bench.err
python - <<'PY'
import json
l=[x for x in open('gpurun_out/u_bench.json').read().splitlines() if x.startswith('{')][-1]
d=json.loads(l); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'])
PY
python bench.py --only c3 > gpurun_out/u_c3.json 2> gpurun_out/u_c3.err; tail -c 600 gpurun_out/u_c3.json
