python -m pytest tests/test_gpu_parity.py tests/test_gpu_garbler.py tests/test_gpu_configs.py -x -q -m gpu -k "RESCALE or rescale or CIFAR or cifar or scaled or live" 2>&1 | tail -3
python bench.py --only c3 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['inferences_per_s'], d['bit_exact_vs_oracle'], d['busy_ms_by_layer_type'])"
