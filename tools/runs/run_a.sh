set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -x -q -m gpu > gpurun_out/r2_t2.log 2>&1; echo "rc=$?" >> gpurun_out/r2_t2.log
tail -5 gpurun_out/r2_t2.log
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench1.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['layer_ms'])"
ncu --set full --clock-control none --import-source on -k regex:k_sign -s 8 -c 1 -o gpurun_out/r2_ksign_a -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_a.log 2>&1
ls -la gpurun_out/r2_ksign_a.ncu-rep
