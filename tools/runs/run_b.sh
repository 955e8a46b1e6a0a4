set -x
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "direct_ingest or golden_final or full_size" > gpurun_out/r2_t3.log 2>&1; echo "rc=$?" >> gpurun_out/r2_t3.log
tail -5 gpurun_out/r2_t3.log
for L in 4 8; do
DASHB_LANES=$L python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench2_l$L.json 2> gpurun_out/r2_bench2_l$L.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench2_l$L.json')); print('lanes $L', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['layer_ms'])"
done
DASHB_DIRECT_INGEST=0 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench2_staged.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2_bench2_staged.json')); print('staged', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])"
DASHB_TRACE=1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 >/dev/null | grep "dashb trace" | tail -8
ncu --set full --clock-control none --import-source on -k regex:k_sign -s 4 -c 1 -o gpurun_out/r2_ksign_b -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_b.log 2>&1
