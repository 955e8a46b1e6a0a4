python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 > gpurun_out/x_bench8_final.json 2> gpurun_out/x_bench8_final.err
python - <<'PY'
import json
l=[x for x in open('gpurun_out/x_bench8_final.json').read().splitlines() if x.startswith('{')][-1]
d=json.loads(l)
print(d['value'], d['e2e']['value'], d['e2e_compressed_pipelined']['value'], d['c3_batch_sharded']['inferences_per_s'], d['c4_channel_sharded']['ms_per_step'], d['c4_channel_sharded']['bit_exact_vs_oracle'])
PY
