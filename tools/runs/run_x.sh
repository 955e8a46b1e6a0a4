python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --no-extras > gpurun_out/x_bench8b.json 2> gpurun_out/x_bench8b.err
python - <<'PY'
import json
l=[x for x in open('gpurun_out/x_bench8b.json').read().splitlines() if x.startswith('{')][-1]
d=json.loads(l)
print(d['value'], d['e2e']['value'], d['e2e_sync']['value'], d['e2e_compressed_pipelined']['value'], d.get('host_affinity'))
PY
nproc; lscpu | grep -i "numa\|socket\|model name" | head
