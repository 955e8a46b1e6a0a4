T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512"
$T bench.py --gpus 8 --only c4 > gpurun_out/c48_comp.json 2> gpurun_out/c48_comp.err
DASHB_SHARD_COMPRESS=0 $T bench.py --gpus 8 --only c4 > gpurun_out/c48_plain.json 2> gpurun_out/c48_plain.err
for f in comp plain; do tail -1 gpurun_out/c48_$f.json | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$f', d['ms_per_step'], d['nvlink_bytes_received_per_rank_per_step'], d['bit_exact_vs_oracle'])"; done
