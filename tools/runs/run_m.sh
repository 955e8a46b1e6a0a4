run() { echo "== $*"; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 4 --only c4 > gpurun_out/m.out 2> gpurun_out/m.err || tail -5 gpurun_out/m.err; tail -1 gpurun_out/m.out | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d.get('bit_exact_vs_oracle'), d['exchange'][:60])" || tail -8 gpurun_out/m.err; }
run X=1
run DASHB_PUSH_CTAS_PER_SM=6
run DASHB_C4_LANES=3
run DASHB_C4_LANES=4 DASHB_PUSH_CTAS_PER_SM=4
run DASHB_QUAD_MAX=300000
run DASHB_C4_EXCHANGE=nccl
