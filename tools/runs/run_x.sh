# 8-GPU bench (headline + c3 + c4 legs)
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 > gpurun_out/x_bench8.json 2> gpurun_out/x_bench8.err
tail -c 1500 gpurun_out/x_bench8.json
