timeout 900 python -m pytest tests/test_gpu_p2p_shard.py tests/test_gpu_channel_shard.py -x -q -m gpu 2>&1 | tail -3
DASHB_SHARD_COMPRESS=0 timeout 900 python -m pytest tests/test_gpu_p2p_shard.py -x -q -m gpu 2>&1 | tail -2
