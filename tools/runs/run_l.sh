python -m pytest tests/test_gpu_p2p_shard.py tests/test_gpu_channel_shard.py -x -q -m gpu 2>&1 | tail -15
python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -x -q -m gpu 2>&1 | tail -3
