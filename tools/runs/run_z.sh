timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -x -q -m gpu 2>&1 | tail -2
python tools/probe_async.py 2>&1 | tail -1
python tools/probe_async.py 2>&1 | tail -1
DASHB_INGEST_CTAS=96 python tools/probe_async.py 2>&1 | tail -1
DASHB_LANES=8 python tools/probe_async.py 2>&1 | tail -1
DASHB_LANES=2 python tools/probe_async.py 2>&1 | tail -1
