timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
python tools/probe_async.py 2>&1 | tail -1
DASHB_ASYNC_MODE=sets python tools/probe_async.py 2>&1 | tail -1
DASHB_DIRECT_EGRESS=0 python tools/probe_async.py 2>&1 | tail -1
DASHB_TRACE_ASYNC=1 python tools/probe_async.py > gpurun_out/z_inorder.log 2>&1
tail -4 gpurun_out/z_inorder.log | cut -c1-420
