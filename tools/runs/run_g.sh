DASHB_DEBUG_SKIP_INPUTS=1 python tools/probe_async.py 2>&1 | tail -1
DASHB_DEBUG_SKIP_INPUTS=1 DASHB_LANES=8 python tools/probe_async.py 2>&1 | tail -1
python tools/h2d_bandwidth.py 2>&1 | tail -5
