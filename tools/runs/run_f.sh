python tools/probe_async.py 2>&1 | tail -1
DASHB_INGEST_CTAS=32 python tools/probe_async.py 2>&1 | tail -1
DASHB_INGEST_CTAS=64 python tools/probe_async.py 2>&1 | tail -1
DASHB_INGEST_CTAS=160 python tools/probe_async.py 2>&1 | tail -1
DASHB_LANES=2 python tools/probe_async.py 2>&1 | tail -1
DASHB_LANES=8 python tools/probe_async.py 2>&1 | tail -1
DASHB_DIRECT_INGEST=0 python tools/probe_async.py 2>&1 | tail -1
DASHB_DIRECT_INGEST=0 DASHB_LANES=2 python tools/probe_async.py 2>&1 | tail -1
