python tools/probe_async.py 2>&1 | tail -1
DASHB_INGEST_CTAS=128 python tools/probe_async.py 2>&1 | tail -1
DASHB_INGEST_CTAS=296 python tools/probe_async.py 2>&1 | tail -1
DASHB_INGEST_CTAS=592 python tools/probe_async.py 2>&1 | tail -1
