# link rate of this box vs the pipelined e2e rate
python tools/h2d_bandwidth.py
python tools/probe_async.py
DASHB_DIRECT_INGEST=0 python tools/probe_async.py
DASHB_INGEST_CTAS=296 python tools/probe_async.py
DASHB_INGEST_CTAS=592 python tools/probe_async.py
DASHB_INGEST_CTAS=96 python tools/probe_async.py
