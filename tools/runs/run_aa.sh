python tools/probe_h2d_under_load.py 2>&1 | tail -6
