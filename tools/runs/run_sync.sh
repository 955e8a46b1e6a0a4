DASHB_DIRECT_INGEST=0 python bench.py --no-extras --no-cpu-baseline > gpurun_out/sync_staged.json 2>/dev/null
python - <<'PY'
import json
l=[x for x in open('gpurun_out/sync_staged.json').read().splitlines() if x.startswith('{')][-1]
d=json.loads(l); print('staged sync', d['e2e_sync']['value'], d['e2e_sync']['ms_per_step'], 'e2e', d['e2e']['value'])
PY
