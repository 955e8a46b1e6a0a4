run() { echo "== $*"; env "$@" python tools/probe_ingest.py $MODE 2>&1 | grep -E "wall|move_inputs" | tail -3; }
MODE=direct run X=1
MODE=direct run DASHB_LANE_SPLIT=1,3,4,4,3,1
MODE=direct run DASHB_LANE_SPLIT=1,2,3,4,3,2,1
MODE=direct run DASHB_LANE_SPLIT=2,3,4,4,2,1
MODE=direct run DASHB_LANE_SPLIT=1,3,4,4,3,1 DASHB_INGEST_CTAS=48
MODE=direct run DASHB_LANE_SPLIT=1,3,4,4,3,1 DASHB_INGEST_CTAS=192
MODE=staged run DASHB_LANE_SPLIT=1,3,4,4,3,1
MODE=direct DASHB_LANE_SPLIT=1,3,4,4,3,1 python tools/probe_ingest.py direct 2>&1 | grep trace | tail -6
DASHB_LANE_SPLIT=1,3,4,4,3,1 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench4.json 2> gpurun_out/r2_bench4.err; tail -3 gpurun_out/r2_bench4.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench4.json')); print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'comp', d['e2e_compressed']['value'], 'vals', d['e2e_values']['value'])"
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "direct_ingest" 2>&1 | tail -2
DASHB_LANE_SPLIT=1,3,4,4,3,1 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "direct_ingest or full_size or golden_final" 2>&1 | tail -2
