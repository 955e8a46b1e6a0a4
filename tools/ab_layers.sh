#!/bin/bash
# Per-layer-type device times (lanes serialised) of the builds in build_variants/.
#   gpurun -- 'bash tools/ab_layers.sh > gpurun_out/ab_layers.log 2>&1'
cd "$(dirname "$0")/.."
cp dash_b200/libdash_b200.so /tmp/_default.so
for v in ${@:-build_variants/*.so}; do
    cp "$v" dash_b200/libdash_b200.so
    echo "# $v"
    python tools/layer_times_gpu_garbled.py MODEL_B_POOL_REPL 256 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['model'], d['batch'], d['by_type_ms'], d['total_ms'])"
    python tools/layer_times_gpu_garbled.py MODEL_F_GNNP_POOL_REPL 16 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['model'], d['batch'], d['by_type_ms'], d['total_ms'])"
done
cp /tmp/_default.so dash_b200/libdash_b200.so
