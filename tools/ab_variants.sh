#!/bin/bash
# A/B timing of the builds in build_variants/ on a B200 box: throughput of the
# benchmark model (ReLU-bound) and of the CIFAR model with CRT scaling
# (ReLU + rescale + tcgen05 convolutions), material from the GPU garbler.
#   gpurun -- 'bash tools/ab_variants.sh > gpurun_out/ab.log 2>&1'
cd "$(dirname "$0")/.."
cp dash_b200/libdash_b200.so /tmp/_default.so
for v in ${@:-build_variants/*.so}; do
    cp "$v" dash_b200/libdash_b200.so
    echo "# $v"
    python tools/batch_scaling.py MODEL_B_POOL_REPL 256 4096
    python tools/batch_scaling.py MODEL_F_GNNP_POOL_REPL 16
done
cp /tmp/_default.so dash_b200/libdash_b200.so
