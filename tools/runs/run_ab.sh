cp dash_b200/libdash_b200.so /tmp/_default.so
for v in base sign7 sign6 sign5; do
  cp build_variants/$v.so dash_b200/libdash_b200.so
  echo "# $v"
  python tools/batch_scaling.py MODEL_B_POOL_REPL 256 256 | tail -1
  python tools/batch_scaling.py MODEL_F_GNNP_POOL_REPL 64 | tail -1
done 2>&1 | grep -v "^$" | tee gpurun_out/ab_minb2.log
cp /tmp/_default.so dash_b200/libdash_b200.so
