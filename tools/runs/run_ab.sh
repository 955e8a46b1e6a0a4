cp dash_b200/libdash_b200.so /tmp/_default.so
for v in lin_u1 lin_u2 lin_u4; do
  cp build_variants/$v.so dash_b200/libdash_b200.so
  echo "# $v"
  python tools/batch_scaling.py MODEL_B_POOL_REPL 256 256
done 2>&1 | grep -v "^$" | tee gpurun_out/ab_lin.log
cp /tmp/_default.so dash_b200/libdash_b200.so
