# A/B: L1 prefetches across hash calls in k_sign
cp dash_b200/libdash_b200.so /tmp/_default.so
for v in pf_none pf_rows pf_rows_x pf_all; do
  cp build_variants/$v.so dash_b200/libdash_b200.so
  echo "# $v"
  python tools/batch_scaling.py MODEL_B_POOL_REPL 256 256 1024
  python tools/batch_scaling.py MODEL_F_GNNP_POOL_REPL 64
done 2>&1 | grep -v "^$" | tee gpurun_out/v_ab.log
cp /tmp/_default.so dash_b200/libdash_b200.so
