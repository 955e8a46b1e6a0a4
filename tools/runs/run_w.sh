# A/B: hash() inlined (rolled rounds) so that table-row loads stay in flight across it
cp dash_b200/libdash_b200.so /tmp/_default.so
for v in base hash_inline; do
  cp build_variants/$v.so dash_b200/libdash_b200.so
  echo "# $v"
  python tools/batch_scaling.py MODEL_B_POOL_REPL 256 256 1024
  python tools/batch_scaling.py MODEL_F_GNNP_POOL_REPL 64
done 2>&1 | grep -v "^$" | tee gpurun_out/w_ab.log
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "golden" 2>&1 | tail -2
cp /tmp/_default.so dash_b200/libdash_b200.so
