for q in 0 75000 100000 200000 400000 1000000; do
  echo "# DASHB_QUAD_MAX=$q"
  DASHB_QUAD_MAX=$q python tools/batch_scaling.py MODEL_B_POOL_REPL 256 256 | tail -1
done 2>&1 | tee gpurun_out/quad.log
