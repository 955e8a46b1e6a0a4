for l in 3 4 5 6; do
  echo "# DASHB_LANES=$l"
  DASHB_LANES=$l python tools/batch_scaling.py MODEL_B_POOL_REPL 256 256 | tail -1
done 2>&1 | tee gpurun_out/lanes.log
for sp in "5,4,4,3" "3,4,4,5" "6,5,3,2"; do
  echo "# DASHB_LANE_SPLIT=$sp"
  DASHB_LANE_SPLIT=$sp python tools/batch_scaling.py MODEL_B_POOL_REPL 256 256 | tail -1
done 2>&1 | tee -a gpurun_out/lanes.log
