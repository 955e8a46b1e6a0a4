python -m pytest tests -x -q -m gpu > gpurun_out/r2_t_full.log 2>&1; echo "rc=$?" >> gpurun_out/r2_t_full.log; tail -4 gpurun_out/r2_t_full.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
