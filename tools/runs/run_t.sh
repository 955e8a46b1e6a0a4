# paired-halves k_rescale + compressed async: parity, c3/c4 timing, headline bench, k_sign source-level capture
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -x -q -m gpu -k "RESCALE or rescale or CIFAR or cifar or BASE_EXT or async or SCALE or golden" > gpurun_out/t_pytest.log 2>&1; tail -3 gpurun_out/t_pytest.log
python bench.py --only c3 > gpurun_out/t_c3.json 2> gpurun_out/t_c3.err; tail -c 700 gpurun_out/t_c3.json
python bench.py --only c4 > gpurun_out/t_c4.json 2> gpurun_out/t_c4.err; tail -c 500 gpurun_out/t_c4.json
timeout 600 python bench.py --no-extras > gpurun_out/t_bench.json 2> gpurun_out/t_bench.err
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_sign -s 48 -c 1 -o gpurun_out/r2_ksign_src -f python bench.py --no-extras > gpurun_out/t_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
