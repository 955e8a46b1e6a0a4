# full GPU suite, smoke, default bench, conv launch lists with and without strips
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r_pytest.log
tail -3 gpurun_out/r_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r_smoke.log 2>&1; tail -2 gpurun_out/r_smoke.log
timeout 900 python bench.py > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err; tail -c 600 gpurun_out/r_bench.json
M="gpu__time_duration.sum,smsp__inst_executed.sum,lts__t_bytes.sum,dram__bytes_read.sum"
timeout 600 ncu --metrics $M --clock-control none -k regex:"k_conv_tc" -s 16 -c 24 --csv --log-file gpurun_out/r_c3_conv_strip.csv python bench.py --only c3 > gpurun_out/r_ncu1.log 2>&1
DASHB_CONV_STRIP=0 timeout 600 ncu --metrics $M --clock-control none -k regex:"k_conv_tc" -s 16 -c 24 --csv --log-file gpurun_out/r_c3_conv_nostrip.csv python bench.py --only c3 > gpurun_out/r_ncu2.log 2>&1
