# round-end evidence: full GPU suite, smoke, the k_sign capture (ncu --set full) and the default bench of the same tree
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log
tail -3 gpurun_out/f_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/f_smoke.log 2>&1; tail -1 gpurun_out/f_smoke.log
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_sign -s 48 -c 1 -o gpurun_out/r2_ksign_final -f python bench.py --no-extras --no-cpu-baseline > gpurun_out/f_ncu_full.log 2>&1
ls -la gpurun_out/r2_ksign_final.ncu-rep
python tools/make_traffic.py gpurun_out/r2_ksign_final.ncu-rep 2880 64 > gpurun_out/f_traffic.log 2>&1
cp profiles/r2_traffic.json gpurun_out/f_r2_traffic.json
timeout 900 python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; tail -c 300 gpurun_out/f_bench.json
