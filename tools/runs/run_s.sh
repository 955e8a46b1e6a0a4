# e2e with the copy-engine ingest path instead of the zero-copy one; source-level capture of k_rescale
DASHB_DIRECT_INGEST=0 timeout 600 python bench.py --no-extras > gpurun_out/s_bench_staged.json 2> gpurun_out/s_bench_staged.err
tail -c 900 gpurun_out/s_bench_staged.json
timeout 600 python bench.py --no-extras > gpurun_out/s_bench_direct.json 2> gpurun_out/s_bench_direct.err
tail -c 900 gpurun_out/s_bench_direct.json
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_rescale -s 6 -c 1 -o gpurun_out/r2_rescale_src -f python bench.py --only c3 > gpurun_out/s_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
