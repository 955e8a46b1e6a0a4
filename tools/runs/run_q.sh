# c3/c4 with and without the strip convolution kernel
python bench.py --only c3 > gpurun_out/q_c3_strip.json 2> gpurun_out/q_c3_strip.err
DASHB_CONV_STRIP=0 python bench.py --only c3 > gpurun_out/q_c3_nostrip.json 2> gpurun_out/q_c3_nostrip.err
python bench.py --only c4 > gpurun_out/q_c4_strip.json 2> gpurun_out/q_c4_strip.err
DASHB_CONV_STRIP=0 python bench.py --only c4 > gpurun_out/q_c4_nostrip.json 2> gpurun_out/q_c4_nostrip.err
tail -c 1500 gpurun_out/q_c3_strip.json; echo; tail -c 1500 gpurun_out/q_c3_nostrip.json
