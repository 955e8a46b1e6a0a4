python tools/probe_ingest.py direct 2>&1 | grep -v Warn | tail -12
python tools/probe_ingest.py staged 2>&1 | tail -12
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench3.json 2> gpurun_out/r2_bench3.err; tail -3 gpurun_out/r2_bench3.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench3.json')); print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value']); print(json.dumps(d['roofline'], indent=1))"
ncu --set full --clock-control none --import-source on -k regex:k_sign -s 48 -c 1 -o gpurun_out/r2_ksign_big -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_c.log 2>&1
ncu -i gpurun_out/r2_ksign_big.ncu-rep --page raw --csv | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]
for r in rows[2:]:
    print(r[h.index('Grid Size')], r[h.index('gpu__time_duration.sum')])"
