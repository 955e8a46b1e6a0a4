ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,dram__bytes_read.sum --clock-control none -k regex:"k_rescale|k_sign|k_dense_tc|k_conv_tc|k_linear" -s 120 -c 80 --csv --log-file gpurun_out/r2_c3_launches.csv python bench.py --only c3 > gpurun_out/ncu_c3.log 2>&1
python - <<'PY'
import csv,collections
rows=list(csv.reader(open('gpurun_out/r2_c3_launches.csv')))
i0=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[i0]; data=rows[i0+1:]
ik=h.index('Kernel Name'); im=h.index('Metric Name'); iv=h.index('Metric Value'); ig=h.index('Grid Size')
agg=collections.OrderedDict()
for r in data:
    if len(r)<=iv: continue
    key=(r[0],r[ik][:40],r[ig])
    agg.setdefault(key,{})[r[im]]=r[iv]
for (i,k,g),m in list(agg.items())[:60]:
    print(i,k,g,m.get('gpu__time_duration.sum'),m.get('smsp__inst_executed.sum'),m.get('smsp__issue_active.avg.pct_of_peak_sustained_active'),m.get('sm__warps_active.avg.pct_of_peak_sustained_active'),m.get('launch__registers_per_thread'),m.get('dram__bytes_read.sum'))
PY
