B="python bench.py --steps 1 --warmup 0 --individuals-per-gpu 8"
$B > gpurun_out/r1_plain3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 44000 --csv --log-file gpurun_out/r1_launches3.csv $B > gpurun_out/r1_ncu_l3.log 2>&1
python - <<'PY'
import csv, collections, re
lines=[l for l in open('gpurun_out/r1_launches3.csv') if not l.startswith('==')]
agg=collections.defaultdict(lambda:[0,0.0])
for row in csv.DictReader(lines):
    if row.get('Metric Name')!='gpu__time_duration.sum': continue
    name=re.sub(r'\(.*','',row['Kernel Name']); g=row['Grid Size']
    v=float(row['Metric Value'].replace(',','')); u=row['Metric Unit']
    v = v/1000 if u=='ns' else (v*1000 if u=='ms' else v)
    agg[(name,g)][0]+=1; agg[(name,g)][1]+=v
tot=sum(v[1] for v in agg.values())
with open('gpurun_out/r1_launches3_summary.txt','w') as out:
    out.write(f'total us {tot:.1f} launches {sum(v[0] for v in agg.values())}\n')
    for k,v in sorted(agg.items(), key=lambda kv:-kv[1][1]):
        out.write(f"{v[1]:10.1f} us {100*v[1]/tot:5.1f}%  n={v[0]:5d} avg={v[1]/v[0]:8.2f}  {k[0][:70]} {k[1]}\n")
PY
rm -f gpurun_out/r1_launches3.csv
head -60 gpurun_out/r1_launches3_summary.txt
