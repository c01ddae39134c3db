# durations of the refinement + rescan launches (serialised) for option values: tools/refine_ab.py OPTION v...
mkdir -p gpurun_out/n
timeout 600 ncu --metrics gpu__time_duration.sum,lts__t_bytes.sum,l1tex__t_sector_hit_rate.pct --clock-control none -k regex:"k_tc_refine|k_tc_rescan" --csv --log-file gpurun_out/n/reflist.csv python tools/refine_ab.py "$@" > gpurun_out/n/reflist.log 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/n/reflist.csv')))
h=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
d={}
for r in rows[h+1:]:
    if len(r)>14: d.setdefault(int(r[0]),{'k':r[4][:34]})[r[12]]=r[14]
ids=sorted(d)
for i in ids[:4]+ids[len(ids)//2:len(ids)//2+4]: print(i, d[i])
PY
