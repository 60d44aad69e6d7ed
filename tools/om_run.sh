mkdir -p gpurun_out; timeout 400 python -m pytest tests/test_gpu_objmask.py tests/test_zz_no_alloc.py -x -q > gpurun_out/om_tests.log 2>&1; tail -2 gpurun_out/om_tests.log; timeout 120 python tools/time_objmask.py 1024 10 2>&1 | tail -1; timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/om_l.csv python tools/time_objmask.py 1024 2 > /dev/null 2>&1; python - <<PY
import csv,collections
d=collections.defaultdict(list)
rows=[r for r in csv.reader(open("gpurun_out/om_l.csv")) if len(r)>5]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
for r in rows[1:]:
    try: d[r[ki][:40]].append(float(r[vi].replace(",","")))
    except: pass
for k,v in d.items(): print(k, len(v), sum(v)/len(v)/1000, "us")
PY
