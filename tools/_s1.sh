set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
for m in 0 1 0 1; do
CXN_OVERLAP_WGRAD=$m CXN_STEP_TRACE=1 timeout 300 python bench.py --steps 30 --warmup 5 --batch 32 > $OUT/t_$m.json 2> $OUT/t_$m.err
python - <<PY
import json,re
d=json.loads([l for l in open("$OUT/t_$m.json") if l.startswith('{')][-1])
rows=[l.split() for l in open("$OUT/t_$m.err") if re.match(r'^[fb]\d+:|^metric|^end|^whole',l)]
t={r[0]:float(r[1]) for r in rows}
fwd=sum(v for k,v in t.items() if k.startswith('f')); bconv=sum(v for k,v in t.items() if k.startswith('b') and int(k[1:].split(':')[0])<=14); bfc=sum(v for k,v in t.items() if k.startswith('b') and int(k[1:].split(':')[0])>14)
print("overlap=$m ms/step %.4f | trace: fwd %.1f bwd-fc %.1f bwd-conv %.1f metric %.1f end %.1f whole %.1f" % (d['ms_per_step'], fwd, bfc, bconv, t.get('metric',0), t.get('end',0), t.get('whole',0)))
PY
done
