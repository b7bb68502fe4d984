set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_layers.py -x -q -m gpu > $OUT/pytest_conv.log 2>&1; echo "pytest exit $?"; tail -15 $OUT/pytest_conv.log
for m in; do
CXN_CONV_CPAD=$m timeout 300 python bench.py --steps 40 --warmup 5 > $OUT/q_$m.json 2> $OUT/q_$m.err
python - <<PY
import json
d=json.loads([l for l in open("$OUT/q_$m.json") if l.startswith('{')][-1])
print("cpad=$m ms/step %.4f img/s %.0f e2e %.0f" % (d['ms_per_step'], d['value'], d['e2e']['value']))
PY
done
true
true
