set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 600 python -m pytest tests/test_gpu_layers.py tests/test_gpu_graph.py -x -q -m gpu > $OUT/pytest_s2d.log 2>&1; echo "pytest exit $?"; tail -15 $OUT/pytest_s2d.log
for m in 0 1 0 1; do
CXN_CONV_S2D=$m timeout 300 python bench.py --steps 40 --warmup 5 > $OUT/q_$m.json 2> $OUT/q_$m.err
python - <<PY
import json
d=json.loads([l for l in open("$OUT/q_$m.json") if l.startswith('{')][-1])
print("s2d=$m ms/step %.4f img/s %.0f e2e %.0f" % (d['ms_per_step'], d['value'], d['e2e']['value']))
PY
done
CXN_STEP_TRACE=1 timeout 300 python bench.py --steps 30 --warmup 5 > $OUT/qt_4.json 2> $OUT/qt_4.err
grep -E "^([fb][0-9]+:conv|whole)" $OUT/qt_4.err | tr '\n' ' ' | tr -s ' '; echo
