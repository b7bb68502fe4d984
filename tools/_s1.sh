set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_layers.py -x -q -m gpu > $OUT/pytest_pool.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_pool.log
for m in 0 1 0 1; do
CXN_POOL_QUAD=$m timeout 300 python bench.py --steps 40 --warmup 5 > $OUT/q_$m.json 2> $OUT/q_$m.err
python - <<PY
import json
d=json.loads([l for l in open("$OUT/q_$m.json") if l.startswith('{')][-1])
print("quad=$m ms/step %.4f img/s %.0f e2e %.0f" % (d['ms_per_step'], d['value'], d['e2e']['value']))
PY
done
for m in 0 1; do
CXN_POOL_QUAD=$m CXN_STEP_TRACE=1 timeout 300 python bench.py --steps 30 --warmup 5 > $OUT/qt_$m.json 2> $OUT/qt_$m.err
echo "trace quad=$m"; grep -E "^(b[0-9]+:(max_pooling|lrn)|f[0-9]+:(max_pooling|lrn))" $OUT/qt_$m.err | tr '\n' ' ' | tr -s ' '; echo
done
