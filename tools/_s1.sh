set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_layers.py -x -q -m gpu > $OUT/pytest_conv.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_conv.log
for cfg in "256 1" "32 0" "32 1" "32 0" "32 1"; do
set -- $cfg
CXN_CONV_CPAD=$2 timeout 300 python bench.py --steps 40 --warmup 5 --batch $1 > $OUT/q_$1_$2.json 2> $OUT/q_$1_$2.err
python - <<PY
import json
d=json.loads([l for l in open("$OUT/q_$1_$2.json") if l.startswith('{')][-1])
print("batch $1 cpad=$2 ms/step %.4f img/s %.0f e2e %.0f" % (d['ms_per_step'], d['value'], d['e2e']['value']))
PY
done
