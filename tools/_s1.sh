set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 900 python -m pytest tests/test_gpu_conv.py -q > $OUT/pytest_conv.log 2>&1; echo "conv tests exit $?"; grep -E "passed|failed|^FAILED|Error: |rror" $OUT/pytest_conv.log | head -12
b() { name=$1; shift; timeout 300 python bench.py "$@" > $OUT/b1_$name.json 2>$OUT/b1_$name.err; python -c "
import json
try:
    d=json.loads([l for l in open('$OUT/b1_$name.json') if l.startswith('{')][-1]); print('$name', round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms')
except Exception as e: print('$name failed', e)
"; }
CXN_CONV_FSPLIT=0 b b32_nosplit --steps 30 --warmup 5 --batch 32
b b32_split --steps 30 --warmup 5 --batch 32
CXN_CONV_FSPLIT=0 b b64_nosplit --steps 30 --warmup 5 --batch 64
b b64_split --steps 30 --warmup 5 --batch 64
b b256 --steps 20 --warmup 5
CXN_STEP_TRACE=1 timeout 300 python bench.py --steps 10 --warmup 5 --batch 32 > $OUT/b1_trace32.json 2> $OUT/b1_trace32.err; grep -A60 "^rank 0" $OUT/b1_trace32.err | grep -E "conv|whole"
