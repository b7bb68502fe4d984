set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
b() { name=$1; shift; timeout 300 python bench.py "$@" > $OUT/b1_$name.json 2>$OUT/b1_$name.err; python -c "
import json
try:
    d=json.loads([l for l in open('$OUT/b1_$name.json') if l.startswith('{')][-1]); print('$name', round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms  e2e', round(d['e2e']['value']))
except Exception as e: print('$name failed', e)
"; }
CXN_CONV_WSPLIT=1 b b32_old --steps 30 --warmup 5 --batch 32
b b32_new --steps 30 --warmup 5 --batch 32
CXN_CONV_WSPLIT=1 b b64_old --steps 30 --warmup 5 --batch 64
b b64_new --steps 30 --warmup 5 --batch 64
CXN_CONV_WSPLIT=1 b b128_old --steps 30 --warmup 5 --batch 128
b b128_new --steps 30 --warmup 5 --batch 128
CXN_CONV_WSPLIT=1 b b256_old --steps 20 --warmup 5
b b256_new --steps 20 --warmup 5
CXN_STEP_TRACE=1 timeout 300 python bench.py --steps 10 --warmup 5 --batch 32 > $OUT/b1_trace32.json 2> $OUT/b1_trace32.err; grep -A60 "^rank 0" $OUT/b1_trace32.err | grep -E "conv|whole|end"
