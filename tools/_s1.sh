set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 1500 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; grep -E "passed|failed|^FAILED|Error: " $OUT/pytest_gpu.log | head -20
timeout 300 python bench.py --steps 20 --warmup 5 > $OUT/b1.json 2>$OUT/b1.err; python -c "
import json; d=json.loads([l for l in open('$OUT/b1.json') if l.startswith('{')][-1]); print('cur b256', round(d['value']), round(d['ms_per_step'],4), 'ms launches', d['gpu_launches'], 'err', d['train_error_running'])"
timeout 600 python bench/gemm_sweep.py > $OUT/gemm_sweep.txt 2>&1; python - <<PY
import json
for l in open("$OUT/gemm_sweep.txt"):
    if l.startswith('{'):
        d=json.loads(l); print(d['shape'], 'auto', d['auto_us'], 'bn256', d['bn256_us'], 'pair', d.get('pair_us'), 'cublas', d['cublas_us'])
PY
