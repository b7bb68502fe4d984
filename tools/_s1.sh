set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_layers.py tests/test_gpu_kernels.py -x -q -m gpu > $OUT/pytest_fuse.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_fuse.log
for m in 0 1 0 1; do
CXN_FUSE_UPDATE=$m timeout 300 python bench.py --steps 40 --warmup 5 > $OUT/f_$m.json 2> $OUT/f_$m.err
python - <<PY
import json
d=json.loads([l for l in open("$OUT/f_$m.json") if l.startswith('{')][-1])
print("fuse=$m ms/step %.4f img/s %.0f e2e %.0f launches %s" % (d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches']))
PY
done
for m in 0 1; do
CXN_FUSE_UPDATE=$m CXN_STEP_TRACE=1 timeout 300 python bench.py --steps 30 --warmup 5 > $OUT/ft_$m.json 2> $OUT/ft_$m.err
echo "trace fuse=$m"; grep -E "^(b1[5-9]|b2[0-9]|end|whole|metric)" $OUT/ft_$m.err | tr '\n' ' '; echo
done
