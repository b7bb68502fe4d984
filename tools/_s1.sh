set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_layers.py -x -q -m gpu > $OUT/pytest_fuse.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_fuse.log
CXN_GEMM_WIDE=1 timeout 200 python bench/wgrad_epilogue.py 2>&1 | tee $OUT/wgrad_epilogue_pipelined.txt
for cfg in "1 0" "1 1" "1 0" "1 1"; do
set -- $cfg
CXN_GEMM_WIDE=$1 CXN_FUSE_UPDATE=$2 timeout 300 python bench.py --steps 40 --warmup 5 > $OUT/f_$1$2.json 2> $OUT/f_$1$2.err
python - <<PY
import json
d=json.loads([l for l in open("$OUT/f_$1$2.json") if l.startswith('{')][-1])
print("wide=$1 fuse=$2 ms/step %.4f img/s %.0f e2e %.0f" % (d['ms_per_step'], d['value'], d['e2e']['value']))
PY
done
