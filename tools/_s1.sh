set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_layers.py -q -k "strided or stride or shared" > $OUT/pytest_strided.log 2>&1; echo "exit $?"; grep -E "passed|failed|^FAILED|Error: |rror" $OUT/pytest_strided.log | head -20
