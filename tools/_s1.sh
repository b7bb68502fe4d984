set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu -k parameter_server > $OUT/pytest_ps.log 2>&1; echo "pytest exit $?"; tail -25 $OUT/pytest_ps.log
