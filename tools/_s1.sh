set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
echo "== gpu tests"
timeout 1500 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; grep -E "passed|failed|^FAILED|Error: " $OUT/pytest_gpu.log | head -20
echo "== bench 1 GPU"
timeout 300 python bench.py --steps 20 --warmup 5 > $OUT/b1.json 2>$OUT/b1.err; cat $OUT/b1.json | cut -c1-300
timeout 300 python bench.py --steps 20 --warmup 5 --impl library > $OUT/b1_library.json 2>$OUT/b1_library.err; cat $OUT/b1_library.json | cut -c1-300
echo "== config benches"
timeout 300 python bench/config_bench.py --model lenet > $OUT/cfg_lenet_1.json 2>$OUT/cfg_lenet_1.err; cat $OUT/cfg_lenet_1.json | cut -c1-400
timeout 300 python bench/config_bench.py --model bowl > $OUT/cfg_bowl_1.json 2>$OUT/cfg_bowl_1.err; cat $OUT/cfg_bowl_1.json | cut -c1-400
echo "== memcheck (small shapes)"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_conv.py tests/test_gpu_exchange_single.py -q -x -k "basic or cta_pairs or odd_shapes or single_rank or 3x3_wide" > $OUT/sanitizer_memcheck.log 2>&1; echo "memcheck exit $?"; grep -E "ERROR SUMMARY|passed|failed|Invalid|error" $OUT/sanitizer_memcheck.log | head -10
