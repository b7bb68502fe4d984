set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 200 python bench/conv1_ring.py 2>&1 | tee $OUT/conv1_ring.txt
