set -u
OUT=gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -k "shard" > $OUT/pytest_shard.log 2>&1; echo "shard tests exit $?"; tail -30 $OUT/pytest_shard.log | grep -E "rel err|passed|failed|Error" 
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 30 --warmup 5 > $OUT/bench_2.json 2> $OUT/bench_2.err; echo "bench2 exit $?"; cat $OUT/bench_2.json; grep verify $OUT/bench_2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29556 bench.py --gpus 2 --steps 30 --warmup 5 --extra "fullc_gather = 1" > $OUT/bench_2_gather.json 2> $OUT/bench_2_gather.err; echo "bench2 gather exit $?"; cat $OUT/bench_2_gather.json; grep verify $OUT/bench_2_gather.err
