set -u
OUT=gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
python - <<'PY'
import torch
print("devices", torch.cuda.device_count())
PY
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > $OUT/pytest_multi_nvls.log 2>&1; echo "multi tests (CXN_SYMM=auto) exit $?"; tail -5 $OUT/pytest_multi_nvls.log | cut -c1-300
CXN_SYMM=ipc timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -k "shard or lenet" > $OUT/pytest_multi_ipc.log 2>&1; echo "multi tests (CXN_SYMM=ipc) exit $?"; tail -3 $OUT/pytest_multi_ipc.log | cut -c1-300
for mode in auto ipc; do
CXN_SYMM=$mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 30 --warmup 5 --no-verify > $OUT/bench_2_$mode.json 2> $OUT/bench_2_$mode.err; echo "bench2 $mode exit $?"; python -c "
import json; d=json.loads([l for l in open('$OUT/bench_2_$mode.json') if l.startswith('{')][-1]); print('$mode', round(d['value']), round(d['ms_per_step'],4), d.get('symmetric_memory'), d.get('replica_max_diff'))"; grep -iE "error|Traceback" $OUT/bench_2_$mode.err | head -5
CXN_SYMM=$mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29556 bench.py --gpus 2 --steps 30 --warmup 5 --no-verify --extra "fullc_shard = 1" > $OUT/bench_2_tp_$mode.json 2> $OUT/bench_2_tp_$mode.err; echo "bench2 tp $mode exit $?"; python -c "
import json; d=json.loads([l for l in open('$OUT/bench_2_tp_$mode.json') if l.startswith('{')][-1]); print('tp $mode', round(d['value']), round(d['ms_per_step'],4), d.get('symmetric_memory'), d.get('replica_max_diff'))"; grep -iE "error|Traceback" $OUT/bench_2_tp_$mode.err | head -5
done
