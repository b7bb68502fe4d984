set -u
N=${1:-4}
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
run() { # name extra...
  name=$1; shift
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --warmup 5 --steps 40 "$@" > $OUT/bench_${N}_$name.json 2> $OUT/bench_${N}_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$OUT/bench_${N}_$name.json") if l.startswith('{')][-1])
    print("N=$N $name", round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms', 'e2e', round(d.get('e2e',{}).get('value',0)), d.get('symmetric_memory'), d.get('replica_max_diff'), d.get('verify'))
except Exception as e:
    print("$name failed", e)
PY
}
run final
