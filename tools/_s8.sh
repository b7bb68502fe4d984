set -u
N=${1:-8}
OUT=$PWD/gpurun_out; mkdir -p $OUT
run() { # name extra...
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --warmup 5 --steps 40 "$@" > $OUT/bench_${N}_$name.json 2> $OUT/bench_${N}_$name.err; echo "$name exit $?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$OUT/bench_${N}_$name.json") if l.startswith('{')][-1])
    print("N=$N $name", round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms', 'e2e', round(d.get('e2e',{}).get('value',0)), d.get('symmetric_memory'), d.get('replica_max_diff'))
except Exception as e:
    print("$name failed", e)
PY
}
CXN_OVERLAP_WGRAD=0 run seq --no-verify
run overlap --no-verify
