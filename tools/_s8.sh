set -u
N=${1:-8}
OUT=$PWD/gpurun_out; mkdir -p $OUT
run() { # name dir extra...
  name=$1; dir=$2; shift 2
  ( cd $dir && timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 40 --warmup 5 "$@" > $OUT/bench_${N}_$name.json 2> $OUT/bench_${N}_$name.err ); echo "$name exit $?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$OUT/bench_${N}_$name.json") if l.startswith('{')][-1])
    print("$name", round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms  e2e', round(d['e2e']['value']), 'host', round(d['host_enqueue_ms_per_step'],3), d.get('replica_max_diff'))
except Exception as e:
    print("$name failed", e)
PY
}
run r1_gather build/r1
run cur_gather . --no-verify --extra "fullc_gather = 1"
run cur_tp . --no-verify
run r1_gather_again build/r1
