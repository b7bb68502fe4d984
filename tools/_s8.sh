set -u
N=${1:-8}
OUT=$PWD/gpurun_out; mkdir -p $OUT
run() { # name dir extra...
  name=$1; dir=$2; shift 2
  ( cd $dir && timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --warmup 5 "$@" > $OUT/bench_${N}_$name.json 2> $OUT/bench_${N}_$name.err ); echo "$name exit $?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$OUT/bench_${N}_$name.json") if l.startswith('{')][-1])
    print("N=$N $name", round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms', 'e2e', d.get('e2e',{}).get('value'), d.get('replica_max_diff'), d.get('verify'))
except Exception as e:
    print("$name failed", e)
PY
}
CXN_STEP_TRACE=1 run trace . --steps 10 --no-verify
grep -A70 "^rank 0" $OUT/bench_${N}_trace.err | head -75
run final . --steps 40
run library . --steps 20 --impl library
