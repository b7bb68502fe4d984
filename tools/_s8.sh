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
    print("N=$N $name", round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms  e2e', round(d['e2e']['value']), round(d['e2e']['ms_per_step'],4), 'launches/step', d['gpu_launches']/d['steps'], d.get('replica_max_diff'))
except Exception as e:
    print("$name failed", e)
PY
}
if [ "$N" = "4" ]; then
run r1_gather build/r1
CXN_XCHG_MIN_BUCKET_MB=4 run tp_3buckets . --no-verify
run tp_1bucket . --no-verify
run tp_1bucket_u8 . --no-verify --input u8
else
run tp_1bucket_u8 . --input u8
fi
