set -u
N=${1:-8}
OUT=$PWD/gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
    print(sys.argv[2], round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms e2e', round(d.get('e2e',{}).get('value',0)), d.get('symmetric_memory'), d.get('replica_max_diff'), d.get('verify'))
except Exception as e:
    print(sys.argv[2], 'failed', e)
PY
}
timeout 600 $TR 29560 bench/exchange_sweep.py --iters 10 > $OUT/xsweep_${N}_nvls.json 2> $OUT/xsweep_${N}_nvls.err; echo "sweep nvls exit $?"
if [ "$N" = "8" ]; then
CXN_SYMM=ipc timeout 600 $TR 29561 bench/exchange_sweep.py --iters 10 > $OUT/xsweep_${N}_ipc.json 2> $OUT/xsweep_${N}_ipc.err; echo "sweep ipc exit $?"
CXN_XCHG_MC_U=1 timeout 600 $TR 29566 bench/exchange_sweep.py --iters 10 > $OUT/xsweep_${N}_nvls_u1.json 2> $OUT/xsweep_${N}_nvls_u1.err; echo "sweep nvls u1 exit $?"
else
timeout 600 $TR 29562 bench.py --gpus $N --steps 40 --warmup 5 > $OUT/final_$N.json 2> $OUT/final_$N.err; show $OUT/final_$N.json "final N=$N"
timeout 300 $TR 29563 bench/config_bench.py --model bowl --gpus $N > $OUT/cfg_bowl_$N.json 2> $OUT/cfg_bowl_$N.err; show $OUT/cfg_bowl_$N.json "bowl N=$N"
fi
if [ "$N" = "2" ]; then
timeout 300 $TR 29564 bench/config_bench.py --model lenet --gpus 2 > $OUT/cfg_lenet_2.json 2> $OUT/cfg_lenet_2.err; show $OUT/cfg_lenet_2.json "lenet N=2"
timeout 600 $TR 29565 bench.py --gpus 2 --steps 20 --warmup 5 --impl library > $OUT/library_2.json 2> $OUT/library_2.err; show $OUT/library_2.json "library N=2"
fi
python tools/show_sweep.py $OUT/xsweep_${N}_nvls.json $OUT/xsweep_${N}_ipc.json $OUT/xsweep_${N}_nvls_u1.json
