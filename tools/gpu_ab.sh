#!/bin/bash
# A/B of one environment switch on the headline benchmark, back to back on one box (each value twice, interleaved):
#   gpurun --timeout 600 -- 'bash tools/gpu_ab.sh CXN_CONV_CPAD 0 1'            # 1 GPU
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_ab.sh CXN_OVERLAP_WGRAD 0 1 8'
# Prints ms/step, images/s and the end-to-end figure per run; JSON lines land in gpurun_out/ab_<var>_<value>_<i>.json.
set -u
VAR=$1; A=$2; B=$3; N=${4:-1}
OUT=gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
i=0
for v in $A $B $A $B; do
  i=$((i + 1))
  f=$OUT/ab_${VAR}_${v}_$i.json
  if [ "$N" -gt 1 ]; then
    env $VAR=$v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
        --master-port 29555 bench.py --gpus $N --steps 40 --warmup 5 --no-verify > $f 2> ${f%.json}.err
  else
    env $VAR=$v timeout 300 python bench.py --steps 40 --warmup 5 > $f 2> ${f%.json}.err
  fi
  python - "$f" "$VAR=$v" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
    print("%-28s %.4f ms/step  %8.0f img/s  e2e %8.0f" % (sys.argv[2], d['ms_per_step'], d['value'], d['e2e']['value']))
except Exception as e:
    print(sys.argv[2], "failed:", e)
PY
done
