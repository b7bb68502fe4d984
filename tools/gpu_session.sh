#!/bin/bash
# One GPU-box session in the order that yields evidence even if the call is cut short:
#   tests -> bench line -> launch list -> one full ncu capture of the top kernel.
# Usage (from the repository root, through the remote runner):
#   gpurun --timeout 1500 -- 'bash tools/gpu_session.sh'            # 1 GPU
#   gpurun --gpus 2 --timeout 1500 -- 'bash tools/gpu_session.sh 2'  # also the 2-GPU tests and bench
# Everything lands in gpurun_out/ (scratch); copy what should be judged into profiles/.
# Opt-in experiments that have not yet run on hardware (each under its own timeout, failures do not stop the rest):
#   CXN_TEST_COLLECTIVE=1   torch.distributed exchange on GPUs (tests/test_gpu_multi.py, needs 2 GPUs)
#   CXN_TEST_GEMM_CLUSTER=1 activation-multicast GEMM clusters (tests/test_gpu_gemm.py)
set -u
N=${1:-1}
OUT=gpurun_out
mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }

echo "== pytest -m gpu"
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -3 $OUT/pytest_gpu.log

echo "== opt-in experiments"
timeout 300 python -m pytest tests/test_gpu_conv.py -q -k cluster > $OUT/pytest_cluster.log 2>&1
rc_cluster=$?
echo "cluster conv exit $rc_cluster"; tail -3 $OUT/pytest_cluster.log
if [ $rc_cluster -eq 0 ]; then
  CXN_CONV_CLUSTER=2 timeout 300 python bench.py --steps 20 --warmup 5 > $OUT/bench_cluster.json 2>/dev/null && \
    python -c "import json;d=json.load(open('$OUT/bench_cluster.json'));print('CXN_CONV_CLUSTER=2:', round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms')"
  EXTRA=1 WHICH=fprop timeout 600 python bench/conv_sweep.py > $OUT/conv_sweep_extra.txt 2>&1; tail -12 $OUT/conv_sweep_extra.txt
fi
timeout 300 python -m pytest tests/test_gpu_graph.py -q -k u8 > $OUT/pytest_u8.log 2>&1
echo "u8 input batches exit $?"; tail -3 $OUT/pytest_u8.log
timeout 300 python bench.py --input u8 --steps 20 --warmup 5 > $OUT/bench_u8.json 2>/dev/null && python -c "import json;d=json.load(open('$OUT/bench_u8.json'));print('u8 input: e2e', round(d['e2e']['value']), 'img/s, device', round(d['value']))"
timeout 300 python -m pytest tests/test_gpu_conv.py -q -k halo > $OUT/pytest_halo.log 2>&1
rc_halo=$?
echo "halo-tile conv exit $rc_halo"; tail -3 $OUT/pytest_halo.log
if [ $rc_halo -eq 0 ]; then
  CXN_CONV_HALO=1 timeout 300 python bench.py --steps 20 --warmup 5 > $OUT/bench_halo.json 2>/dev/null && \
    python -c "import json;d=json.load(open('$OUT/bench_halo.json'));print('CXN_CONV_HALO=1:', round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms')"
fi
CXN_TEST_GEMM_CLUSTER=1 timeout 300 python -m pytest tests/test_gpu_gemm.py -q -k clusters > $OUT/pytest_gemm_cluster.log 2>&1
rc_gc=$?
echo "cluster GEMM exit $rc_gc"; tail -3 $OUT/pytest_gemm_cluster.log
if [ $rc_gc -eq 0 ]; then
  for cl in 2 4; do
    CXN_GEMM_CLUSTER=$cl timeout 300 python bench.py --steps 20 --warmup 5 > $OUT/bench_gemm_cl$cl.json 2>/dev/null && \
      python -c "import json;d=json.load(open('$OUT/bench_gemm_cl$cl.json'));print('CXN_GEMM_CLUSTER=$cl:', round(d['value']), 'img/s', round(d['ms_per_step'],4), 'ms')"
  done
fi
timeout 120 python bench/desc_probe.py > $OUT/desc_probe.txt 2>&1
echo "descriptor probe exit $?"; tail -18 $OUT/desc_probe.txt
if [ "$N" -ge 2 ]; then
  CXN_TEST_COLLECTIVE=1 timeout 300 python -m pytest tests/test_gpu_multi.py -q -k collective > $OUT/pytest_collective.log 2>&1
  echo "collective exchange exit $?"; tail -3 $OUT/pytest_collective.log
fi

echo "== bench"
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $OUT/bench_1.json 2> $OUT/bench_1.err && cat $OUT/bench_1.json
if [ "$N" -ge 2 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 \
      bench.py --gpus $N --steps 30 --warmup 5 > $OUT/bench_$N.json 2> $OUT/bench_$N.err && cat $OUT/bench_$N.json
fi

echo "== small-batch A/B (what one GPU computes in the 8-GPU run)"
timeout 300 python bench.py --batch 32 --steps 30 --warmup 5 > $OUT/bench_b32.json 2>/dev/null && python -c "import json;d=json.load(open('$OUT/bench_b32.json'));print('batch 32:', round(d['ms_per_step'],4), 'ms')"
CXN_CONV_FILL=1 timeout 300 python bench.py --batch 32 --steps 30 --warmup 5 > $OUT/bench_b32_fill.json 2>/dev/null && python -c "import json;d=json.load(open('$OUT/bench_b32_fill.json'));print('batch 32, CXN_CONV_FILL=1:', round(d['ms_per_step'],4), 'ms')"

echo "== launch list (eager steps, so that every kernel is visible)"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 2 --warmup 3 --extra "cuda_graph = 0" > $OUT/ncu_launches.log 2>&1
python tools/launch_profile.py $OUT/launches.csv > $OUT/step_launches.txt 2>&1 && head -30 $OUT/step_launches.txt

echo "== full capture of the convolution kernels of one step"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_ -c 14 -o $OUT/conv_full \
    python bench.py --steps 1 --warmup 3 --extra "cuda_graph = 0" > $OUT/ncu_full.log 2>&1
ls -la $OUT | tail -12
