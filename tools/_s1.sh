set -u
OUT=$PWD/gpurun_out; mkdir -p $OUT
python build.py > $OUT/build.log 2>&1 || { tail -20 $OUT/build.log; exit 1; }
timeout 600 python -m pytest tests/test_gpu_exchange_single.py tests/test_zz_no_alloc.py -q 2>&1 | tail -3
echo "== new barriers"; timeout 300 python bench/barrier_latency.py 2>&1 | tail -5
echo "== round-1 tree (old barriers; fused exchange only)"; (cd build/r1 && timeout 300 python - <<'PY'
import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import ctypes, torch
from cxxnet_b200 import ops
import test_zz_no_alloc as Z
lib = ctypes.CDLL(Z.LIB)
Pt, I, LL, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint
lib.cxn_fused_rs_update_ag.argtypes = [Pt, I, I, LL, LL, LL, LL, LL, LL, Pt, Pt, LL, LL, I, U, I, Pt, I, Pt]
blk = Z.Block(lib, [65536])
counter = torch.zeros(4, dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def run(n, s0):
    for i in range(n):
        lib.cxn_fused_rs_update_ag(*blk.args(s0 + i + 1, 2), counter.data_ptr(), 16, st)
run(10, 0); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); run(200, 10); e1.record(); torch.cuda.synchronize()
print("fused exchange, 64 K parameters (round-1 kernel): %.2f us per launch" % (e0.elapsed_time(e1) / 200 * 1e3))
PY
)
