#!/usr/bin/env python
"""Where does the CTA-pair GEMM differ from torch?  Prints, per shape, the largest error and its location."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cxxnet_b200 import ops
L = ops.lib()
r8 = lambda x: ops.round_up(x, 8)


def run(M, N, K, pair, bias=False, f32=False):
    torch.manual_seed(1)
    A = torch.zeros(M, r8(K), dtype=torch.bfloat16, device="cuda"); A[:, :K] = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    B = torch.zeros(N, r8(K), dtype=torch.bfloat16, device="cuda"); B[:, :K] = torch.randn(N, K, device="cuda").to(torch.bfloat16)
    D = torch.zeros(M, r8(N), dtype=torch.float32 if f32 else torch.bfloat16, device="cuda")
    bt = torch.randn(N, device="cuda") if bias else None
    L.cxn_gemm_pair(2 if pair else 0)
    ops.gemm(A.data_ptr(), r8(K), 0, 0, B.data_ptr(), r8(K), 0, 0, D.data_ptr(), r8(N), 0, bt.data_ptr() if bias else None, 0,
             M, N, K, flags=0 if f32 else ops.GEMM_OUT_BF16, bn=256 if pair else 0)
    torch.cuda.synchronize()
    L.cxn_gemm_pair(0)
    ref = A[:, :K].float() @ B[:, :K].float().t()
    if bias:
        ref = ref + bt
    err = (D[:, :N].float() - ref).abs()
    i = int(err.argmax())
    bad = (err > 0.05 * max(1.0, float(ref.abs().max()))).nonzero()
    rows = sorted(set(int(r) for r in bad[:, 0].tolist()))
    cols = sorted(set(int(c) for c in bad[:, 1].tolist()))
    print("f32=%d " % f32, end="")
    print("M=%d N=%d K=%d pair=%d bias=%d: max err %.4g at (%d,%d) ref max %.3g; bad rows %s cols %s" % (
        M, N, K, pair, bias, float(err.max()), i // N, i % N, float(ref.abs().max()),
        (rows[:3], rows[-3:], len(rows)) if rows else "-", (cols[:3], cols[-3:], len(cols)) if cols else "-"))


for shape in [(512, 512, 256), (300, 512, 256), (512, 333, 256), (512, 512, 200), (300, 333, 200), (256, 512, 64), (384, 512, 256),
              (512, 384, 256), (512, 512, 192), (512, 512, 320)]:
    run(*shape, pair=1)
run(300, 333, 200, pair=1, bias=True)
run(300, 333, 200, pair=0, bias=True)

for pair in (1, 0):
    for shape in [(512, 512, 256), (384, 1000, 512), (512, 512, 512), (256, 256, 64), (4096, 9216, 256)]:
        run(*shape, pair=pair, f32=True)
