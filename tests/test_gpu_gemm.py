"""tcgen05 GEMM (csrc/kernels/gemm_sm100.cu) against a plain PyTorch fp32 reference."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(rows, cols, ld, batch=1, scale=1.0):
    t = torch.zeros(batch, rows, ld, dtype=torch.bfloat16, device="cuda")
    t[:, :, :cols] = (torch.randn(batch, rows, cols, device="cuda") * scale).to(torch.bfloat16)
    return t


def _run(M, N, K, a_mn, b_mn, batch=1, splits=1, relu=False, bias=False, out_bf16=True, atomic=False, beta1=False,
         bn=0):
    from cxxnet_b200 import ops
    r8 = lambda x: ops.round_up(x, 8)
    torch.manual_seed(M * 7 + N * 3 + K)
    if a_mn:
        A = _mk(K, M, r8(M), batch)          # stored [K][M]
        Af = A[:, :, :M].float().transpose(1, 2)
        lda, sa = r8(M), K * r8(M)
    else:
        A = _mk(M, K, r8(K), batch)
        Af = A[:, :, :K].float()
        lda, sa = r8(K), M * r8(K)
    if b_mn:
        B = _mk(K, N, r8(N), batch)          # stored [K][N]
        Bf = B[:, :, :N].float().transpose(1, 2)
        ldb, sb = r8(N), K * r8(N)
    else:
        B = _mk(N, K, r8(K), batch)
        Bf = B[:, :, :K].float()
        ldb, sb = r8(K), N * r8(K)
    ref = torch.bmm(Af, Bf.transpose(1, 2))
    bias_t = None
    if bias:
        bias_t = torch.randn(batch, N, device="cuda")
        ref = ref + bias_t[:, None, :]
    if relu:
        ref = torch.relu(ref)
    ldd = r8(N)
    flags = 0
    if relu: flags |= ops.GEMM_RELU
    if out_bf16:
        D = torch.zeros(batch, M, ldd, dtype=torch.bfloat16, device="cuda")
        flags |= ops.GEMM_OUT_BF16
    else:
        D = torch.randn(batch, M, ldd, device="cuda") if (atomic or beta1) else torch.zeros(batch, M, ldd, device="cuda")
        if atomic or beta1:
            ref = ref + D[:, :, :N]
        if atomic: flags |= ops.GEMM_ATOMIC
        if beta1: flags |= ops.GEMM_BETA1
    ops.gemm(A.data_ptr(), lda, sa, int(a_mn), B.data_ptr(), ldb, sb, int(b_mn), D.data_ptr(), ldd, M * ldd,
             bias_t.data_ptr() if bias else None, N, M, N, K, batch=batch, splits=splits, flags=flags, bn=bn)
    torch.cuda.synchronize()
    got = D[:, :, :N].float()
    tol = 2e-2 * max(1.0, float(ref.abs().max())) if out_bf16 else 2e-3 * max(1.0, float(ref.abs().max()))
    err = float((got - ref).abs().max())
    assert err <= tol, "M=%d N=%d K=%d a_mn=%d b_mn=%d batch=%d splits=%d: err %g > tol %g" % (
        M, N, K, a_mn, b_mn, batch, splits, err, tol)
    if D.shape[2] > N:
        assert float(D[:, :, N:].abs().max()) == 0.0 or (atomic or beta1), "wrote outside the valid columns"


@pytest.mark.parametrize("bn", [64, 128, 256])
def test_gemm_tn_basic(bn):
    _run(256, 256, 512, False, False, bn=bn)


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (100, 10, 784), (300, 200, 72), (256, 4096, 1024), (1000, 96, 363),
                                   (129, 257, 65)])
def test_gemm_tn_shapes(M, N, K):
    _run(M, N, K, False, False, bias=True, relu=True)
    _run(M, N, K, False, False, out_bf16=False)


@pytest.mark.parametrize("M,N,K", [(256, 512, 128), (100, 784, 100), (333, 72, 96)])
def test_gemm_k_major_a_mn_major_b(M, N, K):
    _run(M, N, K, False, True)


@pytest.mark.parametrize("M,N,K", [(128, 128, 256), (100, 784, 100), (96, 363, 3000), (10, 100, 100)])
def test_gemm_mn_major_both(M, N, K):
    _run(M, N, K, True, True, out_bf16=False)
    _run(M, N, K, True, True, out_bf16=False, beta1=True)


def test_gemm_mn_major_a_k_major_b():
    _run(192, 160, 320, True, False)


def test_gemm_batched_groups():
    _run(500, 128, 1200, False, False, batch=2, bias=True, relu=True)
    _run(128, 1200, 500, True, True, batch=2, out_bf16=False)


@pytest.mark.parametrize("splits", [2, 5, 16])
def test_gemm_split_k_atomic(splits):
    _run(96, 363, 8192, True, True, splits=splits, out_bf16=False, atomic=True)
    _run(128, 1728, 4096, True, True, batch=2, splits=splits, out_bf16=False, atomic=True)


def test_gemm_many_tiles_persistent():
    # more tiles than SMs: exercises the persistent loop and both TMEM accumulator stages
    _run(128 * 20, 128 * 12, 256, False, False, bn=128)
    _run(128 * 40, 256, 192, False, False, bn=64, bias=True)


@pytest.mark.skipif(__import__("os").environ.get("CXN_TEST_GEMM_CLUSTER", "0") != "1",
                    reason="activation-multicast GEMM clusters: written without access to a GPU, opt-in with "
                           "CXN_TEST_GEMM_CLUSTER=1 until the route has run on hardware")
@pytest.mark.parametrize("cl", [2, 4])
def test_gemm_clusters_share_the_activation_tile(cl):
    """cxn_gemm_cluster(cl): skinny K-major-A problems run as clusters of cl column tiles that multicast the
    activation tile; fully connected forward (K-major weights, bias + ReLU, bf16) and backward-data (MN-major
    weights) shapes, column-tile counts that do not divide by the cluster size, M below a tile"""
    from cxxnet_b200 import ops
    L = ops.lib()
    L.cxn_gemm_cluster(cl)
    try:
        _run(256, 4096, 1024, 0, 0, relu=True, bias=True)         # fc forward: 2 row tiles x 64 column tiles
        _run(256, 4096, 1024, 0, 1)                               # fc backward-data: weights MN-major
        _run(256, 64 * 7, 512, 0, 0, bias=True)                   # 7 column tiles: the last cluster is partly empty
        _run(100, 640, 320, 0, 0, relu=True)                      # M below one tile
        _run(256, 2048, 768, 0, 0, out_bf16=False)                # fp32 output
        _run(384, 1000, 512, 0, 1, bias=True)                     # N not a multiple of the tile
    finally:
        L.cxn_gemm_cluster(0)


def test_memory_is_local_guard():
    """cxn_memory_is_local: device tensors of the current GPU qualify for the multicast routes, pinned host memory
    and null pointers do not (the multicast tensor maps must cover this GPU's own memory only)"""
    import ctypes
    from cxxnet_b200 import ops
    L = ops.lib()
    L.cxn_memory_is_local.argtypes = [ctypes.c_void_p, ctypes.c_ulonglong]
    L.cxn_memory_is_local.restype = ctypes.c_int
    t = torch.zeros(1 << 20, dtype=torch.bfloat16, device="cuda")
    assert L.cxn_memory_is_local(t.data_ptr(), t.numel() * 2) == 1
    h = torch.zeros(1 << 16).pin_memory()
    assert L.cxn_memory_is_local(h.data_ptr(), h.numel() * 4) == 0
    assert L.cxn_memory_is_local(None, 16) == 0


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1), (1, 0)])
def test_gemm_cta_pairs(a_mn, b_mn):
    """cxn_gemm_pair(2): 256 x 256 tiles computed by CTA pairs with cta_group::2 MMAs (each CTA stages its rows of A
    and half of the B tile) in every operand layout: more tiles than pairs (persistent loop, both accumulator stages),
    M / N / K tails, bias + ReLU, fp32 output, batches, split-K with fp32 reductions"""
    from cxxnet_b200 import ops
    L = ops.lib()
    L.cxn_gemm_pair.argtypes = [__import__("ctypes").c_int]
    assert L.cxn_gemm_pair(2) == 0
    try:
        _run(512, 512, 256, a_mn, b_mn)
        _run(256 * 14, 256 * 9, 192, a_mn, b_mn, relu=True, bias=True)      # 126 tiles on 74 pairs
        _run(300, 333, 200, a_mn, b_mn, bias=True)                          # tails in M, N and K
        _run(384, 1000, 512, a_mn, b_mn, out_bf16=False)
        _run(256, 512, 320, a_mn, b_mn, batch=3, out_bf16=False)
        _run(256, 768, 4096, a_mn, b_mn, splits=4, out_bf16=False, atomic=True)
        _run(129, 130, 64, a_mn, b_mn)                                       # barely two CTAs' worth of rows
    finally:
        L.cxn_gemm_pair(-1)


@pytest.mark.parametrize("kind", [0, 1, 2])
@pytest.mark.parametrize("M,N,K", [(256, 512, 64), (4096, 1152, 256), (200, 100, 48), (130, 72, 40)])
def test_gemm_wgrad_with_fused_updater(kind, M, N, K):
    """cxn_gemm_wgrad_update: the weight-gradient GEMM whose epilogue applies sgd / nag / adam to the fp32 master and
    the momentum and rewrites the bf16 copy - against g = dy^T x in fp64 followed by the closed-form updater (full
    tiles, several tiles per CTA, and ragged rows / columns that take the element-wise path)"""
    import ctypes
    from cxxnet_b200 import ops
    L = ops.lib()
    gen = torch.Generator().manual_seed(M + 7 * N + K + kind)
    ldy, ldx = ops.round_up(M, 8), ops.round_up(N, 8)
    dy = torch.zeros(K, ldy, dtype=torch.bfloat16, device="cuda")
    x = torch.zeros(K, ldx, dtype=torch.bfloat16, device="cuda")
    dy[:, :M] = (torch.randn(K, M, generator=gen) * 0.5).cuda().to(torch.bfloat16)
    x[:, :N] = torch.randn(K, N, generator=gen).cuda().to(torch.bfloat16)
    w0 = torch.randn(M, N, generator=gen).cuda()
    m0 = (torch.randn(M, N, generator=gen) * 0.1).cuda()
    v0 = (torch.rand(M, N, generator=gen) * 0.1).cuda()
    lr, wd, mom, clip, adam_lr_t, b1, b2 = 0.05, 0.01, 0.9, (2.0 if kind == 0 else 0.0), 0.002, 0.1, 0.001
    hyper = torch.tensor([lr, wd, mom, clip, adam_lr_t, b1, b2], dtype=torch.float32, device="cuda")
    sld = ops.round_up(N, 8)
    for with_shadow in (True, False):
        w, m, v = w0.clone(), m0.clone(), v0.clone()
        shadow = torch.zeros(M, sld, dtype=torch.bfloat16, device="cuda")
        rc = L.cxn_gemm_wgrad_update(dy.data_ptr(), ldy, x.data_ptr(), ldx, M, N, K, w.data_ptr(), m.data_ptr(),
                                     v.data_ptr() if kind == 2 else None, shadow.data_ptr() if with_shadow else None,
                                     N, sld, hyper.data_ptr(), kind, ops.stream())
        assert rc == 0
        torch.cuda.synchronize()
        g = (dy[:, :M].double().t() @ x[:, :N].double())
        wr, mr, vr = w0.double(), m0.double(), v0.double()
        if kind == 0:
            g = g.clamp(-clip, clip)
            mr = mom * mr - lr * (g + wd * wr)
            wr = wr + mr
        elif kind == 1:
            old = mr
            mr = mom * old - lr * (g + wd * wr)
            wr = wr + (1 + mom) * mr - mom * old
        else:
            g = g - wd * wr
            mr = mr + b1 * (g - mr)
            vr = vr + b2 * (g * g - vr)
            wr = wr - adam_lr_t * (mr / (vr.sqrt() + 1e-8))
        assert torch.allclose(w.double(), wr, atol=2e-5, rtol=2e-5), (kind, float((w.double() - wr).abs().max()))
        assert torch.allclose(m.double(), mr, atol=2e-5, rtol=2e-5)
        if kind == 2:
            assert torch.allclose(v.double(), vr, atol=2e-5, rtol=2e-5)
        else:
            assert torch.equal(v, v0)
        if with_shadow:
            assert torch.equal(shadow[:, :N], w.to(torch.bfloat16))
            assert float(shadow[:, N:].float().abs().max()) == 0.0 if sld > N else True
        else:
            assert float(shadow.float().abs().max()) == 0.0
    # arguments the fused path does not take
    assert L.cxn_gemm_wgrad_update(dy.data_ptr(), ldy, x.data_ptr(), ldx, M, N, K, w.data_ptr(), m.data_ptr(), None,
                                   None, N, sld, hyper.data_ptr(), 2, ops.stream()) == -7      # adam without m2
    assert L.cxn_gemm_wgrad_update(dy.data_ptr(), ldy, x.data_ptr(), ldx, M, N, K, w.data_ptr(), m.data_ptr(), None,
                                   None, N, sld, None, 0, ops.stream()) == -7                  # no hyper-parameters
