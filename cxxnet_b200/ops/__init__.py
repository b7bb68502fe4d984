"""ctypes bindings to the hand-written sm_100a kernel library (``csrc/kernels``).

The library is built in-tree by ``build.py`` / ``__graft_entry__.build()`` into
``cxxnet_b200/lib/libcxxnet_b200.so`` and exposes a plain C ABI declared in
``include/cxxnet_b200.h``: every entry point takes raw device pointers plus a
``cudaStream_t`` and returns 0 on success.  There is no eager/PyTorch fallback here: if a
CUDA device is present and the library is missing, ``lib()`` raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_float, c_int, c_longlong, c_ulonglong, c_void_p
from typing import Optional

import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libcxxnet_b200.so")
_lib = None

# gemm flags (csrc/kernels/gemm_sm100.cu)
GEMM_RELU = 1
GEMM_OUT_BF16 = 2
GEMM_ATOMIC = 4
GEMM_BETA1 = 8

ACT_RELU, ACT_SIGMOID, ACT_TANH, ACT_XELU = 0, 1, 2, 3
POOL_MAX, POOL_SUM, POOL_AVG = 0, 1, 2
UPD_SGD, UPD_NAG, UPD_ADAM = 0, 1, 2


class KernelLibraryError(RuntimeError):
    pass


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def _sig(fn, argtypes):
    fn.argtypes = argtypes
    fn.restype = c_int


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise KernelLibraryError(
            "cxxnet_b200 kernel library not built: %s missing (run `python build.py`)" % _LIB_PATH)
    L = ctypes.CDLL(_LIB_PATH)
    P, I, LL, F, ULL = c_void_p, c_int, c_longlong, c_float, c_ulonglong
    _sig(L.cxn_gemm_bf16, [P, LL, LL, I, P, LL, LL, I, P, LL, LL, P, LL, I, I, I, I, I, I, I, P])
    _sig(L.cxn_gemm_wgrad_update, [P, LL, P, LL, I, I, I, P, P, P, P, LL, LL, P, I, P])
    if hasattr(L, "cxn_gemm_cluster"):
        _sig(L.cxn_gemm_cluster, [I])
    _sig(L.cxn_nchw_f32_to_nhwc_bf16, [P, P, I, I, I, I, P])
    _sig(L.cxn_nhwc_bf16_to_nchw_f32, [P, P, I, I, I, I, P])
    _sig(L.cxn_bf16_hwc_chw, [P, P, I, I, I, LL, LL, I, I, I, P])
    _sig(L.cxn_copy2d, [P, LL, P, LL, LL, I, I, I, P])
    _sig(L.cxn_act_fwd, [P, P, LL, I, LL, I, F, P])
    _sig(L.cxn_act_bwd, [P, P, P, LL, I, LL, I, F, P])
    _sig(L.cxn_dropout, [P, P, LL, I, LL, F, ULL, P, P])
    _sig(L.cxn_colsum_bf16, [P, LL, I, LL, P, I, P])
    _sig(L.cxn_fill_f32, [P, LL, F, P])
    _sig(L.cxn_im2col_nhwc, [P, P, I, I, I, I, I, I, I, I, I, I, I, I, I, LL, P])
    _sig(L.cxn_col2im_nhwc, [P, P, I, I, I, I, I, I, I, I, I, I, I, I, I, LL, P])
    _sig(L.cxn_conv_fprop, [P, I, I, I, I, I, I, I, P, LL, I, I, I, I, I, I, I, I, P, LL, P, I, P])
    _sig(L.cxn_conv_wgrad, [P, I, I, I, I, I, I, I, P, LL, I, I, I, I, I, I, I, I, P, LL, I, P])
    _sig(L.cxn_conv_wgrad_masked, [P, I, I, I, I, I, I, I, P, LL, I, I, I, I, I, I, I, I, P, LL, I, I, I, P])
    _sig(L.cxn_nchw_f32_to_nhwc4_pitched, [P, P, I, I, I, I, I, P])
    _sig(L.cxn_nchw_bf16_to_nhwc4_pitched, [P, P, I, I, I, I, I, P])
    if hasattr(L, "cxn_nchw_u8_to_nhwc4_pitched"):
        _sig(L.cxn_nchw_u8_to_nhwc4_pitched, [P, P, I, I, I, I, I, F, F, F, F, F, P])
    _sig(L.cxn_conv_fprop_masked, [P, I, I, I, I, I, I, I, P, LL, I, I, I, I, I, I, I, I, P, LL, P, P])
    _sig(L.cxn_conv_fprop_view, [P, I, I, I, I, I, I, P, LL, I, I, I, I, I, I, I, I, P, LL, P, I, LL, LL, LL, P])
    _sig(L.cxn_conv_wgrad_view, [P, I, I, I, I, I, I, P, LL, I, I, I, I, I, I, I, I, P, LL, I, I, I, LL, LL, LL, P])
    _sig(L.cxn_conv_flip_weights, [P, P, I, I, I, I, I, I, I, P])
    _sig(L.cxn_conv_flip_weights_pitch, [P, P, I, I, I, I, I, I, I, I, P])
    _sig(L.cxn_nchw_to_s2d4, [P, I, P, I, I, I, I, F, F, F, F, F, P])
    _sig(L.cxn_mask_cols_f32, [P, LL, I, LL, P, P])
    _sig(L.cxn_conv_dgrad_strided, [P, I, I, I, LL, I, I, P, LL, I, I, I, I, I, I, I, I, P, LL, P, P, P])
    _sig(L.cxn_conv_dgrad_strided_ok, [I, I, I, I, I])
    _sig(L.cxn_conv_wgrad_grid_cap, [I])
    _sig(L.cxn_conv_view_group_stride, [I])
    L.cxn_conv_flat_launches.argtypes = []
    L.cxn_conv_flat_launches.restype = c_longlong
    L.cxn_conv_dgrad_strided_scratch.argtypes = [I, I, I, I, I, I]
    L.cxn_conv_dgrad_strided_scratch.restype = c_longlong
    _sig(L.cxn_conv_tune, [I, I])
    _sig(L.cxn_bias_add, [P, P, LL, I, LL, P, P])
    _sig(L.cxn_batchnorm_fwd, [P, P, P, LL, I, LL, LL, LL, P, P, F, P, P, P, P])
    _sig(L.cxn_batchnorm_bwd, [P, P, P, LL, I, LL, LL, LL, P, P, P, P, P, I, P])
    _sig(L.cxn_prelu_fwd, [P, P, LL, I, LL, LL, P, F, ULL, P, P])
    _sig(L.cxn_prelu_bwd, [P, P, P, LL, I, LL, LL, P, P, F, ULL, P, P])
    _sig(L.cxn_insanity_fwd, [P, P, LL, I, LL, LL, F, F, I, ULL, P, P])
    _sig(L.cxn_insanity_bwd, [P, P, P, LL, I, LL, LL, F, F, I, ULL, P, P])
    _sig(L.cxn_insanity_displace, [P, P, I, I, I, I, LL, F, ULL, P, P])
    _sig(L.cxn_conv_set_prof, [P])
    _sig(L.cxn_pool_fwd, [P, P, I, I, I, I, I, I, I, I, I, I, I, P])
    _sig(L.cxn_pool_bwd, [P, P, P, P, I, I, I, I, I, I, I, I, I, I, I, P])
    _sig(L.cxn_pool_bwd_bias, [P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, I, P])
    _sig(L.cxn_lrn_fwd, [P, P, LL, I, I, F, F, F, P])
    _sig(L.cxn_lrn_bwd, [P, P, P, LL, I, I, F, F, F, P])
    _sig(L.cxn_metric_error, [P, LL, P, LL, I, I, P, P])
    _sig(L.cxn_softmax_rows, [P, I, I, LL, P])
    _sig(L.cxn_sigmoid_rows, [P, LL, I, LL, P])
    _sig(L.cxn_loss_grad, [P, LL, P, LL, P, LL, LL, I, F, I, P])
    _sig(L.cxn_update, [P, P, P, P, P, P, P, LL, LL, I, P])
    _sig(L.cxn_refresh_shadow, [P, P, P, P, LL, P])
    _sig(L.cxn_counter_add, [P, ULL, P])
    _sig(L.cxn_timestamp, [P, I, P])
    L.cxn_launch_count.restype = c_ulonglong
    L.cxn_launch_count_add.restype = c_ulonglong
    L.cxn_launch_count_add.argtypes = [c_ulonglong]
    L.cxn_segment_bytes.restype = c_int
    L.cxn_update_chunk.restype = c_int
    _lib = L
    return L


def launch_count() -> int:
    """kernel launches issued by the native library so far in this process."""
    return int(lib().cxn_launch_count())


_device_index = None


def set_device_index(index) -> None:
    """device whose current stream `stream()` reports (one process drives one GPU)."""
    global _device_index
    _device_index = index


def stream() -> int:
    """raw handle of the current CUDA stream.  torch.cuda.current_stream() costs ~13 us of Python per
    call - with ~70 launches per step that was 45 % of the host time of a step."""
    global _device_index
    if _device_index is None:
        _device_index = torch.cuda.current_device()
    return torch._C._cuda_getCurrentRawStream(_device_index)


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise KernelLibraryError("%s failed with code %d" % (what, rc))


def round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class DevMat:
    """A 2-D bf16 device matrix ``[rows][ld]`` with ``cols`` valid entries per row.

    4-D nodes are stored NHWC: rows = b*h*w, cols = ld = c.  Matrix nodes ``(b,1,1,n)``:
    rows = b, cols = n, ld = round_up(n, 8) so TMA strides stay 16-byte aligned.
    """
    __slots__ = ("t", "rows", "cols", "ld")

    def __init__(self, rows: int, cols: int, ld: Optional[int] = None, device=None, dtype=torch.bfloat16,
                 tensor: Optional[torch.Tensor] = None):
        self.rows, self.cols = rows, cols
        self.ld = cols if ld is None else ld
        if tensor is None:
            tensor = torch.zeros((max(rows, 1), self.ld), dtype=dtype, device=device)
        self.t = tensor

    @property
    def ptr(self) -> int:
        return self.t.data_ptr()

    def head(self, rows: int) -> "DevMat":
        """first ``rows`` rows (same storage)."""
        return DevMat(rows, self.cols, self.ld, tensor=self.t[:rows])

    def valid(self) -> torch.Tensor:
        return self.t[: self.rows, : self.cols]


def gemm(a_ptr, lda, stride_a, a_mn, b_ptr, ldb, stride_b, b_mn, d_ptr, ldd, stride_d, bias_ptr, stride_bias,
         M, N, K, batch=1, splits=1, flags=0, bn=0):
    check(lib().cxn_gemm_bf16(a_ptr, lda, stride_a, a_mn, b_ptr, ldb, stride_b, b_mn, d_ptr, ldd, stride_d,
                              bias_ptr, stride_bias, M, N, K, batch, splits, flags, bn, stream()), "cxn_gemm_bf16")
