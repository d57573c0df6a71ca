"""Out-of-bounds writes, checked by hand: compute-sanitizer is closed on this GPU pool, so the new kernels are called
through the raw C ABI with every OUTPUT placed between sentinel-filled guard bands; after the call the bands must be
untouched and the payload fully written (no sentinel left inside, for dense outputs)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda") if torch.cuda.is_available() else None
GUARD = 4096          # elements on each side


class Guarded:
    def __init__(self, numel, dtype, fill):
        self.buf = torch.full((numel + 2 * GUARD,), fill, dtype=dtype, device=cuda)
        self.fill, self.numel = fill, numel
        self.view = self.buf[GUARD: GUARD + numel]

    def ptr(self):
        return ctypes.c_void_p(self.view.data_ptr())

    def check(self, name, dense=True):
        lo, hi = self.buf[:GUARD], self.buf[GUARD + self.numel:]
        if self.fill != self.fill:      # NaN sentinel
            assert torch.isnan(lo).all() and torch.isnan(hi).all(), f"{name}: wrote outside its output"
            if dense:
                assert not torch.isnan(self.view).any(), f"{name}: left part of its output unwritten"
        else:
            assert (lo == self.fill).all() and (hi == self.fill).all(), f"{name}: wrote outside its output"


@pytest.mark.parametrize("B,H,S,d,with_bias", [(3, 4, 197, 64, False), (5, 6, 64, 10, True), (3, 2, 49, 32, True), (2, 2, 130, 32, False)])
def test_attention_kernels_stay_inside_their_outputs(B, H, S, d, with_bias):
    from pytorch_distributedtraining_b200.ops._lib import require_library
    from pytorch_distributedtraining_b200.ops.fused import DT, stream_ptr
    lib = require_library()
    nan = float("nan")
    qkv = torch.randn(B, S, 3 * H * d, device=cuda).bfloat16()
    dout = torch.randn(B, S, H * d, device=cuda).bfloat16()
    bias = torch.randn(H, S, S, device=cuda) if with_bias else None
    out, lse = Guarded(B * S * H * d, torch.bfloat16, nan), Guarded(B * H * S, torch.float32, nan)
    p = lambda t: ctypes.c_void_p(0 if t is None else t.data_ptr())
    rc = lib.b200dp_attention_fwd(p(qkv), out.ptr(), lse.ptr(), p(bias), None, 0, B, H, S, d, ctypes.c_float(d ** -0.5),
                                  DT[torch.bfloat16], stream_ptr(cuda))
    assert rc == 0
    torch.cuda.synchronize()
    out.check("attention_fwd out"); lse.check("attention_fwd lse")
    dqkv = Guarded(qkv.numel(), torch.bfloat16, nan)
    dbias = Guarded(H * S * S, torch.float32, 0.0) if with_bias else None
    rc = lib.b200dp_attention_bwd(p(qkv), out.ptr(), p(dout), lse.ptr(), dqkv.ptr(), p(bias), None, 0,
                                  dbias.ptr() if dbias else None, B, H, S, d, ctypes.c_float(d ** -0.5), DT[torch.bfloat16],
                                  stream_ptr(cuda))
    assert rc == 0
    torch.cuda.synchronize()
    dqkv.check("attention_bwd dqkv")
    if dbias is not None:
        dbias.check("attention_bwd dbias", dense=False)
        assert torch.isfinite(dbias.view).all()


@pytest.mark.parametrize("N,H,W,Cin,Cout,R,stride,pad", [(2, 224, 224, 3, 64, 7, 2, 3), (3, 33, 29, 3, 24, 3, 1, 1),
                                                         (2, 40, 36, 4, 64, 5, 1, 2), (1, 9, 300, 1, 8, 3, 2, 0)])
def test_first_layer_kernels_stay_inside_their_outputs(N, H, W, Cin, Cout, R, stride, pad):
    from pytorch_distributedtraining_b200.ops._lib import require_library
    from pytorch_distributedtraining_b200.ops.fused import DT, stream_ptr
    lib = require_library()
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    x = torch.randn(N, H, W, Cin, device=cuda).half()
    w = torch.randn(Cout, R, R, Cin, device=cuda).half()
    y = Guarded(N * P * Q * Cout, torch.float16, float("nan"))
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.b200dp_conv2d_smallc_fwd(p(x), p(w), y.ptr(), None, N, H, W, Cin, Cout, R, R, stride, pad, DT[torch.float16],
                                      stream_ptr(cuda))
    assert rc == 0
    torch.cuda.synchronize()
    y.check("conv2d_smallc_fwd y")
    dy = torch.randn(N, P, Q, Cout, device=cuda).half()
    dw = Guarded(Cout * R * R * Cin, torch.float32, 0.0)
    rc = lib.b200dp_conv2d_smallc_wgrad(p(x), p(dy), dw.ptr(), N, H, W, Cin, Cout, R, R, stride, pad, DT[torch.float16],
                                        stream_ptr(cuda))
    assert rc == 0
    torch.cuda.synchronize()
    dw.check("conv2d_smallc_wgrad dw", dense=False)
    assert torch.isfinite(dw.view).all() and dw.view.abs().sum() > 0
