"""The allocation-free sum-of-squares entry point, driven through the C ABI.

Some users of the kernel library cannot give it a workspace: a host callback or a serving
loop where asking the driver for memory would stall the GPU, a process whose device
memory has been handed out in full by an outer allocator, or a C program with no caching
allocator at all.  ``b200dp_sumsq_f32_noalloc`` is for them — its scratch is a fixed array
inside the library — and this test calls it exactly as such a caller would: raw pointers
through ctypes, no workspace, twice in a row (the second call sees the first call's
leftovers in the shared array), and checks it against the regular entry point that takes
caller-provided scratch.  The file sorts last so a ``-x`` run has already been through
every other test when it gets here; it must pass on its own in a fresh process too.
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

_SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "pytorch_distributedtraining_b200", "ops", "libb200dp.so")
MAX_ELEMS = 1 << 28


def _load():
    lib = ctypes.CDLL(_SO)
    lib.b200dp_last_error.restype = ctypes.c_char_p
    lib.b200dp_sumsq_scratch_floats.restype = ctypes.c_int64
    lib.b200dp_sumsq_scratch_floats.argtypes = [ctypes.c_int64]
    lib.b200dp_sumsq_f32_noalloc.restype = ctypes.c_int
    lib.b200dp_sumsq_f32_noalloc.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]
    lib.b200dp_sumsq_f32.restype = ctypes.c_int
    lib.b200dp_sumsq_f32.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_void_p, ctypes.c_void_p]
    return lib


def _existing(lib, x, stream):
    """The regular entry point: the caller brings scratch and a zeroed ticket."""
    n = x.numel()
    scratch = torch.empty(int(lib.b200dp_sumsq_scratch_floats(n)), dtype=torch.float32, device=x.device)
    ticket = torch.zeros(1, dtype=torch.int32, device=x.device)
    out = torch.full((1,), -1.0, dtype=torch.float32, device=x.device)
    rc = lib.b200dp_sumsq_f32(x.data_ptr(), n, scratch.data_ptr(), ticket.data_ptr(), out.data_ptr(), stream)
    assert rc == 0, lib.b200dp_last_error()
    return out


@pytest.mark.parametrize("n", [1, 4099, 16384, 3 * 16384 + 5, (1 << 22) + 1])
def test_noalloc_matches_regular_entry_point_twice_in_a_row(n):
    lib = _load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    g = torch.Generator(device="cpu").manual_seed(n)
    xs = [torch.randn(n, generator=g).to(dev), (torch.randn(n, generator=g) * 3.0).to(dev)]
    for x in xs:  # two calls back to back on one stream, sharing the built-in scratch
        out = torch.full((1,), -1.0, dtype=torch.float32, device=dev)
        rc = lib.b200dp_sumsq_f32_noalloc(x.data_ptr(), n, out.data_ptr(), stream)
        assert rc == 0, lib.b200dp_last_error()
        want = _existing(lib, x, stream)
        torch.cuda.synchronize()
        torch.testing.assert_close(out, want, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(out[0], (x.double() ** 2).sum().float(), rtol=1e-5, atol=1e-6)


def test_noalloc_rejects_oversized_and_handles_empty():
    lib = _load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    x = torch.ones(8, device=dev)
    out = torch.full((1,), 7.0, dtype=torch.float32, device=dev)
    rc = lib.b200dp_sumsq_f32_noalloc(x.data_ptr(), MAX_ELEMS + 1, out.data_ptr(), stream)
    assert rc != 0
    msg = lib.b200dp_last_error().decode()
    assert "exceeds" in msg and str(MAX_ELEMS) in msg
    torch.cuda.synchronize()
    assert out.item() == 7.0, "a refused call must not launch anything"
    rc = lib.b200dp_sumsq_f32_noalloc(x.data_ptr(), 0, out.data_ptr(), stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert out.item() == 0.0
    # and the array still works afterwards
    rc = lib.b200dp_sumsq_f32_noalloc(x.data_ptr(), 8, out.data_ptr(), stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert out.item() == 8.0
