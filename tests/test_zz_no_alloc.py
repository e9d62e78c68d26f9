"""The allocation-free cross-entropy entry point, driven through the C ABI.

Some users of libb200ddp cannot allocate device memory: an operator plugged into
a serving process whose memory was handed out at start-up, a latency-critical
loop that must never hit the device-wide stall of cudaMalloc, a C host that owns
nothing but the tensors it was given.  ``b200_cross_entropy_static`` serves them
by keeping its scratch in a fixed array inside the library.  This file checks,
with plain ``ctypes`` calls and no helper from the Python package in between,
that it returns what the regular entry point returns — twice in a row, because
the array is shared and keeps the previous call's contents — and that it refuses
problems larger than its stated limit.

The file name sorts last on purpose: a run that stops at the first failure has
then already executed every other test.
"""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "b200ddp", "_native", "libb200ddp.so")
MAX_BLOCKS = 1024  # B200_CE_STATIC_MAX_BLOCKS


@pytest.fixture(scope="module")
def L():
    lib = C.CDLL(LIB)
    lib.b200_last_error.restype = C.c_char_p
    lib.b200_ce_scratch_bytes.restype = C.c_int64
    lib.b200_ce_scratch_bytes.argtypes = [C.c_int64]
    common = [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_float,
              C.c_void_p]
    lib.b200_cross_entropy.argtypes = common + [C.c_void_p, C.c_void_p]
    lib.b200_cross_entropy.restype = C.c_int
    lib.b200_cross_entropy_static.argtypes = common + [C.c_void_p]
    lib.b200_cross_entropy_static.restype = C.c_int
    return lib


def _both(L, logits, tgt):
    """Runs the regular and the allocation-free entry point on the same inputs."""
    dev = logits.device
    rows, classes = logits.shape
    is_bf16 = int(logits.dtype == torch.bfloat16)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    scratch = torch.zeros(int(L.b200_ce_scratch_bytes(rows)), dtype=torch.uint8, device=dev)
    want_dl, got_dl = torch.empty_like(logits), torch.empty_like(logits)
    want, got = torch.zeros(2, device=dev), torch.zeros(2, device=dev)
    rc = L.b200_cross_entropy(logits.data_ptr(), is_bf16, logits.stride(0), tgt.data_ptr(), rows,
                              classes, want_dl.data_ptr(), 1.0 / rows, want.data_ptr(),
                              scratch.data_ptr(), stream)
    assert rc == 0, L.b200_last_error()
    rc = L.b200_cross_entropy_static(logits.data_ptr(), is_bf16, logits.stride(0), tgt.data_ptr(),
                                     rows, classes, got_dl.data_ptr(), 1.0 / rows, got.data_ptr(),
                                     stream)
    assert rc == 0, L.b200_last_error()
    torch.cuda.synchronize()
    return (want, want_dl), (got, got_dl)


@pytest.mark.parametrize("rows,classes,dtype", [(64, 10, torch.float32), (1000, 777, torch.float32),
                                                (16, 50257, torch.bfloat16)])
def test_matches_regular_entry_point_twice(L, rows, classes, dtype):
    dev = torch.device("cuda")
    for call in range(2):  # second call finds the array as the first one left it
        torch.manual_seed(10 + call)
        logits = (torch.randn(rows, classes, device=dev) * 3).to(dtype)
        tgt = torch.randint(0, classes, (rows,), device=dev)
        (want, want_dl), (got, got_dl) = _both(L, logits, tgt)
        # tolerances of tests/test_gpu_kernels.py::test_cross_entropy_kernel
        assert abs(got[0].item() - want[0].item()) < 2e-4 * max(1.0, abs(want[0].item()))
        assert got[1].item() == want[1].item()
        err = (got_dl.float() - want_dl.float()).norm() / want_dl.float().norm()
        assert err.item() < (1e-5 if dtype == torch.float32 else 1e-2)
        # and both agree with PyTorch
        ref = torch.nn.functional.cross_entropy(logits.float(), tgt)
        assert abs(got[0].item() - ref.item()) < 2e-4 * max(1.0, abs(ref.item()))


def test_rejects_sizes_beyond_the_built_in_scratch(L):
    dev = torch.device("cuda")
    out = torch.zeros(2, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    for rows, classes in ((8 * MAX_BLOCKS + 1, 10), (MAX_BLOCKS + 1, 4096)):
        logits = torch.zeros(rows, classes, device=dev)
        tgt = torch.zeros(rows, dtype=torch.int64, device=dev)
        rc = L.b200_cross_entropy_static(logits.data_ptr(), 0, classes, tgt.data_ptr(), rows, classes,
                                         None, 1.0, out.data_ptr(), stream)
        assert rc == -1
        assert b"built-in scratch" in L.b200_last_error()
    # the largest accepted size still works
    rows, classes = 8 * MAX_BLOCKS, 10
    logits = torch.randn(rows, classes, device=dev)
    tgt = torch.randint(0, classes, (rows,), device=dev)
    rc = L.b200_cross_entropy_static(logits.data_ptr(), 0, classes, tgt.data_ptr(), rows, classes, None,
                                     1.0, out.data_ptr(), stream)
    assert rc == 0, L.b200_last_error()
    torch.cuda.synchronize()
    ref = torch.nn.functional.cross_entropy(logits, tgt)
    assert abs(out[0].item() - ref.item()) < 2e-4 * max(1.0, abs(ref.item()))
