"""The allocation-free column-sum entry point, driven through the raw C ABI.

``tony_colsum_bf16_static`` exists for host programs that cannot hand the library any scratch
memory and cannot let it allocate either — e.g. a job whose framework has already claimed the
whole GPU for its own memory pool, so a ``cudaMalloc`` made by a dlopen'ed helper fails.  Its
scratch is a fixed device array inside the library, shared by every call.  This test loads
the library with ctypes (no wrapper layer), checks the result against the ordinary entry point
``tony_colsum_bf16`` on two back-to-back calls of different shapes (so the second call runs on
top of whatever the first left in the shared scratch), then checks that a matrix wider than the
compiled-in limit is refused with a clear error.  It lives in its own file, named to be
collected last, so a failure here cannot hide the result of any other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_COLS = 8192          # TONY_COLSUM_STATIC_MAX_COLS in include/tony_b200_kernels.h
ERR_TOO_LARGE = 4        # TONY_ERR_TOO_LARGE


def _abi():
    from tony_b200.ops._lib import LIB_PATH

    lib = ctypes.CDLL(str(LIB_PATH))
    vp = ctypes.c_void_p
    lib.tony_colsum_bf16_static.argtypes = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_longlong, vp,
                                            ctypes.c_int, vp]
    lib.tony_colsum_bf16_static.restype = ctypes.c_int
    lib.tony_colsum_bf16.argtypes = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_longlong, vp, vp,
                                     ctypes.c_size_t, ctypes.c_int, vp]
    lib.tony_colsum_bf16.restype = ctypes.c_int
    lib.tony_colsum_scratch_bytes.argtypes = [ctypes.c_int]
    lib.tony_colsum_scratch_bytes.restype = ctypes.c_size_t
    lib.tony_last_error.restype = ctypes.c_char_p
    return lib


def _close(got, want, rel=1e-2):
    err = (got.float() - want.float()).abs().max().item()
    tol = rel * want.float().abs().max().item() + 1e-3
    assert err <= tol, f"max err {err} > tol {tol}"


def test_static_scratch_colsum_matches_regular_entry_point_twice_in_a_row():
    lib = _abi()
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.manual_seed(0)
    # two calls in a row; the second is narrower and shorter, then accumulates into `out`
    for rows, cols, accumulate in ((4096, 4096, 0), (777, 1032, 1)):
        x = (torch.randn(rows, cols, device=dev) * 0.5).to(torch.bfloat16)
        init = (torch.randn(cols, device=dev)).to(torch.bfloat16)
        got, want = init.clone(), init.clone()
        nbytes = lib.tony_colsum_scratch_bytes(cols)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        scratch[:4096].zero_()      # the regular entry point's contract: its counters start at zero
        rc = lib.tony_colsum_bf16(x.data_ptr(), rows, cols, x.stride(0), want.data_ptr(), scratch.data_ptr(),
                                  nbytes, accumulate, stream)
        assert rc == 0, lib.tony_last_error()
        rc = lib.tony_colsum_bf16_static(x.data_ptr(), rows, cols, x.stride(0), got.data_ptr(), accumulate, stream)
        assert rc == 0, lib.tony_last_error()
        torch.cuda.synchronize()
        _close(got, want)
        ref = x.float().sum(0) + (init.float() if accumulate else 0)
        _close(got, ref)


def test_static_scratch_colsum_rejects_oversized_input():
    lib = _abi()
    dev = torch.device("cuda", torch.cuda.current_device())
    cols = MAX_COLS + 8
    x = torch.zeros(64, cols, dtype=torch.bfloat16, device=dev)
    out = torch.zeros(cols, dtype=torch.bfloat16, device=dev)
    rc = lib.tony_colsum_bf16_static(x.data_ptr(), 64, cols, x.stride(0), out.data_ptr(), 0,
                                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == ERR_TOO_LARGE
    assert b"TONY_COLSUM_STATIC_MAX_COLS" in lib.tony_last_error()
    # exactly at the limit is accepted
    x = torch.ones(64, MAX_COLS, dtype=torch.bfloat16, device=dev)
    out = torch.zeros(MAX_COLS, dtype=torch.bfloat16, device=dev)
    rc = lib.tony_colsum_bf16_static(x.data_ptr(), 64, MAX_COLS, x.stride(0), out.data_ptr(), 0,
                                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.tony_last_error()
    torch.cuda.synchronize()
    assert out.float().min().item() == 64 and out.float().max().item() == 64


def test_static_scratch_colsum_allocates_nothing_and_handles_strided_rows():
    """The reason this entry point exists: after its first call on a device (which resolves the built-in
    array's address) further calls must not change the device's free memory at all.  Also a row stride
    wider than the row (a column slice of a bigger matrix), and the argument checks a caller without any
    error channel but the return code relies on."""
    lib = _abi()
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.manual_seed(1)
    big = (torch.randn(1000, 2048, device=dev) * 0.5).to(torch.bfloat16)
    x = big[:, 256:256 + 1024]                      # ld = 2048 > cols = 1024, 16-byte aligned start
    out = torch.empty(1024, dtype=torch.bfloat16, device=dev)
    assert lib.tony_colsum_bf16_static(x.data_ptr(), 1000, 1024, x.stride(0), out.data_ptr(), 0, stream) == 0
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]
    for _ in range(3):
        rc = lib.tony_colsum_bf16_static(x.data_ptr(), 1000, 1024, x.stride(0), out.data_ptr(), 0, stream)
        assert rc == 0, lib.tony_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free_before, "the allocation-free entry point changed free memory"
    _close(out, x.float().sum(0))
    # refused, with TONY_ERR_INVALID (1): a column count / stride that is not a multiple of 8, a
    # misaligned base pointer, null pointers
    ERR_INVALID = 1
    assert lib.tony_colsum_bf16_static(x.data_ptr(), 1000, 1020, x.stride(0), out.data_ptr(), 0, stream) == ERR_INVALID
    assert lib.tony_colsum_bf16_static(x.data_ptr(), 1000, 1024, 2044, out.data_ptr(), 0, stream) == ERR_INVALID
    assert lib.tony_colsum_bf16_static(x.data_ptr() + 2, 1000, 1024, x.stride(0), out.data_ptr(), 0, stream) == ERR_INVALID
    assert lib.tony_colsum_bf16_static(None, 1000, 1024, x.stride(0), out.data_ptr(), 0, stream) == ERR_INVALID
    assert lib.tony_colsum_bf16_static(x.data_ptr(), 1000, 1024, x.stride(0), None, 0, stream) == ERR_INVALID
