"""``dgb_colsum_bf16_noalloc`` -- the column sum for callers that may not allocate device memory.

A host application that gave the whole GPU to its own pool, or a loop that must never hit a device-wide stall from
``cudaMalloc``/``cudaFree``, cannot supply the scratch buffer ``dgb_colsum_bf16`` asks for.  For them the library keeps a
fixed workspace inside its own image; this test drives that entry point through the plain C ABI (ctypes, no Python
wrapper) and checks it against the regular entry point.  Two calls are made back to back with different shapes, so the
second call starts with the first call's leftovers in the shared workspace -- the result must not depend on them.

The file name sorts last on purpose: with ``-x`` every other test has already run when this one is reached.
"""
import ctypes

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(120)]


def _lib():
    from dgraph_b200.ops import _native

    assert _native.available(), "native library must be built"
    lib = ctypes.CDLL(str(_native.lib_path()))  # C ABI only
    lib.dgb_colsum_scratch_bytes.restype = ctypes.c_int64
    lib.dgb_colsum_bf16.restype = ctypes.c_int
    lib.dgb_colsum_bf16_noalloc.restype = ctypes.c_int
    return lib


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _regular(lib, x, with_sq):
    M, N = x.shape
    nbytes = lib.dgb_colsum_scratch_bytes(ctypes.c_int64(M), ctypes.c_int(N), ctypes.c_int(with_sq))
    scratch = torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device="cuda")
    out = torch.empty(N * (2 if with_sq else 1), dtype=torch.float32, device="cuda")
    rc = lib.dgb_colsum_bf16(ctypes.c_void_p(x.data_ptr()), ctypes.c_int64(x.stride(0)), ctypes.c_int64(M), ctypes.c_int(N),
                             ctypes.c_int(with_sq), ctypes.c_void_p(scratch.data_ptr()), ctypes.c_void_p(out.data_ptr()),
                             _stream())
    assert rc == 0
    return out


def _noalloc(lib, x, with_sq, out):
    M, N = x.shape
    return lib.dgb_colsum_bf16_noalloc(ctypes.c_void_p(x.data_ptr()), ctypes.c_int64(x.stride(0)), ctypes.c_int64(M),
                                       ctypes.c_int(N), ctypes.c_int(with_sq), ctypes.c_void_p(out.data_ptr()), _stream())


def test_noalloc_colsum_matches_regular_entry_point_twice_in_a_row():
    lib = _lib()
    torch.manual_seed(0)
    # call 1: wide and tall, fills most of the built-in workspace; call 2: narrower and shorter, sees call 1's leftovers
    cases = [(700_000, 512, 1), (3_001, 104, 0)]
    xs = [(torch.randn(M, N, device="cuda") * 3 + 1).bfloat16() for M, N, _ in cases]
    outs = [torch.empty(N * (2 if sq else 1), dtype=torch.float32, device="cuda") for _, N, sq in cases]
    for x, (M, N, sq), out in zip(xs, cases, outs):  # back to back on one stream, nothing in between
        assert _noalloc(lib, x, sq, out) == 0
    torch.cuda.synchronize()
    for x, (M, N, sq), out in zip(xs, cases, outs):
        want = _regular(lib, x, sq)
        torch.cuda.synchronize()
        assert torch.allclose(out, want, rtol=1e-3, atol=1e-2 * max(1.0, M ** 0.5))
        xf = x.float()
        ref = torch.cat([xf.sum(0), (xf * xf).sum(0)]) if sq else xf.sum(0)
        assert torch.allclose(out, ref, rtol=1e-3, atol=1e-2 * max(1.0, M ** 0.5))


def test_noalloc_colsum_rejects_oversize_and_handles_empty():
    lib = _lib()
    x = torch.zeros(16, 2048, device="cuda", dtype=torch.bfloat16)
    out = torch.full((2048,), 7.0, dtype=torch.float32, device="cuda")
    assert _noalloc(lib, x, 0, out) != 0  # 2048 columns of partials > 1024: refused, nothing written
    assert _noalloc(lib, x[:, :1024], 1, out) != 0  # 1024 columns with squares = 2048 > 1024
    torch.cuda.synchronize()
    assert (out == 7.0).all()
    assert _noalloc(lib, x[:, :1024], 0, out) == 0  # exactly at the limit
    empty = torch.zeros(0, 64, device="cuda", dtype=torch.bfloat16)
    assert _noalloc(lib, empty, 0, out) == 0
    torch.cuda.synchronize()
    assert (out[:64] == 0).all()
