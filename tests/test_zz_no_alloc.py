"""mdgen_gemm_noalloc (include/mdgen_b200.h): the allocation-free GEMM entry point added
-- no counterpart in the reference.  Its bf16 operand scratch is a static __device__ array (8 MiB) instead of
cudaMalloc'ed buffers.  Checked through the C ABI against the existing entry point mdgen_test_gemm(use_tc = 1), which
runs the same kernels on cudaMalloc'ed scratch, on two calls in a row (the shared array keeps the first call's contents).

Kept in a file pytest collects LAST; passes alone in a fresh process and after the
other tests."""
import ctypes

import pytest
import torch

from mdgen_finetune_b200 import _lib

pytestmark = pytest.mark.gpu


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _problem(M, N, K, seed):
    g = torch.Generator().manual_seed(seed)
    A = torch.randn(M, K, generator=g)
    W = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    return A.cuda(), W.cuda(), b.cuda()


def _existing(M, N, K, A, W, b, st):
    C = torch.empty(M, N, device="cuda")
    _lib.check(_lib.lib().mdgen_test_gemm(M, N, K, _p(A), _p(W), _p(b), _p(C), 1, 0, st), "mdgen_test_gemm")
    return C


def _noalloc(M, N, K, A, W, b, st, C):
    _lib.check(_lib.lib().mdgen_gemm_noalloc(M, N, K, _p(A), _p(W), _p(b), _p(C), st), "mdgen_gemm_noalloc")


def test_noalloc_gemm_matches_existing_entry_point_twice_in_a_row():
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # second problem is SMALLER and differently shaped: stale rows of the first call stay in the scratch beyond it
    shapes = [(4000, 384, 384, 1), (777, 1152, 384, 2)]
    outs = [torch.empty(M, N, device="cuda") for M, N, K, _ in shapes]
    probs = [_problem(M, N, K, seed) for M, N, K, seed in shapes]
    torch.cuda.synchronize()
    # back to back on one stream, no synchronisation in between: same-stream calls are ordered
    for (M, N, K, _), (A, W, b), C in zip(shapes, probs, outs):
        _noalloc(M, N, K, A, W, b, st, C)
    torch.cuda.synchronize()
    for (M, N, K, _), (A, W, b), C in zip(shapes, probs, outs):
        ref = _existing(M, N, K, A, W, b, st)
        torch.cuda.synchronize()
        assert torch.isfinite(C).all()
        err = float((C - ref).norm() / ref.norm())
        assert err < 1e-5, err                      # same kernels, same operands: the tolerance of tests/test_gpu_kernels.py
        # and against plain arithmetic on the bf16-rounded operands (what both entry points compute)
        exact = (A.bfloat16().double() @ W.bfloat16().double().T + b.double()).float()
        assert float((C - exact).norm() / exact.norm()) < 1e-5


def test_noalloc_gemm_allocates_nothing_and_rejects_oversize():
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    M, N, K = 2048, 384, 1024                       # M * K = 2^21: exactly the limit
    assert M * K == _lib.NOALLOC_MAX_OPERAND_ELEMS
    A, W, b = _problem(M, N, K, 3)
    C = torch.empty(M, N, device="cuda")
    _noalloc(M, N, K, A, W, b, st, C)               # first call on this device: symbol address lookup, kernel attributes
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    alloc0 = torch.cuda.memory_allocated()
    for _ in range(3):
        _noalloc(M, N, K, A, W, b, st, C)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0 and torch.cuda.memory_allocated() == alloc0      # no cudaMalloc, no torch allocation
    exact = (A.bfloat16().double() @ W.bfloat16().double().T + b.double()).float()
    assert float((C - exact).norm() / exact.norm()) < 1e-5
    # one element past the limit: refused with a message that names the limit, nothing launched
    before = C.clone()
    rc = _lib.lib().mdgen_gemm_noalloc(M + 8, N, K, _p(A), _p(W), _p(b), _p(C), st)
    assert rc == -1
    msg = _lib.lib().mdgen_last_error().decode()
    assert "too large" in msg and str(_lib.NOALLOC_MAX_OPERAND_ELEMS) in msg, msg
    with pytest.raises(ValueError, match="MDGEN_NOALLOC_MAX_OPERAND_ELEMS"):
        _lib.check(_lib.lib().mdgen_gemm_noalloc(8, 4096, 1024, _p(A), _p(W), _p(b), _p(C), st), "mdgen_gemm_noalloc")
    torch.cuda.synchronize()
    assert torch.equal(C, before)


def test_noalloc_gemm_is_capturable():
    """The point of the entry point: usable where cudaMalloc is forbidden, e.g. under stream capture."""
    M, N, K = 1000, 384, 384
    A, W, b = _problem(M, N, K, 4)
    C = torch.zeros(M, N, device="cuda")
    st0 = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _noalloc(M, N, K, A, W, b, st0, C)              # warm: per-device set-up happens outside the capture
    torch.cuda.synchronize()
    ref = C.clone()
    C.zero_()
    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            _noalloc(M, N, K, A, W, b, ctypes.c_void_p(side.cuda_stream), C)
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(C, ref)
