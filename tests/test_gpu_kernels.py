"""Kernel-level parity on the B200, through the C-ABI test entry points."""
import ctypes

import pytest
import torch

from mdgen_finetune_b200 import _lib

pytestmark = pytest.mark.gpu


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _gemm(A, W, b, use_tc, block_n=0):
    M, K = A.shape
    N = W.shape[0]
    C = torch.full((M, N), float("nan"), device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().mdgen_test_gemm(M, N, K, _p(A), _p(W), _p(b), _p(C), use_tc, block_n, st), "test_gemm")
    torch.cuda.synchronize()
    return C


@pytest.mark.parametrize("M,N,K", [(64, 64, 16), (100, 72, 40), (257, 672, 384), (33, 384, 256), (5, 20, 7)])
def test_gemm_simt_fp32(M, N, K):
    g = torch.Generator().manual_seed(M + N + K)
    A, W, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g), torch.randn(N, generator=g)
    C = _gemm(A.cuda(), W.cuda(), b.cuda(), 0).cpu()
    ref = (A.double() @ W.double().T + b.double()).float()
    assert float((C - ref).norm() / ref.norm()) < 2e-6


@pytest.mark.parametrize("M,N,K,bn", [(128, 128, 64, 128), (128, 128, 384, 128), (256, 256, 384, 128),
                                      (4000, 1152, 384, 0), (4000, 384, 1536, 0), (1000, 384, 384, 64),
                                      (777, 1536, 384, 128), (300, 21, 384, 0), (129, 48, 384, 64)])
def test_gemm_tcgen05_bf16(M, N, K, bn):
    g = torch.Generator().manual_seed(M + N + K)
    A, W, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / K ** 0.5, torch.randn(N, generator=g)
    C = _gemm(A.cuda(), W.cuda(), b.cuda(), 1, bn).cpu()
    ref = (A.bfloat16().double() @ W.bfloat16().double().T + b.double()).float()   # same bf16 operands
    assert torch.isfinite(C).all(), "unwritten / non-finite outputs"
    err = float((C - ref).norm() / ref.norm())
    assert err < 1e-5, err          # fp32 accumulation of identical bf16 products


@pytest.mark.parametrize("M,N,K,bn", [(4000, 1152, 384, 0), (777, 1536, 384, 128), (1000, 384, 384, 64), (129, 48, 384, 64),
                                      (300, 72, 64, 0), (64, 24, 384, 32),
                                      # block_n 256 = the CTA-pair (cta_group::2) kernel, 256 x 256 tiles
                                      (256, 256, 64, 256), (4000, 1152, 384, 256), (777, 1536, 384, 256), (300, 128, 64, 256),
                                      (20000, 384, 384, 256)])
@pytest.mark.parametrize("mode", [2, 3])
def test_gemm_tcgen05_bf16_out_staged_epilogue(M, N, K, bn, mode):
    g = torch.Generator().manual_seed(M + N + K)
    A, W, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / K ** 0.5, torch.randn(N, generator=g)
    C = _gemm(A.cuda(), W.cuda(), b.cuda(), mode, bn).cpu()
    ref = A.bfloat16().double() @ W.bfloat16().double().T + b.double()
    if mode == 3:
        ref = ref * 0.5 * (1.0 + torch.erf(ref / 2 ** 0.5))          # layers.py:78-85
    assert torch.isfinite(C).all()
    err = float((C - ref.float()).norm() / ref.norm())
    assert err < 4e-3, err          # one bf16 rounding of the output


@pytest.mark.parametrize("offset", [0.3, 40.0])
@pytest.mark.parametrize("M,K", [(128, 384), (300, 384), (1000, 384), (4000, 384), (4000, 1536), (6400, 384), (777, 64)])
def test_gemm_tcgen05_fused_residual_layernorm_cluster_of_four(M, K, offset):
    """gemm_ln4.cuh (small token counts): the row split over a cluster of 4 CTAs, LayerNorm statistics exchanged through
    distributed shared memory.  Same operands and reference as the pair kernel's test; the residual stream must equal the
    pair kernel's bit for bit (same accumulation order), the normalised rows to bf16 rounding."""
    N = 384
    g = torch.Generator().manual_seed(M + K)
    A, W = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / K ** 0.5
    vec = torch.randn(7, N, generator=g) * 0.5
    bias, gate, shift, scale = vec[0], vec[1:3], vec[3:5], vec[5:7]
    x0 = torch.randn(M, N, generator=g) * 2.0 + offset
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dA, dW, dv = A.cuda(), W.cuda(), vec.contiguous().cuda()
    L = _lib.lib()
    idx = _lib.VARIANTS.index("gemm_ln4")
    before = L.mdgen_variant_launches(idx)
    C = torch.cat([x0, torch.zeros(M, N)]).cuda()
    _lib.check(L.mdgen_test_gemm(M, N, K, _p(dA), _p(dW), _p(dv), _p(C), 5, 4, st), "test_gemm")
    torch.cuda.synchronize()
    assert L.mdgen_variant_launches(idx) == before + 1
    b = (torch.arange(M) // 200) % 2
    acc = A.bfloat16().double() @ W.bfloat16().double().T + bias.double()
    x = x0.double() + gate[b].double() * acc
    h = torch.nn.functional.layer_norm(x, (N,), eps=1e-6) * (1 + scale[b].double()) + shift[b].double()
    out = C.cpu()
    assert torch.isfinite(out).all()
    assert float((out[:M] - x.float()).norm() / x.norm()) < 1e-5
    assert float((out[M:] - h.float()).norm() / h.norm()) < 4e-3
    if M >= 256:
        P = torch.cat([x0, torch.zeros(M, N)]).cuda()
        _lib.check(L.mdgen_test_gemm(M, N, K, _p(dA), _p(dW), _p(dv), _p(P), 5, 0, st), "test_gemm")
        torch.cuda.synchronize()
        assert torch.equal(P[:M], C[:M])
        assert float((P[M:] - C[M:]).abs().max()) <= 0.0625       # at most one bf16 ulp apart at |h| < 8
    for _ in range(4):                                             # repeatable (the exchange has no race)
        R = torch.cat([x0, torch.zeros(M, N)]).cuda()
        _lib.check(L.mdgen_test_gemm(M, N, K, _p(dA), _p(dW), _p(dv), _p(R), 5, 4, st), "test_gemm")
        torch.cuda.synchronize()
        assert torch.equal(R, C)


@pytest.mark.parametrize("env", ["MDGEN_PAIR16_STORE", "MDGEN_NO_PAIR16"])
@pytest.mark.parametrize("M,N,K", [(4000, 1152, 384), (777, 1536, 384), (300, 128, 64), (20000, 1536, 384)])
@pytest.mark.parametrize("mode", [2, 3])
def test_gemm_pair_kernel_epilogue_warp_variants(M, N, K, mode, env, monkeypatch):
    """gemm_tc2.cuh: 16 epilogue warps (default for GELU; MDGEN_PAIR16_STORE=1 also for plain stores) against 8
    (MDGEN_NO_PAIR16=1) -- same tile arithmetic, so the outputs must be bit-identical to each other."""
    g = torch.Generator().manual_seed(M + N + K)
    A, W, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / K ** 0.5, torch.randn(N, generator=g)
    base = _gemm(A.cuda(), W.cuda(), b.cuda(), mode, 256).cpu()
    monkeypatch.setenv(env, "1")
    C = _gemm(A.cuda(), W.cuda(), b.cuda(), mode, 256).cpu()
    assert torch.isfinite(C).all()
    assert torch.equal(C, base)
    ref = A.bfloat16().double() @ W.bfloat16().double().T + b.double()
    if mode == 3:
        ref = ref * 0.5 * (1.0 + torch.erf(ref / 2 ** 0.5))
    assert float((C - ref.float()).norm() / ref.norm()) < 4e-3


@pytest.mark.parametrize("M,N,K,bn", [(4000, 384, 1536, 0), (4000, 384, 384, 0), (777, 384, 384, 128), (129, 48, 384, 64),
                                      (300, 21, 384, 0), (100, 70, 64, 64),
                                      (256, 256, 384, 256), (4000, 384, 1536, 256), (777, 384, 384, 256), (20000, 1152, 384, 256)])
def test_gemm_tcgen05_gated_residual_epilogue(M, N, K, bn):
    g = torch.Generator().manual_seed(M + N + K)
    A, W, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / K ** 0.5, torch.randn(N, generator=g)
    x0 = torch.randn(M, N, generator=g)
    M_, K_ = A.shape
    C = x0.clone().cuda()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dA, dW, db = A.cuda(), W.cuda(), b.cuda()
    _lib.check(_lib.lib().mdgen_test_gemm(M, N, K, _p(dA), _p(dW), _p(db), _p(C), 4, bn, st), "test_gemm")
    torch.cuda.synchronize()
    acc = A.bfloat16().double() @ W.bfloat16().double().T + b.double()
    ref = (x0.double() + b.double() * acc).float()
    assert float((C.cpu() - ref).norm() / ref.norm()) < 1e-5


@pytest.mark.parametrize("offset", [0.3, 40.0])        # 40: row mean 20 standard deviations away from zero (shifted one-pass variance)
@pytest.mark.parametrize("M,K", [(256, 384), (1000, 384), (4000, 384), (4000, 1536), (20000, 384), (777, 64)])
def test_gemm_tcgen05_fused_residual_layernorm(M, K, offset):
    """gemm_ln.cuh: x += gate * (A W^T + b); h = LN(x) * (1 + scale) + shift, batch element switching every 200 rows."""
    N = 384
    g = torch.Generator().manual_seed(M + K)
    A, W = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / K ** 0.5
    # four DISTINCT operands (a swap of gate / shift / scale / bias inside the kernel changes the expected value):
    # vec = [linear bias (1, N) | gate (2, N) | shift (2, N) | scale (2, N)], row b of each pair = batch element b
    vec = torch.randn(7, N, generator=g) * 0.5
    bias, gate, shift, scale = vec[0], vec[1:3], vec[3:5], vec[5:7]
    x0 = torch.randn(M, N, generator=g) * 2.0 + offset
    C = torch.cat([x0, torch.zeros(M, N)]).cuda()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dA, dW, dv = A.cuda(), W.cuda(), vec.contiguous().cuda()
    _lib.check(_lib.lib().mdgen_test_gemm(M, N, K, _p(dA), _p(dW), _p(dv), _p(C), 5, 0, st), "test_gemm")
    torch.cuda.synchronize()
    b = (torch.arange(M) // 200) % 2              # batch boundaries fall inside the 256-row slabs
    acc = A.bfloat16().double() @ W.bfloat16().double().T + bias.double()
    x = x0.double() + gate[b].double() * acc
    h = torch.nn.functional.layer_norm(x, (N,), eps=1e-6) * (1 + scale[b].double()) + shift[b].double()
    out = C.cpu()
    assert torch.isfinite(out).all()
    assert float((out[:M] - x.float()).norm() / x.norm()) < 1e-5
    assert float((out[M:] - h.float()).norm() / h.norm()) < 4e-3      # one bf16 rounding
    # sanity of the test itself: swapping shift and scale must be visible at this tolerance
    h_sw = torch.nn.functional.layer_norm(x, (N,), eps=1e-6) * (1 + shift[b].double()) + scale[b].double()
    assert float((h_sw - h).norm() / h.norm()) > 0.1


@pytest.mark.parametrize("M", [256, 1000, 4000, 20000, 38000])
def test_gemm_tcgen05_fused_mlp_block(M):
    """gemm_mlp.cuh: hid = GELU(A W1^T + b1); x += gate * (hid W2^T + b2); h' = LN(x) * (1 + scale) + shift, batch element
    switching every 200 rows.  38000 rows = 149 tiles: every CTA pair runs more than two tiles (ring / barrier phases wrap)."""
    C, F = 384, 1536
    g = torch.Generator().manual_seed(M)
    A = torch.randn(M, C, generator=g)
    W1, W2 = torch.randn(F, C, generator=g) / C ** 0.5, torch.randn(C, F, generator=g) / F ** 0.5
    b1 = torch.randn(F, generator=g) * 0.5
    vec = torch.randn(7, C, generator=g) * 0.5
    b2, gate, shift, scale = vec[0], vec[1:3], vec[3:5], vec[5:7]
    x0 = torch.randn(M, C, generator=g) * 2.0 + 0.3
    Cbuf = torch.cat([x0, torch.zeros(M, C)]).cuda()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dA, dW = A.cuda(), torch.cat([W1.flatten(), W2.flatten()]).cuda()
    dv = torch.cat([b1, vec.flatten()]).contiguous().cuda()
    _lib.check(_lib.lib().mdgen_test_gemm(M, C, C, _p(dA), _p(dW), _p(dv), _p(Cbuf), 6, 0, st), "test_gemm")
    torch.cuda.synchronize()
    b = (torch.arange(M) // 200) % 2
    hid = A.bfloat16().double() @ W1.bfloat16().double().T + b1.double()
    hid = (hid * 0.5 * (1.0 + torch.erf(hid / 2 ** 0.5))).bfloat16().double()          # the hidden activations are bf16
    acc = hid @ W2.bfloat16().double().T + b2.double()
    x = x0.double() + gate[b].double() * acc
    hn = torch.nn.functional.layer_norm(x, (C,), eps=1e-6) * (1 + scale[b].double()) + shift[b].double()
    out = Cbuf.cpu()
    assert torch.isfinite(out).all()
    # the GELU rounding to bf16 can differ by one ulp from torch's erf on a few elements: 1e-4 on the fp32 stream
    assert float((out[:M] - x.float()).norm() / x.norm()) < 1e-4
    assert float((out[M:] - hn.float()).norm() / hn.norm()) < 4e-3


@pytest.mark.parametrize("M", [256, 300, 1000, 5001, 20000, 37999])
def test_fused_mlp_block_bit_identical_to_the_two_kernel_path(M):
    """The one-kernel MLP block accumulates in the same order as fc1 + GELU (pair GEMM) followed by fc2 + residual +
    LayerNorm (gemm_ln.cuh): both the residual stream and the normalised rows must agree BIT FOR BIT, ragged M included."""
    C, F = 384, 1536
    g = torch.Generator().manual_seed(7 * M)
    A = torch.randn(M, C, generator=g)
    W1, W2 = torch.randn(F, C, generator=g) / C ** 0.5, torch.randn(C, F, generator=g) / F ** 0.5
    b1 = torch.randn(F, generator=g) * 0.5
    vec = torch.randn(7, C, generator=g) * 0.5
    x0 = torch.randn(M, C, generator=g)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    L = _lib.lib()
    dA = A.cuda()
    fused = torch.cat([x0, torch.zeros(M, C)]).cuda()
    dW = torch.cat([W1.flatten(), W2.flatten()]).cuda()
    dv = torch.cat([b1, vec.flatten()]).contiguous().cuda()
    _lib.check(L.mdgen_test_gemm(M, C, C, _p(dA), _p(dW), _p(dv), _p(fused), 6, 0, st), "test_gemm")
    hid = _gemm(dA, W1.cuda(), b1.cuda(), 3, 0)                       # bf16 values, returned as fp32: exact A operand of fc2
    two = torch.cat([x0, torch.zeros(M, C)]).cuda()
    dW2, dvec = W2.contiguous().cuda(), vec.contiguous().cuda()
    _lib.check(L.mdgen_test_gemm(M, C, F, _p(hid), _p(dW2), _p(dvec), _p(two), 5, 0, st), "test_gemm")
    torch.cuda.synchronize()
    assert torch.isfinite(fused).all()
    assert torch.equal(fused[:M], two[:M]), float((fused[:M] - two[:M]).abs().max())
    assert torch.equal(fused[M:], two[M:]), float((fused[M:] - two[M:]).abs().max())
    for _ in range(8):                    # the same launch again and again: a race in the hand-offs would show as a changed bit
        again = torch.cat([x0, torch.zeros(M, C)]).cuda()
        _lib.check(L.mdgen_test_gemm(M, C, C, _p(dA), _p(dW), _p(dv), _p(again), 6, 0, st), "test_gemm")
        torch.cuda.synchronize()
        assert torch.equal(again, fused)


def _attn_ref(qkv, bk, bv, valid, nb, S, inner, H, hd, rope):
    from oracle import mdgen_oracle as O
    C = H * hd
    rows = torch.tensor([[(b // inner) * S * inner + s * inner + (b % inner) for s in range(S)] for b in range(nb)])
    x = qkv[rows]                                           # [nb,S,3C]
    q, k, v = x[..., :C], x[..., C:2 * C], x[..., 2 * C:]
    k = torch.cat([k, bk.view(1, 1, C).expand(nb, 1, C)], 1)
    v = torch.cat([v, bv.view(1, 1, C).expand(nb, 1, C)], 1)
    pad = torch.cat([valid[rows] == 0, torch.zeros(nb, 1, dtype=torch.bool)], 1)
    q = q.view(nb, S, H, hd).transpose(1, 2)
    k = k.view(nb, S + 1, H, hd).transpose(1, 2)
    v = v.view(nb, S + 1, H, hd).transpose(1, 2)
    if rope:
        cos, sin = O.rope_tables(S + 1, hd)
        q = q * cos[:S] + O._rot_half(q) * sin[:S]
        k = k * cos + O._rot_half(k) * sin
    a = (q @ k.transpose(-1, -2)).masked_fill(pad[:, None, None, :], float("-inf")).softmax(-1)
    o = (a @ v).transpose(1, 2).reshape(nb, S, C)
    out = torch.zeros(nb * S, C)
    out[rows.reshape(-1)] = o.reshape(nb * S, C)
    return out


@pytest.mark.parametrize("nb,S,inner,rope", [(6, 4, 1, 1), (8, 37, 4, 1), (4, 200, 2, 1), (3, 130, 1, 0), (2, 1000, 2, 1)])
def test_attention_simt(nb, S, inner, rope):
    H, hd = 16, 24
    C = H * hd
    g = torch.Generator().manual_seed(nb * 1000 + S)
    qkv = torch.randn(nb * S, 3 * C, generator=g) * 0.7
    bk, bv = torch.randn(C, generator=g), torch.randn(C, generator=g)
    valid = (torch.rand(nb * S, generator=g) > 0.2).float()
    valid[:S * inner] = 1.0
    if nb > 2:
        valid[attn_rows(1, S, inner)] = 0.0      # one fully padded sequence: only the bias token is attended
    out = torch.empty(nb * S, C, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dq, dk, dv, dm = qkv.cuda(), bk.cuda(), bv.cuda(), valid.cuda()     # keep alive across the call
    _lib.check(_lib.lib().mdgen_test_attention(nb, S, inner, H, hd, _p(dq), _p(dk), _p(dv), _p(dm), _p(out), 0,
                                               rope, st), "test_attention")
    ref = _attn_ref(qkv, bk, bv, valid, nb, S, inner, H, hd, rope)
    assert float((out.cpu() - ref).norm() / ref.norm()) < 1e-5


@pytest.mark.parametrize("nb,S,inner,rope", [(6, 4, 1, 1), (1001, 4, 1, 1), (8000, 4, 1, 1), (10, 4, 1, 0), (12, 4, 4, 1), (9, 3, 1, 1)])
def test_attention_short_bf16(nb, S, inner, rope):
    """bf16 short-axis kernels (residue axis of the 4AA configs): bulk-copy kernel for contiguous S = 4 sequences,
    thread-per-(sequence, head) kernel otherwise."""
    H, hd = 16, 24
    C = H * hd
    g = torch.Generator().manual_seed(nb * 1000 + S)
    qkv = (torch.randn(nb * S, 3 * C, generator=g) * 0.7).bfloat16().float()
    bk, bv = torch.randn(C, generator=g), torch.randn(C, generator=g)
    valid = (torch.rand(nb * S, generator=g) > 0.2).float()
    valid[:S * inner] = 1.0
    if nb > 2:
        valid[attn_rows(1, S, inner)] = 0.0
    out = torch.full((nb * S, C), float("nan"), device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dq, dk, dv, dm = qkv.cuda(), bk.cuda(), bv.cuda(), valid.cuda()
    _lib.check(_lib.lib().mdgen_test_attention(nb, S, inner, H, hd, _p(dq), _p(dk), _p(dv), _p(dm), _p(out), 2,
                                               rope, st), "test_attention")
    ref = _attn_ref(qkv, bk, bv, valid, nb, S, inner, H, hd, rope)
    o = out.cpu()
    assert torch.isfinite(o).all()
    err = float((o - ref).norm() / ref.norm())
    assert err < 8e-3, err      # bf16 q/k/v (K rotated in bf16), fp32 arithmetic, bf16 output


def attn_rows(b, S, inner):
    return torch.tensor([(b // inner) * S * inner + s * inner + (b % inner) for s in range(S)])


@pytest.mark.parametrize("nb,S,inner,rope", [(4, 200, 2, 1), (3, 130, 1, 0), (2, 1000, 2, 1), (5, 256, 1, 1),
                                              (4, 64, 4, 1), (2, 1023, 1, 1), (3, 513, 1, 1), (2, 1500, 2, 1), (3, 95, 1, 1), (40, 100, 4, 1)])
def test_attention_tcgen05(nb, S, inner, rope):
    H, hd = 16, 24
    C = H * hd
    g = torch.Generator().manual_seed(nb * 1000 + S)
    qkv = (torch.randn(nb * S, 3 * C, generator=g) * 0.7).bfloat16().float()
    bk, bv = torch.randn(C, generator=g), torch.randn(C, generator=g)
    valid = (torch.rand(nb * S, generator=g) > 0.2).float()
    valid[:S * inner] = 1.0
    if nb > 2:
        valid[attn_rows(1, S, inner)] = 0.0
    out = torch.full((nb * S, C), float("nan"), device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dq, dk, dv, dm = qkv.cuda(), bk.cuda(), bv.cuda(), valid.cuda()
    _lib.check(_lib.lib().mdgen_test_attention(nb, S, inner, H, hd, _p(dq), _p(dk), _p(dv), _p(dm), _p(out), 1,
                                               rope, st), "test_attention")
    ref = _attn_ref(qkv, bk, bv, valid, nb, S, inner, H, hd, rope)
    o = out.cpu()
    assert torch.isfinite(o).all()
    err = float((o - ref).norm() / ref.norm())
    assert err < 8e-3, err      # bf16 q/k/v/P operands, fp32 accumulation and softmax statistics
