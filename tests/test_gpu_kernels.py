"""Per-kernel parity on a B200: every CUDA kernel, called through the C ABI on raw device pointers,
against the oracle's corresponding function (oracle/ops.py) on the same seeded inputs.

Tolerances: fp32 outputs (residual, split-K slices, logits, routing weights): max abs err 1e-4 x the
value scale; bf16 outputs: within 1 bf16 ulp (2^-8 relative) of the rounded oracle value; integer
outputs (tokens, synthetic weight bits): bit-exact.
"""
import ctypes as C
import math

import pytest
import torch

from butterfly_b200 import _lib
from butterfly_b200.kv_layout import pack_pages, unpack_pages
from butterfly_b200.weights import synth_block, synth_scale, tensor_key
from oracle import ops

pytestmark = pytest.mark.gpu

# attention multiplies bf16-rounded probabilities (<= 1, relative error 2^-9) with |v| <~ 4 on tensor cores
ATTN_ATOL = 1.6e-2


@pytest.fixture(scope="module")
def lib(cuda):
    return _lib.load()


def stream():
    return torch.cuda.current_stream().cuda_stream


def ok(lib, rc):
    _lib.check(rc)
    torch.cuda.synchronize()


def bf16_close(got: torch.Tensor, want_f32: torch.Tensor, extra_abs: float = 0.0):
    want = want_f32.to(torch.bfloat16).float()
    tol = want.abs() * 2.0 ** -7 + 1e-6 + extra_abs
    bad = ((got.float().cpu() - want).abs() > tol)
    assert not bad.any(), f"{int(bad.sum())} of {bad.numel()} elements off by more than 1 bf16 ulp; max abs err " \
                          f"{(got.float().cpu() - want).abs().max().item():.3e}"


_keep = []


def D(t: torch.Tensor) -> int:
    """Device pointer of a CUDA copy of `t` that stays alive for the whole test (a temporary would be
    freed -- and its block reused by the next copy -- before the kernel is even launched)."""
    d = t.to("cuda").contiguous()
    _keep.append(d)
    return d.data_ptr()


@pytest.fixture(autouse=True)
def _release_keepalive():
    yield
    torch.cuda.synchronize()
    _keep.clear()


def gen(shape, seed, scale=1.0, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype)


# ---------------------------------------------------------------------------------------------
def test_fill_synth_bit_exact(lib, cuda):
    key = tensor_key(0, 3, "wq")
    for (r0, r1, c0, c1, ncols, std, off) in [(0, 64, 0, 128, 128, 0.02, 0.0), (5, 77, 3, 250, 4096, 0.02, 0.0),
                                              (0, 512, 0, 1, 1, 0.02, 1.0), (100, 101, 7, 4000, 4096, 0.002, 0.0)]:
        want = synth_block(key, ncols, r0, r1, c0, c1, synth_scale(std), off)
        got = torch.zeros((r1 - r0, c1 - c0), dtype=torch.bfloat16, device=cuda)
        ok(lib, lib.bfly_k_fill_synth(got.data_ptr(), c1 - c0, key, ncols, r0, c0, r1 - r0, c1 - c0,
                                      float(synth_scale(std)), off, stream()))
        assert torch.equal(got.cpu().view(torch.int16), want.view(torch.int16))


def test_embed(lib, cuda):
    V, h, B, S = 300, 256, 3, 5
    table, pos = gen((V, h), 1, 0.02, torch.bfloat16), gen((64, h), 2, 0.01, torch.bfloat16)
    g = torch.Generator().manual_seed(3)
    toks = torch.randint(0, V, (B * S,), generator=g, dtype=torch.int32)
    lens = torch.tensor([0, 7, 11], dtype=torch.int32)
    positions = (lens[:, None] + torch.arange(S)[None, :]).reshape(-1)
    for use_pos in (False, True):
        out = torch.empty((B * S, h), dtype=torch.float32, device=cuda)
        ok(lib, lib.bfly_k_embed(D(toks), D(table),
                                 D(pos) if use_pos else None, D(lens), S,
                                 out.data_ptr(), B * S, h, V, stream()))
        want = ops.embed(toks, table.float(), pos.float() if use_pos else None, positions)
        assert (out.cpu() - want).abs().max().item() < 1e-7


@pytest.mark.parametrize("h,layer_norm,n_part", [(256, False, 0), (4096, False, 3), (768, True, 2), (8192, False, 1)])
def test_add_norm(lib, cuda, h, layer_norm, n_part):
    T = 7
    res, parts = gen((T, h), 1), gen((max(n_part, 1), T, h), 2, 0.3)
    bias, gamma, beta = gen((h,), 3, 0.1, torch.bfloat16), (1 + gen((h,), 4, 0.05)).to(torch.bfloat16), \
        gen((h,), 5, 0.05, torch.bfloat16)
    d_res, d_out = res.to(cuda), torch.empty((T, h), dtype=torch.bfloat16, device=cuda)
    ok(lib, lib.bfly_k_add_norm(d_res.data_ptr(), D(parts) if n_part else None, n_part, T * h,
                                D(bias), D(gamma),
                                D(beta) if layer_norm else None, 1e-5, int(layer_norm),
                                d_out.data_ptr(), T, h, stream()))
    x = res.clone()
    for p in range(n_part):
        x = x + parts[p]
    x = x + bias.float()
    assert (d_res.cpu() - x).abs().max().item() < 1e-5
    want = ops.norm(x, gamma.float(), beta.float() if layer_norm else None, 1e-5, layer_norm)
    bf16_close(d_out, want, extra_abs=2e-5)


GEMM_SHAPES = [
    # M (weight rows), N (tokens), K
    (128, 16, 64), (256, 32, 256), (384, 32, 4096), (1000, 7, 768), (6144, 32, 4096), (4096, 32, 1792),
    (500, 100, 520), (256, 300, 1024), (3072, 64, 768), (128, 1, 8192), (2048, 512, 512),
    (4096, 2048, 512), (1152, 1000, 320),      # persistent kernel: several units per CTA pair, odd tile counts
]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_split_slices(lib, cuda, M, N, K):
    W, X = gen((M, K), 1, 0.05, torch.bfloat16).to(cuda), gen((N, K), 2, 1.0, torch.bfloat16).to(cuda)
    want = (X.float() @ W.float().t()).cpu()                       # [N, M] fp32 reference on the same bf16 inputs
    for splits in (0, 1, 3):
        if splits > (K + 63) // 64:
            continue
        used = C.c_int32(0)
        cap = 16
        out = torch.full((cap, N, M), float("nan"), dtype=torch.float32, device=cuda)
        ok(lib, lib.bfly_k_gemm(W.data_ptr(), X.data_ptr(), M, N, K, K, splits, 0, 0, out.data_ptr(), N * M, None, M,
                                C.byref(used), stream()))
        got = out[:used.value].sum(0).cpu()
        assert not torch.isnan(got).any(), f"unwritten outputs with splits={used.value}"
        err = (got - want).abs().max().item()
        # bf16-exact inputs, fp32 accumulation on both sides: only the summation order differs
        assert err < 2e-4 * want.abs().max().item(), f"splits={used.value} max err {err}"


@pytest.mark.parametrize("N,K,n_blocks,fb,splits,routed", [
    (32, 1024, 1, 512, 1, False), (32, 4096, 1, 1792, 4, False), (32, 2048, 1, 256, 8, False), (7, 512, 3, 200, 2, True),
    (300, 1024, 1, 384, 1, False), (64, 768, 2, 128, 3, True), (20, 4096, 1, 3584, 0, False),
    (1500, 512, 1, 2048, 0, False), (700, 256, 3, 200, 0, True)])      # persistent dual kernel
def test_gemm_fused_swiglu(lib, cuda, N, K, n_blocks, fb, splits, routed):
    """gate/up projection with the SwiGLU (and MoE routing weight) in the epilogue; splits > 1 reduces the
    K-slices inside a thread-block cluster through distributed shared memory."""
    W = gen((n_blocks * 2 * fb, K), 1, 0.03, torch.bfloat16).to(cuda)
    X = gen((N, K), 2, 1.0, torch.bfloat16).to(cuda)
    rw = torch.rand((N, n_blocks), generator=torch.Generator().manual_seed(3))
    rw[rw < 0.3] = 0.0
    rw_d = rw.to(cuda)
    out = torch.full((N, n_blocks * fb), float("nan"), dtype=torch.bfloat16, device=cuda)
    ok(lib, lib.bfly_k_gemm_swiglu(W.data_ptr(), X.data_ptr(), N, K, n_blocks, fb, rw_d.data_ptr() if routed else None,
                                   splits, out.data_ptr(), stream()))
    gu = (X.float() @ W.float().t()).cpu().reshape(N, n_blocks, 2 * fb)
    want = ops.swiglu(gu)
    if routed:
        want = want * rw[:, :, None]
    assert not torch.isnan(out.float()).any()
    bf16_close(out, want.reshape(N, n_blocks * fb), extra_abs=2e-3)


def test_gemm_cluster_reduce_is_deterministic(lib, cuda):
    M, N, K = 4096, 32, 4096
    W, X = gen((M, K), 1, 0.05, torch.bfloat16).to(cuda), gen((N, K), 2, 1.0, torch.bfloat16).to(cuda)
    outs = []
    for _ in range(3):
        out = torch.zeros((N, M), dtype=torch.float32, device=cuda)
        used = C.c_int32(0)
        ok(lib, lib.bfly_k_gemm(W.data_ptr(), X.data_ptr(), M, N, K, K, 8, 0, 0, out.data_ptr(), N * M, None, M,
                                C.byref(used), stream()))
        assert used.value == 1
        outs.append(out.cpu())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    assert (outs[0] - (X.float() @ W.float().t()).cpu()).abs().max().item() < 5e-3


def test_gemm_bf16_out_and_pitched_x(lib, cuda):
    M, N, K, S = 512, 6, 1024, 5
    W = gen((M, K), 1, 0.05, torch.bfloat16).to(cuda)
    Xall = gen((N * S, K), 2, 1.0, torch.bfloat16).to(cuda)
    X_last = Xall.view(N, S, K)[:, S - 1]                           # rows b*S + S-1, pitch S*K
    want = (X_last.float() @ W.float().t()).cpu()
    out = torch.zeros((N, M), dtype=torch.bfloat16, device=cuda)
    used = C.c_int32(0)
    ok(lib, lib.bfly_k_gemm(W.data_ptr(), Xall.data_ptr() + (S - 1) * K * 2, M, N, K, S * K, 1, 0, 0, None, 0,
                            out.data_ptr(), M, C.byref(used), stream()))
    assert used.value == 1
    bf16_close(out, want, extra_abs=2e-3)


@pytest.mark.parametrize("S", [4, 45])          # S >= 32 takes the tiled prefill kernel
@pytest.mark.parametrize("d,nq,nk,theta,use_bias", [(128, 4, 1, 5e5, False), (64, 6, 6, 0.0, True), (32, 4, 2, 1e6, False)])
def test_rope_kv(lib, cuda, d, nq, nk, theta, use_bias, S):
    B, ps, pps, n_part = 3, 32, 3, 2
    T, ld = B * S, (nq + 2 * nk) * d
    lens = torch.tensor([0, 9, 21], dtype=torch.int32)
    qkv = gen((n_part, T, ld), 1, 0.7)
    bias = gen((ld,), 2, 0.1, torch.bfloat16)
    perm = torch.randperm(B * pps, generator=torch.Generator().manual_seed(5)).to(torch.int32).reshape(B, pps)
    pages = torch.zeros((B * pps, 2, nk, ps * d), dtype=torch.bfloat16, device=cuda)
    half = d // 2
    max_pos = 96
    inv = theta ** (-torch.arange(0, half, dtype=torch.float64) * 2.0 / d) if theta > 0 else None
    table = None
    if theta > 0:
        ang = torch.arange(max_pos, dtype=torch.float64)[:, None] * inv[None, :]
        table = torch.stack([ang.cos().float(), ang.sin().float()], dim=-1).contiguous().to(cuda)   # [pos, half, 2]
    q_out = torch.zeros((T, nq * d), dtype=torch.bfloat16, device=cuda)
    ok(lib, lib.bfly_k_rope_kv(D(qkv), n_part, T * ld, D(bias) if use_bias else None,
                               table.data_ptr() if table is not None else None, D(lens), S, T, nq, nk,
                               d, q_out.data_ptr(), pages.data_ptr(), D(perm), pps, ps, stream()))
    x = qkv.sum(0) + (bias.float() if use_bias else 0.0)
    positions = (lens[:, None] + torch.arange(S)[None, :]).reshape(-1)
    q = x[:, :nq * d].reshape(T, nq, d)
    k = x[:, nq * d:(nq + nk) * d].reshape(T, nk, d)
    v = x[:, (nq + nk) * d:].reshape(T, nk, d)
    if theta > 0:
        q, k = ops.rope(q, positions, theta), ops.rope(k, positions, theta)
    bf16_close(q_out, q.reshape(T, nq * d), extra_abs=1e-5)
    pk, pv = unpack_pages(pages.cpu(), ps, d)
    for t in range(T):
        b, pos = t // S, int(positions[t])
        page, slot = int(perm[b, pos // ps]), pos % ps
        bf16_close(pk[page, :, slot, :], k[t], extra_abs=1e-5)
        bf16_close(pv[page, :, slot, :], v[t], extra_abs=1e-5)


@pytest.mark.parametrize("d,nq,nk,ctx_lens,n_splits", [
    (128, 4, 1, [1, 33, 257], 1), (128, 8, 2, [100, 5, 64], 4), (128, 8, 1, [300, 17, 96], 3),
    (64, 6, 6, [40, 41, 2], 2), (32, 8, 4, [70, 1, 128], 5), (128, 4, 1, [2048, 1999, 31], 9)])
def test_attn_decode(lib, cuda, d, nq, nk, ctx_lens, n_splits):
    B, ps = len(ctx_lens), 32
    pps = (max(ctx_lens) + ps - 1) // ps
    g = torch.Generator().manual_seed(7)
    perm = torch.randperm(B * pps, generator=g).to(torch.int32).reshape(B, pps)
    kl, vl = gen((B * pps, nk, ps, d), 8, 1.0, torch.bfloat16), gen((B * pps, nk, ps, d), 10, 1.0, torch.bfloat16)
    pages = pack_pages(kl, vl)
    q = gen((B, nq * d), 9, 1.0, torch.bfloat16)
    lens = torch.tensor([c - 1 for c in ctx_lens], dtype=torch.int32)        # tokens cached before the query
    out = torch.zeros((B, nq * d), dtype=torch.bfloat16, device=cuda)
    part_o = torch.zeros((B * nq * n_splits * d,), dtype=torch.float32, device=cuda)
    part_ml = torch.zeros((B * nq * n_splits * 2,), dtype=torch.float32, device=cuda)
    ok(lib, lib.bfly_k_attn_decode(D(q), D(pages), D(perm), pps, ps,
                                   nk, d, D(lens), 1, B, nq, part_o.data_ptr(), part_ml.data_ptr(),
                                   n_splits, out.data_ptr(), stream()))
    for b, ctx in enumerate(ctx_lens):
        pos = torch.arange(ctx)
        pg = perm[b, pos // ps].long()
        kk, vv = kl[pg, :, pos % ps, :].float(), vl[pg, :, pos % ps, :].float()   # [ctx, nk, d]
        want = ops.attention(q[b].float().reshape(1, 1, nq, d), kk[None], vv[None], ctx - 1)
        bf16_close(out[b], want.reshape(-1), extra_abs=ATTN_ATOL)


@pytest.mark.parametrize("d,nq,nk,ctx_lens,n_splits,n_part,theta,use_bias", [
    (128, 4, 1, [1, 33, 257], 1, 3, 5e5, False), (128, 8, 2, [100, 5, 64], 4, 16, 5e5, False),
    (128, 8, 1, [300, 17, 96, 32], 3, 1, 1e6, False), (64, 6, 6, [40, 41, 2], 2, 5, 0.0, True),
    (32, 8, 4, [70, 1, 128], 5, 2, 1e4, True)])
def test_attn_decode_with_fused_rope(lib, cuda, d, nq, nk, ctx_lens, n_splits, n_part, theta, use_bias):
    """bfly_k_attn_decode_rope (RoPE + KV append + attention in one kernel, the engine's decode path) gives the
    same bits as bfly_k_rope_kv followed by bfly_k_attn_decode, writes the same cache, and matches the oracle."""
    B, ps = len(ctx_lens), 32
    pps = (max(ctx_lens) + ps - 1) // ps
    ld = (nq + 2 * nk) * d
    g = torch.Generator().manual_seed(17)
    perm = torch.randperm(B * pps, generator=g).to(torch.int32).reshape(B, pps)
    kl, vl = gen((B * pps, nk, ps, d), 8, 1.0, torch.bfloat16), gen((B * pps, nk, ps, d), 10, 1.0, torch.bfloat16)
    parts = gen((n_part, B, ld), 3, 0.5)
    bias = gen((ld,), 4, 0.1, torch.bfloat16)
    lens = torch.tensor([c - 1 for c in ctx_lens], dtype=torch.int32)        # tokens cached before the new one
    max_pos = max(ctx_lens)
    half = d // 2
    inv = torch.pow(torch.tensor(theta if theta > 0 else 1.0, dtype=torch.float64), -2.0 * torch.arange(half, dtype=torch.float64) / d)
    ang = torch.arange(max_pos, dtype=torch.float64)[:, None] * inv[None, :]
    table = torch.stack([torch.cos(ang), torch.sin(ang)], dim=-1).float().contiguous()       # [pos][half][2]
    d_parts, d_bias, d_table, d_perm, d_lens = parts.to(cuda), bias.to(cuda), table.to(cuda), perm.to(cuda), lens.to(cuda)
    rope_ptr = d_table.data_ptr() if theta > 0 else None
    bias_ptr = d_bias.data_ptr() if use_bias else None
    slots = B * nq * n_splits

    def scratch():
        return (torch.zeros((slots * d,), dtype=torch.float32, device=cuda), torch.zeros((slots * 2,), dtype=torch.float32, device=cuda))

    # reference path: two kernels
    pages_a = pack_pages(kl, vl).to(cuda)
    q_a = torch.zeros((B, nq * d), dtype=torch.bfloat16, device=cuda)
    out_a = torch.zeros((B, nq * d), dtype=torch.bfloat16, device=cuda)
    po, pml = scratch()
    ok(lib, lib.bfly_k_rope_kv(d_parts.data_ptr(), n_part, B * ld, bias_ptr, rope_ptr, d_lens.data_ptr(), 1, B, nq, nk, d,
                               q_a.data_ptr(), pages_a.data_ptr(), d_perm.data_ptr(), pps, ps, stream()))
    ok(lib, lib.bfly_k_attn_decode(q_a.data_ptr(), pages_a.data_ptr(), d_perm.data_ptr(), pps, ps, nk, d, d_lens.data_ptr(), 1, B,
                                   nq, po.data_ptr(), pml.data_ptr(), n_splits, out_a.data_ptr(), stream()))
    # fused path
    pages_b = pack_pages(kl, vl).to(cuda)
    out_b = torch.zeros((B, nq * d), dtype=torch.bfloat16, device=cuda)
    po, pml = scratch()
    ok(lib, lib.bfly_k_attn_decode_rope(d_parts.data_ptr(), n_part, B * ld, bias_ptr, rope_ptr, pages_b.data_ptr(),
                                        d_perm.data_ptr(), pps, ps, nk, d, d_lens.data_ptr(), B, nq, po.data_ptr(),
                                        pml.data_ptr(), n_splits, out_b.data_ptr(), stream()))
    assert torch.equal(pages_a.view(torch.int16), pages_b.view(torch.int16)), "fused kernel appended other K/V bits"
    assert torch.equal(out_a.view(torch.int16), out_b.view(torch.int16)), "fused kernel differs from rope_kv + attn_decode"
    # oracle: fold, bias, rope, append, attend
    x = parts.sum(0) + (bias.float() if use_bias else 0.0)
    qo = x[:, :nq * d].reshape(B, nq, d)
    ko = x[:, nq * d:(nq + nk) * d].reshape(B, nk, d)
    vo = x[:, (nq + nk) * d:].reshape(B, nk, d)
    if theta > 0:
        qo, ko = ops.rope(qo, lens.long(), theta), ops.rope(ko, lens.long(), theta)
    for b, ctx in enumerate(ctx_lens):
        pos = torch.arange(ctx)
        pg = perm[b, pos // ps].long()
        kk, vv = kl[pg, :, pos % ps, :].float().clone(), vl[pg, :, pos % ps, :].float().clone()   # [ctx, nk, d]
        kk[ctx - 1], vv[ctx - 1] = ko[b].to(torch.bfloat16).float(), vo[b].to(torch.bfloat16).float()
        want = ops.attention(qo[b].to(torch.bfloat16).float().reshape(1, 1, nq, d), kk[None], vv[None], ctx - 1)
        bf16_close(out_b[b], want.reshape(-1), extra_abs=ATTN_ATOL)


def test_attn_prefill_rows(lib, cuda):
    """S new tokens per sequence: token s of sequence b sees lens[b] + s + 1 keys (causal)."""
    d, nq, nk, ps, S = 64, 4, 2, 32, 5
    lens0 = [0, 43]
    B = len(lens0)
    pps = 2
    kl, vl = gen((B * pps, nk, ps, d), 8, 1.0, torch.bfloat16), gen((B * pps, nk, ps, d), 10, 1.0, torch.bfloat16)
    pages = pack_pages(kl, vl)
    table = torch.arange(B * pps, dtype=torch.int32).reshape(B, pps)
    q = gen((B * S, nq * d), 9, 1.0, torch.bfloat16)
    out = torch.zeros((B * S, nq * d), dtype=torch.bfloat16, device=cuda)
    n_splits = 2
    part_o = torch.zeros((B * S * nq * n_splits * d,), dtype=torch.float32, device=cuda)
    part_ml = torch.zeros((B * S * nq * n_splits * 2,), dtype=torch.float32, device=cuda)
    ok(lib, lib.bfly_k_attn_decode(D(q), D(pages), D(table), pps, ps,
                                   nk, d, D(torch.tensor(lens0, dtype=torch.int32)), S, B * S, nq,
                                   part_o.data_ptr(), part_ml.data_ptr(), n_splits, out.data_ptr(), stream()))
    for b in range(B):
        ctx = lens0[b] + S
        pos = torch.arange(ctx)
        pg = table[b, pos // ps].long()
        kk, vv = kl[pg, :, pos % ps, :].float(), vl[pg, :, pos % ps, :].float()
        want = ops.attention(q[b * S:(b + 1) * S].float().reshape(1, S, nq, d), kk[None], vv[None], lens0[b])
        bf16_close(out[b * S:(b + 1) * S], want.reshape(S, nq * d), extra_abs=ATTN_ATOL)


@pytest.mark.parametrize("d,nq,nk,S,lens0", [(128, 8, 2, 70, [0, 45]), (64, 12, 12, 33, [5, 0, 64]), (32, 8, 4, 100, [31]),
                                              (128, 8, 1, 40, [0, 200]), (128, 4, 2, 129, [3])])
def test_attn_prefill_tensor_core_kernel(lib, cuda, d, nq, nk, S, lens0):
    """K7: causal attention of S new tokens over a (possibly non-empty) paged cache."""
    ps, B = 32, len(lens0)
    pps = (max(lens0) + S + ps - 1) // ps
    kl, vl = gen((B * pps, nk, ps, d), 8, 1.0, torch.bfloat16), gen((B * pps, nk, ps, d), 10, 1.0, torch.bfloat16)
    table = torch.randperm(B * pps, generator=torch.Generator().manual_seed(4)).to(torch.int32).reshape(B, pps)
    q = gen((B * S, nq * d), 9, 1.0, torch.bfloat16)
    out = torch.zeros((B * S, nq * d), dtype=torch.bfloat16, device=cuda)
    ok(lib, lib.bfly_k_attn_prefill(D(q), D(pack_pages(kl, vl)), D(table), pps, ps, nk, d,
                                    D(torch.tensor(lens0, dtype=torch.int32)), S, B, nq, out.data_ptr(), stream()))
    for b in range(B):
        pos = torch.arange(lens0[b] + S)
        pg = table[b, pos // ps].long()
        kk, vv = kl[pg, :, pos % ps, :].float(), vl[pg, :, pos % ps, :].float()
        want = ops.attention(q[b * S:(b + 1) * S].float().reshape(1, S, nq, d), kk[None], vv[None], lens0[b])
        bf16_close(out[b * S:(b + 1) * S], want.reshape(S, nq * d), extra_abs=ATTN_ATOL)


@pytest.mark.parametrize("mode,n_blocks,fb,n_part,routed", [(0, 1, 512, 2, False), (1, 1, 3072, 1, False), (0, 4, 256, 3, True)])
def test_act(lib, cuda, mode, n_blocks, fb, n_part, routed):
    T = 5
    f_out = n_blocks * fb
    ld = f_out * (2 if mode == 0 else 1)
    x = gen((n_part, T, ld), 1, 1.5)
    bias = gen((f_out,), 2, 0.1, torch.bfloat16)
    rw = torch.rand((T, n_blocks), generator=torch.Generator().manual_seed(3))
    rw[rw < 0.4] = 0.0
    out = torch.zeros((T, f_out), dtype=torch.bfloat16, device=cuda)
    ok(lib, lib.bfly_k_act(D(x), n_part, T * ld, D(bias) if mode == 1 else None, mode,
                           n_blocks, fb, D(rw) if routed else None, out.data_ptr(), T, stream()))
    s = x.sum(0)
    if mode == 0:
        want = ops.swiglu(s.reshape(T, n_blocks, 2 * fb)).reshape(T, f_out)
    else:
        want = ops.gelu_tanh(s + bias.float())
    if routed:
        want = (want.reshape(T, n_blocks, fb) * rw[:, :, None]).reshape(T, f_out)
    bf16_close(out, want, extra_abs=1e-5)


def test_lm_finalize_and_argmax(lib, cuda):
    T, V, n_part, v0 = 6, 5003, 3, 1000
    parts = gen((n_part, T, V), 1, 0.3)
    parts[:, 2, 77] = 1.0
    parts[:, 2, 4000] = 1.0            # exact tie: lowest index must win
    logits = torch.zeros((T, V + 5), dtype=torch.float32, device=cuda)
    keys = torch.zeros((T,), dtype=torch.int64, device=cuda)
    toks = torch.zeros((T,), dtype=torch.int32, device=cuda)
    ok(lib, lib.bfly_k_lm_finalize(D(parts), n_part, T * V, T, V, v0, logits.data_ptr(), V + 5,
                                   keys.data_ptr(), toks.data_ptr(), stream()))
    want = parts[0] + parts[1] + parts[2]
    assert (logits[:, :V].cpu() - want).abs().max().item() < 1e-6
    assert torch.equal(toks.cpu(), ops.argmax(logits[:, :V].cpu()) + v0)
    assert int(toks[2]) == v0 + 77


def test_moe_router(lib, cuda):
    T, h, E, k = 9, 512, 8, 2
    xn, wr = gen((T, h), 1, 1.0, torch.bfloat16), gen((E, h), 2, 0.05, torch.bfloat16)
    idx, w = ops.moe_router(xn.float(), wr.float(), k)
    for first, n_local in ((0, 8), (4, 4), (6, 2)):
        out = torch.zeros((T, n_local), dtype=torch.float32, device=cuda)
        ok(lib, lib.bfly_k_moe_router(D(xn), D(wr), T, h, E, k, first, n_local,
                                      out.data_ptr(), stream()))
        want = torch.zeros((T, E))
        want.scatter_(1, idx, w)
        assert (out.cpu() - want[:, first:first + n_local]).abs().max().item() < 1e-5
