"""bfly_k_attn_decode_static: the decode-attention entry point for callers that cannot allocate.

Some users of the library cannot hand the attention kernel its split-context scratch: the memory pool
of the application they live in has reserved the whole GPU, or their hot loop must never reach
cudaMalloc/cudaFree (both synchronise the device), or they are a thin binding that only passes the
caller's inputs and outputs through.  For them the scratch is a fixed 4 MiB array inside the library.
This file checks, through the C ABI, that the static variant returns bit for bit what the ordinary
entry point returns (it runs the same kernels on the same inputs; only the owner of the scratch differs),
on calls in a row that share -- and overwrite -- that array; that the result also agrees with the CPU
oracle's attention directly; that no call after the first changes the amount of free device memory (the
point of the entry point: such callers have none to give); and that an oversized request is refused.

Named test_zz_* so that pytest collects it last: a run that stops at the first failure has then
already executed every other test.  It passes alone in a fresh process and after the other tests.
"""
import pytest
import torch

from butterfly_b200 import _lib
from butterfly_b200.kv_layout import pack_pages
from oracle import ops

pytestmark = pytest.mark.gpu

ORACLE_ATOL = 1.6e-2   # vs the fp32 oracle: same bound as the attention tests in test_gpu_kernels.py (bf16 probabilities)


def _case(seed, B, nq, nk, d, ctx_lens, n_splits, dev):
    g = torch.Generator().manual_seed(seed)
    ps = 32
    pps = (max(ctx_lens) + ps - 1) // ps
    k = torch.randn((B * pps, nk, ps, d), generator=g).to(torch.bfloat16)
    v = torch.randn((B * pps, nk, ps, d), generator=g).to(torch.bfloat16)
    table = torch.randperm(B * pps, generator=g).to(torch.int32).reshape(B, pps)
    q = torch.randn((B, nq * d), generator=g).to(torch.bfloat16)
    t = {
        "pages": pack_pages(k, v).to(dev),
        "table": table.to(dev),
        "q": q.to(dev),
        "lens": torch.tensor([c - 1 for c in ctx_lens], dtype=torch.int32).to(dev),
        "host": (k, v, table, q, ctx_lens),
    }
    return t, dict(B=B, nq=nq, nk=nk, d=d, ps=ps, pps=pps, n_splits=n_splits)


def _oracle(t, s):
    """oracle.ops.attention on the logical (unpacked, unpaged) K/V of every sequence."""
    k, v, table, q, ctx_lens = t["host"]
    rows = []
    for b, ctx in enumerate(ctx_lens):
        pos = torch.arange(ctx)
        pg = table[b, pos // s["ps"]].long()
        kk, vv = k[pg, :, pos % s["ps"], :].float(), v[pg, :, pos % s["ps"], :].float()      # [ctx, nk, d]
        rows.append(ops.attention(q[b].float().reshape(1, 1, s["nq"], s["d"]), kk[None], vv[None], ctx - 1).reshape(-1))
    return torch.stack(rows)


def _run_both(lib, t, s, dev):
    st = torch.cuda.current_stream().cuda_stream
    slots = s["B"] * s["nq"] * s["n_splits"]
    part_o = torch.empty((slots * s["d"],), dtype=torch.float32, device=dev)
    part_ml = torch.empty((slots * 2,), dtype=torch.float32, device=dev)
    want = torch.zeros((s["B"], s["nq"] * s["d"]), dtype=torch.bfloat16, device=dev)
    got = torch.zeros_like(want)
    _lib.check(lib.bfly_k_attn_decode(t["q"].data_ptr(), t["pages"].data_ptr(), t["table"].data_ptr(), s["pps"], s["ps"],
                                      s["nk"], s["d"], t["lens"].data_ptr(), 1, s["B"], s["nq"], part_o.data_ptr(),
                                      part_ml.data_ptr(), s["n_splits"], want.data_ptr(), st))
    _lib.check(lib.bfly_k_attn_decode_static(t["q"].data_ptr(), t["pages"].data_ptr(), t["table"].data_ptr(), s["pps"],
                                             s["ps"], s["nk"], s["d"], t["lens"].data_ptr(), 1, s["B"], s["nq"],
                                             s["n_splits"], got.data_ptr(), st))
    torch.cuda.synchronize()
    return want.float().cpu(), got.float().cpu()


def test_static_scratch_variant_matches_on_two_calls_in_a_row(cuda):
    lib = _lib.load()
    # call 1: large splits (fills much of the array); call 2: other shape and sizes, so stale contents
    # of call 1 sit exactly where call 2 works
    lens1 = [700, 33, 512, 1, 257, 64] + [int(x) for x in torch.randint(1, 700, (18,), generator=torch.Generator().manual_seed(9))]
    cases = [_case(1, 24, 32, 8, 128, lens1, 8, cuda),          # 24*32*8*130 floats = 3.2 of the 4 MiB
             _case(2, 3, 4, 4, 64, [90, 300, 31], 3, cuda),
             _case(3, 6, 8, 2, 128, [64, 640, 5, 129, 333, 96], 4, cuda),
             _case(5, 2, 4, 2, 128, [45, 200], 1, cuda)]          # one split: direct output, no scratch at all
    free_after_first = None
    for i, (t, s) in enumerate(cases):
        want, got = _run_both(lib, t, s, cuda)
        assert torch.isfinite(got).all()
        assert torch.equal(want, got), "static-scratch attention differs bitwise from bfly_k_attn_decode"
        assert want.abs().max().item() > 0.05          # the comparison is not vacuous
        err = (got - _oracle(t, s).to(torch.bfloat16).float()).abs().max().item()
        assert err <= ORACLE_ATOL, f"static-scratch attention differs from oracle.ops.attention by {err}"
        free_now = torch.cuda.mem_get_info()[0]        # cudaMemGetInfo
        if i == 0:
            free_after_first = free_now                # the first call may have placed the library's array
    # every later call, through the static entry point alone, leaves the free device memory untouched
    st = torch.cuda.current_stream().cuda_stream
    t, s = cases[1]
    got = torch.zeros((s["B"], s["nq"] * s["d"]), dtype=torch.bfloat16, device=cuda)
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        _lib.check(lib.bfly_k_attn_decode_static(t["q"].data_ptr(), t["pages"].data_ptr(), t["table"].data_ptr(), s["pps"],
                                                 s["ps"], s["nk"], s["d"], t["lens"].data_ptr(), 1, s["B"], s["nq"],
                                                 s["n_splits"], got.data_ptr(), st))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == before, "a call through the static entry point changed free device memory"
    assert free_after_first is not None


def test_single_split_has_no_size_limit(cuda):
    """n_splits == 1 writes its output directly and touches no scratch, so a batch whose split form would not
    fit the array is still served."""
    lib = _lib.load()
    B, nq, nk, d = 130, 64, 8, 128                      # 130*64*1*130 floats > 1,048,576
    t, s = _case(6, B, nq, nk, d, [40 + (i % 23) for i in range(B)], 1, cuda)
    want, got = _run_both(lib, t, s, cuda)
    assert torch.equal(want, got)


def test_oversized_request_is_refused_with_a_message(cuda):
    lib = _lib.load()
    t, s = _case(4, 2, 8, 2, 128, [40, 50], 2, cuda)
    out = torch.zeros((2, 8 * 128), dtype=torch.bfloat16, device=cuda)
    # claim a batch whose scratch need (T*nq*n_splits*(d+2) floats) exceeds the 4 MiB array
    rc = lib.bfly_k_attn_decode_static(t["q"].data_ptr(), t["pages"].data_ptr(), t["table"].data_ptr(), s["pps"], s["ps"],
                                       s["nk"], s["d"], t["lens"].data_ptr(), 1, 4096, 8, 32, out.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream)
    assert rc == 1
    msg = lib.bfly_last_error(None).decode()
    assert "scratch" in msg and "4 MiB" in msg
    torch.cuda.synchronize()
    assert float(out.float().abs().max()) == 0.0        # nothing was launched
