"""bfly_k_attn_decode_static: the decode-attention entry point for callers that cannot allocate.

Some users of the library cannot hand the attention kernel its split-context scratch: the memory pool
of the application they live in has reserved the whole GPU, or their hot loop must never reach
cudaMalloc/cudaFree (both synchronise the device), or they are a thin binding that only passes the
caller's inputs and outputs through.  For them the scratch is a fixed 4 MiB array inside the library.
This file checks, through the C ABI, that the static variant returns what the ordinary entry point
returns (the ordinary one is itself checked against the oracle in test_gpu_kernels.py), on two calls in
a row that share -- and overwrite -- that array, and that an oversized request is refused.

Named test_zz_* so that pytest collects it last: a run that stops at the first failure has then
already executed every other test.  It passes alone in a fresh process and after the other tests.
"""
import pytest
import torch

from butterfly_b200 import _lib
from butterfly_b200.kv_layout import pack_pages

pytestmark = pytest.mark.gpu

ATOL = 1.6e-2          # same bound as the attention tests in test_gpu_kernels.py (bf16 probabilities)


def _case(seed, B, nq, nk, d, ctx_lens, n_splits, dev):
    g = torch.Generator().manual_seed(seed)
    ps = 32
    pps = (max(ctx_lens) + ps - 1) // ps
    k = torch.randn((B * pps, nk, ps, d), generator=g).to(torch.bfloat16)
    v = torch.randn((B * pps, nk, ps, d), generator=g).to(torch.bfloat16)
    t = {
        "pages": pack_pages(k, v).to(dev),
        "table": torch.randperm(B * pps, generator=g).to(torch.int32).reshape(B, pps).to(dev),
        "q": torch.randn((B, nq * d), generator=g).to(torch.bfloat16).to(dev),
        "lens": torch.tensor([c - 1 for c in ctx_lens], dtype=torch.int32).to(dev),
    }
    return t, dict(B=B, nq=nq, nk=nk, d=d, ps=ps, pps=pps, n_splits=n_splits)


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
             _case(3, 6, 8, 2, 128, [64, 640, 5, 129, 333, 96], 4, cuda)]
    for t, s in cases:
        want, got = _run_both(lib, t, s, cuda)
        assert torch.isfinite(got).all()
        err = (want - got).abs().max().item()
        assert err <= ATOL, f"static-scratch attention differs from bfly_k_attn_decode by {err}"
        assert want.abs().max().item() > 0.05          # the comparison is not vacuous


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
