#!/usr/bin/env python
"""Micro-benchmarks of the decode-step kernels on one B200, through the C ABI (run under gpurun).

    python tools/tune_kernels.py gemm [tp] [model]   # split-K / stage sweep for the decode GEMM shapes of one rank
    python tools/tune_kernels.py attn [tp] [model]   # context-split sweep for paged decode attention
(model: llama3-8b (default, B=32 ctx 2048) or llama3-70b (B=64 ctx 4096))

Weights / KV buffers are rotated so that no launch re-reads L2-resident data; times are CUDA-event
averages over the rotation.  Prints GB/s of algorithmic bytes against MEASURED_PEAKS.json.
"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from butterfly_b200 import _lib  # noqa: E402
from butterfly_b200.kv_layout import pack_pages  # noqa: E402


def peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return json.load(open(p))["hbm_gbs"] if os.path.exists(p) else 6650.0


def time_launches(fn, n, warm=3):
    st = torch.cuda.current_stream()
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(n):
        fn(i)
    e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


SHAPES = {  # h, f, nq, nk, d, V, batch, ctx
    "llama3-8b": (4096, 14336, 32, 8, 128, 128256, 32, 2048),
    "llama3-70b": (8192, 28672, 64, 8, 128, 128256, 64, 4096),
}


def tune_gemm(tp=1, model="llama3-8b"):
    lib = _lib.load()
    h, f, nq, nk, d, V, batch, _ = SHAPES[model]
    shapes = {"qkv": ((nq + 2 * nk) * d // tp, h), "wo": (h, nq * d // tp), "gate_up": (2 * f // tp, h),
              "down": (h, f // tp), "lm_head": (V // tp, h)}
    N = batch
    pk = peak()
    for name, (M, K) in shapes.items():
        n_buf = max(2, int(300e6 // (M * K * 2)) + 1)
        n_buf = min(n_buf, 12)
        Ws = [torch.randn((M, K), device="cuda").mul_(0.02).to(torch.bfloat16) for _ in range(n_buf)]
        X = torch.randn((N, K), device="cuda").to(torch.bfloat16)
        out = torch.empty((16, N, M), device="cuda", dtype=torch.float32)
        used = C.c_int32(0)
        ps, pbn, pst = C.c_int32(0), C.c_int32(0), C.c_int32(0)
        lib.bfly_k_gemm_plan(M, N, K, C.byref(ps), C.byref(pbn), C.byref(pst))
        print(f"== {name}: M={M} N={N} K={K} ({M * K * 2 / 1e6:.1f} MB)  planner: splits={ps.value} stages={pst.value}")
        best = None
        nkb = (K + 63) // 64
        for splits in (1, 2, 3, 4, 5, 6, 8, 9, 12, 16):
            if splits > nkb // 2:
                continue
            row = []
            for stages in (3, 4, 6, 8):
                def fn(i):
                    rc = lib.bfly_k_gemm(Ws[i % n_buf].data_ptr(), X.data_ptr(), M, N, K, K, splits, 0, stages,
                                         out.data_ptr(), N * M, None, M, C.byref(used),
                                         torch.cuda.current_stream().cuda_stream)
                    assert rc == 0, lib.bfly_last_error(None)
                ms = time_launches(fn, 3 * n_buf)
                gbs = (M * K * 2 + N * K * 2) / ms / 1e6
                row.append(f"st{stages}:{ms * 1e3:6.1f}us {gbs / pk * 100:4.0f}%")
                if best is None or ms < best[0]:
                    best = (ms, splits, stages)
            print(f"  splits={splits:2d}  " + "  ".join(row))
        print(f"  best: {best[0] * 1e3:.1f} us  splits={best[1]} stages={best[2]}")
        del Ws


def tune_swiglu(tp=1, batch=32):
    """gate/up projection with the fused SwiGLU epilogue: cluster split sweep."""
    lib = _lib.load()
    h, f = 4096, 14336 // tp
    n_buf = 3 if tp == 1 else 8
    Ws = [torch.randn((2 * f, h), device="cuda").mul_(0.02).to(torch.bfloat16) for _ in range(n_buf)]
    X = torch.randn((batch, h), device="cuda").to(torch.bfloat16)
    out = torch.empty((batch, f), device="cuda", dtype=torch.bfloat16)
    pk = peak()
    print(f"== fused gate/up+SwiGLU: f={f} N={batch} K={h} ({2 * f * h * 2 / 1e6:.1f} MB)")
    for splits in (0, 1, 2, 3, 4, 5, 6, 8):
        def fn(i):
            rc = lib.bfly_k_gemm_swiglu(Ws[i % n_buf].data_ptr(), X.data_ptr(), batch, h, 1, f, None, splits,
                                        out.data_ptr(), torch.cuda.current_stream().cuda_stream)
            assert rc == 0, lib.bfly_last_error(None)
        ms = time_launches(fn, 4 * n_buf)
        print(f"  splits={splits} (0=planner): {ms * 1e3:6.1f} us  {2 * f * h * 2 / ms / 1e6 / pk * 100:4.0f}%")


def tune_attn(tp=1, model="llama3-8b"):
    lib = _lib.load()
    _, _, nq_all, nk_all, d, _, batch, ctx = SHAPES[model]
    nq, nk, ps = nq_all // tp, nk_all // tp, 32
    pps = (ctx + ps - 1) // ps
    n_pages = batch * pps
    n_buf = 3
    pages = [torch.randn((n_pages, 2, nk, ps * d), device="cuda").to(torch.bfloat16) for _ in range(n_buf)]
    table = torch.arange(n_pages, dtype=torch.int32, device="cuda").reshape(batch, pps)
    q = torch.randn((batch, nq * d), device="cuda").to(torch.bfloat16)
    lens = torch.full((batch,), ctx - 1, dtype=torch.int32, device="cuda")
    out = torch.empty((batch, nq * d), device="cuda", dtype=torch.bfloat16)
    pk = peak()
    kv_bytes = batch * ctx * nk * d * 2 * 2
    print(f"== attn decode: B={batch} ctx={ctx} nq={nq} nk={nk} d={d}  KV bytes {kv_bytes / 1e6:.1f} MB")
    for n_splits in (1, 2, 3, 4, 5, 6, 7, 8, 9, 12, 16):
        part_o = torch.empty((batch * nq * n_splits * d,), device="cuda", dtype=torch.float32)
        part_ml = torch.empty((batch * nq * n_splits * 2,), device="cuda", dtype=torch.float32)

        def fn(i):
            rc = lib.bfly_k_attn_decode(q.data_ptr(), pages[i % n_buf].data_ptr(), table.data_ptr(), pps, ps, nk, d,
                                        lens.data_ptr(), 1, batch, nq, part_o.data_ptr(), part_ml.data_ptr(),
                                        n_splits, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
            assert rc == 0, lib.bfly_last_error(None)
        ms = time_launches(fn, 12)
        print(f"  n_splits={n_splits:2d}: {ms * 1e3:7.1f} us  {kv_bytes / ms / 1e6:7.0f} GB/s  {kv_bytes / ms / 1e6 / pk * 100:4.0f}% "
              f"(attention + combine kernels)")


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "gemm"
    tp = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    model = sys.argv[3] if len(sys.argv) > 3 else "llama3-8b"
    if what == "gemm":
        tune_gemm(tp, model)
    elif what == "attn":
        tune_attn(tp, model)
    elif what == "swiglu":
        tune_swiglu(tp)
    else:
        tune_gemm(tp, model)
        tune_attn(tp, model)
