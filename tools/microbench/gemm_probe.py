#!/usr/bin/env python
"""Probe: weight-streaming GEMM (bfly_k_gemm, M x 32 x 4096) with different ring depths, i.e. 1 vs 2 CTAs per SM."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from butterfly_b200 import _lib  # noqa: E402

lib = _lib.load()
dev = torch.device("cuda:0")
N, K = 32, 4096
MS = [int(x) for x in os.environ.get("PROBE_M", "128256,28672").split(",")]
STAGES = [int(x) for x in os.environ.get("PROBE_STAGES", "3,4,5,6,8").split(",")]
ITERS = int(os.environ.get("PROBE_ITERS", "20"))
for M in MS:
    Ws = [torch.randn((M, K), device=dev, dtype=torch.float32).mul_(0.02).to(torch.bfloat16) for _ in range(2)]
    X = torch.randn((N, K), device=dev, dtype=torch.float32).to(torch.bfloat16)
    out = torch.empty((4, N, M), device=dev, dtype=torch.float32)
    st = torch.cuda.current_stream()
    for stages in STAGES:
        used = C.c_int32(0)

        def launch(i):
            rc = lib.bfly_k_gemm(Ws[i % 2].data_ptr(), X.data_ptr(), M, N, K, K, 1, 32, stages, out.data_ptr(), N * M, None, M,
                                 C.byref(used), st.cuda_stream)
            assert rc == 0, lib.bfly_last_error(None)
        for i in range(3):
            launch(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for i in range(ITERS):
            launch(i)
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / ITERS
        smem = 1024 + stages * (16384 + 32 * 128) + 200
        print(f"M={M} stages={stages} (~{227 * 1024 // smem} CTAs/SM by smem): {ms * 1e3:.1f} us, {2 * M * K / ms / 1e6:.0f} GB/s")
