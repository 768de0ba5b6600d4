#!/usr/bin/env python
"""CPU calibration of the bf16 numerics contract at depth (no GPU involved).

    python tools/calibrate_bf16_contract.py [n_layers ...]        (default: 2 8 32; 32 needs ~40 GB of host RAM)

Runs the fp32 oracle and the SAME oracle with activations rounded to bf16 exactly where the B200 path stores
bf16 (GEMM operands, q, KV cache, attention / activation outputs: OracleModel(emulate_bf16=True)) on the
Llama-3-8B shape at the given depths -- identical weights, B=4, prompt 64, 3 decode steps -- and prints the
error of the emulation relative to the oracle's logit sigma.  This is the error ANY faithful bf16-operand
pipeline has on this random-init model; the parity gates for deep models (tests/parity_gates.py) and the
table in BASELINE.md come from its output.
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

from butterfly_b200 import LLAMA3_8B, synthetic_tokens  # noqa: E402
from oracle import OracleModel, ops  # noqa: E402


def run(depth: int):
    from test_gpu_model import full_depth_oracle
    cfg = LLAMA3_8B if depth == LLAMA3_8B.n_layers else LLAMA3_8B.with_layers(depth)
    t0 = time.perf_counter()
    om = full_depth_oracle(cfg)
    ob = OracleModel.__new__(OracleModel)            # second view of the same fp32 weights, bf16 emulation on
    ob.dist, ob.local_ranks, ob.plan, ob.cfg, ob.w, ob.rq = None, [0], om.plan, om.cfg, om.w, ops.bf16_round
    ob.k_cache, ob.v_cache, ob.pos, ob.last_tokens = [], [], 0, None
    ob.reset()
    toks = synthetic_tokens(cfg, 4, 64)
    lo, lb = om.prefill(toks), ob.prefill(toks)
    worst_max = worst_rms = 0.0
    margins = []
    for step in range(4):
        d, s = lb - lo, lo.std().item()
        worst_max = max(worst_max, d.abs().max().item() / s)
        worst_rms = max(worst_rms, d.pow(2).mean().sqrt().item() / s)
        top2 = lo.topk(2, dim=-1).values
        margins += (top2[:, 0] - top2[:, 1]).tolist()
        if step == 3:
            break
        t = ops.argmax(lo)
        lo, lb = om.decode(t), ob.decode(t)
    print(f"L={depth:2d}: bf16-contract emulation vs fp32 oracle: rms {worst_rms * 100:.2f} % of sigma, max "
          f"{worst_max * 100:.2f} % of sigma (sigma {s:.3f}, max abs {worst_max * s:.2e}); oracle top-1/top-2 margin over "
          f"{len(margins)} positions: min {min(margins):.2e} median {sorted(margins)[len(margins) // 2]:.2e}; "
          f"{time.perf_counter() - t0:.0f} s", flush=True)


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    for depth in [int(a) for a in sys.argv[1:]] or [2, 8, 32]:
        run(depth)
