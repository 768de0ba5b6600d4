#!/usr/bin/env python
"""Oracle logit scale and greedy margins per BASELINE.json config (CPU only; fills the table in BASELINE.md
"Parity protocol": the oracle's logit sigma and its top-1/top-2 margin over the decode positions).

    python tools/oracle_margins.py [--full-8b]     (--full-8b: all 32 layers, needs ~40 GB of host RAM, minutes)

Where the fp32 oracle of the full model does not fit host memory (70B: 282 GB, Mixtral: 187 GB) the width is full
and the depth reduced, as in the parity tests; the depth used is printed.
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

from butterfly_b200 import GPT2_SMALL, LLAMA3_8B, LLAMA3_70B, MIXTRAL_8X7B, synthetic_tokens  # noqa: E402
from oracle import ops  # noqa: E402


def run(cfg, B, S, steps):
    from test_gpu_model import full_depth_oracle
    t0 = time.perf_counter()
    om = full_depth_oracle(cfg)
    lo = om.prefill(synthetic_tokens(cfg, B, S))
    margins, sigmas = [], []
    for _ in range(steps):
        top2 = lo.topk(2, dim=-1).values
        margins += (top2[:, 0] - top2[:, 1]).tolist()
        sigmas.append(lo.std().item())
        lo = om.decode(ops.argmax(lo))
    m = sorted(margins)
    below = sum(1 for x in m if x <= 2e-2) / len(m)
    print(f"| {cfg.name} | {cfg.n_layers} | {B} x {steps} | {sum(sigmas) / len(sigmas):.3f} | {m[0]:.2e} | "
          f"{m[len(m) // 2]:.2e} | {below * 100:.0f} % | ({time.perf_counter() - t0:.0f} s)", flush=True)


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    print("| config | layers | positions (B x steps) | logit sigma | min margin | median margin | margin <= 2e-2 |")
    run(GPT2_SMALL, 2, 64, 32)
    run(LLAMA3_8B if "--full-8b" in sys.argv else LLAMA3_8B.with_layers(2), 4, 64, 32)
    run(LLAMA3_70B.with_layers(1), 4, 33, 32)
    run(MIXTRAL_8X7B.with_layers(1), 4, 20, 32)
