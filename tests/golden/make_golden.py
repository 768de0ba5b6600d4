#!/usr/bin/env python
"""Generates tests/golden/oracle_tiny.npz from the oracle (run once per deliberate oracle change):
    python tests/golden/make_golden.py
The reference has no golden vectors (it has no code); these fixtures freeze the stand-in oracle."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from butterfly_b200 import SyntheticSource, make_plan, shard_weights, synthetic_tokens, tiny  # noqa: E402
from oracle import OracleModel  # noqa: E402

out = {}
for family in ("llama", "gpt2", "mixtral"):
    cfg = tiny(family)
    src = SyntheticSource(cfg, 0)
    plan = make_plan(cfg, 1)
    m = OracleModel(plan, [shard_weights(src, plan, 0)])
    toks = synthetic_tokens(cfg, 2, 8)
    out[f"{family}_tokens"] = toks.numpy()
    out[f"{family}_logits"] = m.prefill(toks).numpy()
    out[f"{family}_greedy"] = m.decode_greedy(6).numpy()
out["kat_wq_bits"] = SyntheticSource(tiny("llama"), 0).block("wq", 0, 0, 2, 0, 8).view(torch.int16).numpy()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle_tiny.npz"), **out)
print({k: v.shape for k, v in out.items()})
