#!/usr/bin/env python
"""Per-phase times of the fused decode step kernel (csrc/decode_step.cu), from its in-kernel trace.

    BFLY_STEP_TRACE=1 python tools/step_trace.py [--shard-of N] [--model llama3-8b] [--layers L]

Prints, per phase kind, the average time between the completion of the previous phase and of this one as
seen by CTA 0 (averaged over layers of the last of a few decode steps)."""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["BFLY_STEP_TRACE"] = "1"

import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shard-of", type=int, default=1)
    ap.add_argument("--model", default="llama3-8b")
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--ctx", type=int, default=2048)
    args = ap.parse_args()
    if args.shard_of > 1:
        os.environ["BFLY_SKIP_COLLECTIVES"] = "1"
    from butterfly_b200 import PRESETS, Model, SyntheticSource, make_plan, synthetic_tokens
    cfg = PRESETS[args.model]
    if args.layers:
        cfg = cfg.with_layers(args.layers)
    m = Model(make_plan(cfg, args.shard_of), 0, SyntheticSource(cfg, 0), max_batch=args.batch, max_ctx=args.ctx + 64)
    m.prefill(synthetic_tokens(cfg, args.batch, args.ctx), want_logits=False)
    m.decode_enqueue(args.batch, 8)
    m.sync()
    buf = np.zeros(1 << 20, dtype=np.uint64)
    n, splits = C.c_int32(0), C.c_int32(0)
    m._check(m.lib.bfly_read_step_trace(m.ctx, buf.ctypes.data, buf.size, C.byref(n), C.byref(splits)))
    grid = m.stats()["sm_count"]
    n_ph = (n.value - 2) // (1 + grid)
    arrivals = buf[n_ph + 2:n_ph + 2 + n_ph * grid].astype(np.int64).reshape(n_ph, grid)
    n.value = n_ph + 2
    t = buf[:n.value].astype(np.int64)
    names = ["norm_attn", "qkv", "rope", "attn"] + (["combine"] if splits.value > 1 else []) + ["wo", "norm_ffn", "gate_up", "act", "down"]
    per = len(names)
    d = np.diff(t)                                  # d[k] = duration of phase k
    L = cfg.n_layers
    print(f"{cfg.name} tp-shard {args.shard_of}: B={args.batch} ctx={args.ctx} attn_splits={splits.value}; step total {(t[n.value - 1] - t[0]) / 1e3:.1f} us")
    for i, nm in enumerate(names):
        v = d[i:L * per:per] / 1e3
        print(f"  {nm:10s} avg {v.mean():7.2f} us  (min {v.min():.2f}, max {v.max():.2f})")
    tail = d[L * per:] / 1e3
    print("  tail (final norm, lm head, finalize):", np.round(tail, 2).tolist())
    print(f"  per layer {d[:L * per].sum() / L / 1e3:.2f} us")
    # when did each CTA finish its share of a phase, relative to the phase's start (layer 1 and the LM head)?
    for k, nm in [(per + i, names[i]) for i in range(per)] + [(L * per + 1, "lm_head")]:
        a = (arrivals[k] - t[k]) / 1e3
        print(f"  CTA arrival spread {nm:10s}: min {a.min():7.2f} p25 {np.percentile(a, 25):7.2f} median {np.median(a):7.2f} "
              f"p75 {np.percentile(a, 75):7.2f} max {a.max():7.2f} us")
    m.close()


if __name__ == "__main__":
    main()
