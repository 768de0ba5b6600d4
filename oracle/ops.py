"""CPU fp32 oracle -- operator level.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package; butterfly_b200/ never does.

PARITY UNPINNED BY THE REFERENCE: TensorHusker/Butterfly contains no implementation, tests or
golden vectors for this path (ref:CLAUDE.md:13-15 "newly initialized repository ... To Be
Implemented"), so there is no reference file:line any function here could
follow.  Each function instead restates the published architecture of the model family named in
BASELINE.json `configs`, and is pinned by an independent cross-check against HuggingFace
`transformers` 5.5.0 (LlamaForCausalLM / GPT2LMHeadModel / MixtralForCausalLM built offline from
config with the same weights): tests/test_oracle_hf.py.  Rows refer to SURVEY.md section 8a.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import torch


def embed(tokens: torch.Tensor, table: torch.Tensor, pos_table: Optional[torch.Tensor] = None,
          positions: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Row a3: x[t] = table[tokens[t]] (+ pos_table[positions[t]] for learned positions, GPT-2)."""
    x = table[tokens.long()]
    if pos_table is not None:
        x = x + pos_table[positions.long()]
    return x


def norm(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], eps: float, layer_norm: bool) -> torch.Tensor:
    """Row a4: RMSNorm (Llama/Mixtral: x * rsqrt(mean(x^2)+eps) * w) or LayerNorm (GPT-2)."""
    if layer_norm:
        mu = x.mean(-1, keepdim=True)
        var = ((x - mu) ** 2).mean(-1, keepdim=True)
        y = (x - mu) * torch.rsqrt(var + eps) * w
        return y + b if b is not None else y
    return x * torch.rsqrt((x * x).mean(-1, keepdim=True) + eps) * w


def linear(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Rows a5/a9/a11/a12/a14: y = x @ w^T (+ b); w is [out, in]."""
    y = x @ w.t()
    return y + b if b is not None else y


def rope(x: torch.Tensor, positions: torch.Tensor, theta: float) -> torch.Tensor:
    """Row a6: half-split rotary embedding (HF `rotate_half` layout): pairs (i, i + d/2).

    x: [T, heads, d]; positions: [T]."""
    d = x.shape[-1]
    half = d // 2
    inv_freq = theta ** (-torch.arange(0, half, dtype=torch.float64) * 2.0 / d)
    ang = positions.to(torch.float64)[:, None] * inv_freq[None, :]
    cos, sin = ang.cos().float()[:, None, :], ang.sin().float()[:, None, :]
    x1, x2 = x[..., :half], x[..., half:]
    return torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], dim=-1)


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, q_pos0: int) -> torch.Tensor:
    """Rows a7/a8: causal softmax(q k^T / sqrt(d)) v with GQA.

    q: [B, Sq, nq, d]; k, v: [B, Sk, nkv, d] (the cache up to and including the new tokens);
    query i sits at absolute position q_pos0 + i and sees keys 0..q_pos0+i."""
    B, Sq, nq, d = q.shape
    Sk, nkv = k.shape[1], k.shape[2]
    g = nq // nkv
    qh = q.permute(0, 2, 1, 3).reshape(B, nkv, g, Sq, d)
    kh = k.permute(0, 2, 1, 3)[:, :, None]            # [B, nkv, 1, Sk, d]
    vh = v.permute(0, 2, 1, 3)[:, :, None]
    s = (qh @ kh.transpose(-1, -2)) * (1.0 / math.sqrt(d))   # [B, nkv, g, Sq, Sk]
    qi = torch.arange(Sq)[:, None] + q_pos0
    ki = torch.arange(Sk)[None, :]
    s = s.masked_fill(ki > qi, float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = p @ vh                                           # [B, nkv, g, Sq, d]
    return o.reshape(B, nq, Sq, d).permute(0, 2, 1, 3).reshape(B, Sq, nq * d)


def swiglu(gate_up: torch.Tensor) -> torch.Tensor:
    """Row a11: silu(gate) * up; gate_up = [..., gate(f) | up(f)]."""
    f = gate_up.shape[-1] // 2
    g, u = gate_up[..., :f], gate_up[..., f:]
    return g * torch.sigmoid(g) * u


def gelu_tanh(x: torch.Tensor) -> torch.Tensor:
    """Row a11 (GPT-2): 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))  (HF `gelu_new`)."""
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x ** 3)))


def moe_router(x: torch.Tensor, router_w: torch.Tensor, top_k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Row a13: softmax over experts -> top-k -> renormalise.  Returns (idx[T,k], weight[T,k])."""
    probs = torch.softmax(x @ router_w.t(), dim=-1)
    w, idx = torch.topk(probs, top_k, dim=-1)
    return idx, w / w.sum(-1, keepdim=True)


def all_reduce_sum(parts: Sequence[torch.Tensor]) -> torch.Tensor:
    """Row a10: sum over ranks, in rank order."""
    acc = parts[0].clone()
    for p in parts[1:]:
        acc += p
    return acc


def all_gather(parts: Sequence[torch.Tensor]) -> torch.Tensor:
    """Row a14: concatenate vocabulary shards in rank order."""
    return torch.cat(list(parts), dim=-1)


def argmax(logits: torch.Tensor) -> torch.Tensor:
    """Row a14: greedy token; ties resolve to the lowest index (matches the CUDA kernel)."""
    m = logits.max(dim=-1, keepdim=True).values
    V = logits.shape[-1]
    idx = torch.where(logits == m, torch.arange(V)[None, :], torch.full((1, 1), V))
    return idx.min(dim=-1).values.to(torch.int32)


def bf16_round(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32)
