"""CPU fp32 oracle -- the partitioned block forward with tensor parallelism simulated in-process.
TEST INFRASTRUCTURE, NOT PRODUCT CODE (see oracle/ops.py header; parity unpinned by the
reference, pinned against HuggingFace transformers in tests/test_oracle_hf.py).

Same API as butterfly_b200.Model: OracleModel(plan, [RankWeights per rank]).prefill(tokens) /
.decode(tokens) / .decode_greedy(n).  Every rank's work is done in a Python loop; the all-reduce
is a sum over ranks and the all-gather a concatenation.

`emulate_bf16=True` additionally rounds activations to bf16 at exactly the points where the B200
numerics contract stores bf16 (GEMM operands, q, KV cache, attention/activation outputs).  It is a
diagnostic used to calibrate the expected error of the contract on CPU; the parity gate compares
against the plain fp32 mode.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from butterfly_b200.plan import PartitionPlan
from butterfly_b200.weights import RankWeights

from . import ops


class OracleModel:
    def __init__(self, plan: PartitionPlan, rank_weights: List[RankWeights], emulate_bf16: bool = False,
                 dist_group=None):
        """All ranks in-process (default), or -- with `dist_group` (a torch.distributed group of
        plan.tp processes, e.g. gloo) -- only this process's rank, with real collectives."""
        self.dist = None
        if dist_group is not None:
            import torch.distributed as dist
            self.dist, self.group = dist, (None if dist_group is True else dist_group)
            assert len(rank_weights) == 1 and dist.get_world_size(self.group) == plan.tp
            self.local_ranks = [rank_weights[0].rank]
        else:
            assert len(rank_weights) == plan.tp
            self.local_ranks = list(range(plan.tp))
        self.plan, self.cfg = plan, plan.cfg
        self.w = {rw.rank: {k: v.float() for k, v in rw.tensors.items()} for rw in rank_weights}
        self.rq = ops.bf16_round if emulate_bf16 else (lambda t: t)
        self.k_cache: List[List[Optional[torch.Tensor]]] = []
        self.v_cache: List[List[Optional[torch.Tensor]]] = []
        self.pos = 0
        self.last_tokens: Optional[torch.Tensor] = None
        self.reset()

    def reset(self):
        L, tp = self.cfg.n_layers, self.plan.tp
        self.k_cache = [[None] * tp for _ in range(L)]
        self.v_cache = [[None] * tp for _ in range(L)]
        self.pos = 0

    def _append_kv(self, l, r, k, v):
        """Append S new positions to the cache of (layer, rank) and return views of everything cached so far.
        The cache grows geometrically, so a decode step copies only its own token, not the history."""
        B, S = k.shape[0], k.shape[1]
        buf_k, buf_v = self.k_cache[l][r], self.v_cache[l][r]
        need = self.pos + S
        if buf_k is None or buf_k.shape[1] < need:
            cap = max(need, 2 * (buf_k.shape[1] if buf_k is not None else 0), 64)
            nk_, nv_ = k.new_empty((B, cap) + k.shape[2:]), v.new_empty((B, cap) + v.shape[2:])
            if buf_k is not None and self.pos > 0:
                nk_[:, :self.pos], nv_[:, :self.pos] = buf_k[:, :self.pos], buf_v[:, :self.pos]
            buf_k, buf_v = nk_, nv_
            self.k_cache[l][r], self.v_cache[l][r] = buf_k, buf_v
        buf_k[:, self.pos:need], buf_v[:, self.pos:need] = k, v
        return buf_k[:, :need], buf_v[:, :need]

    def _all_reduce(self, parts):
        acc = ops.all_reduce_sum(parts)
        if self.dist is not None:
            self.dist.all_reduce(acc, group=self.group)
        return acc

    def _all_gather(self, parts):
        if self.dist is None:
            return ops.all_gather(parts)
        local = torch.zeros((parts[0].shape[0], self.plan.vocab_shard))
        local[:, :parts[0].shape[1]] = parts[0]
        bufs = [torch.empty_like(local) for _ in range(self.plan.tp)]
        self.dist.all_gather(bufs, local, group=self.group)
        return torch.cat([bufs[r][:, :self.plan.rank(r).n_vocab] for r in range(self.plan.tp)], dim=-1)

    # -- one forward over S new tokens per sequence ------------------------------------------
    def _forward(self, tokens: torch.Tensor) -> torch.Tensor:
        cfg, plan, rq = self.cfg, self.plan, self.rq
        B, S = tokens.shape
        d = cfg.head_dim
        positions = (torch.arange(S) + self.pos).repeat(B)            # [B*S]
        w0 = self.w[self.local_ranks[0]]
        x = ops.embed(tokens.reshape(-1), w0["embed"], w0.get("pos_embed"), positions)   # [T, h] fp32
        for l in range(cfg.n_layers):
            p = f"layers.{l}."
            hn = rq(ops.norm(x, w0[p + "attn_norm.w"], w0.get(p + "attn_norm.b"), cfg.norm_eps, cfg.layer_norm))
            parts = []
            for r in self.local_ranks:
                w = self.w[r]
                nq, nk = plan.q_local, plan.kv_local
                qkv = ops.linear(hn, w[p + "wqkv"], w.get(p + "bqkv"))
                q = qkv[:, :nq * d].reshape(B * S, nq, d)
                k = qkv[:, nq * d:(nq + nk) * d].reshape(B * S, nk, d)
                v = qkv[:, (nq + nk) * d:].reshape(B * S, nk, d)
                if cfg.rope_theta > 0.0:
                    q, k = ops.rope(q, positions, cfg.rope_theta), ops.rope(k, positions, cfg.rope_theta)
                q, k, v = rq(q), rq(k), rq(v)
                k, v = k.reshape(B, S, nk, d), v.reshape(B, S, nk, d)
                k, v = self._append_kv(l, r, k, v)
                a = rq(ops.attention(q.reshape(B, S, nq, d), k, v, self.pos)).reshape(B * S, nq * d)
                parts.append(ops.linear(a, w[p + "wo"]))
            x = x + self._all_reduce(parts)
            if cfg.bias:
                x = x + w0[p + "bo"]
            hn = rq(ops.norm(x, w0[p + "ffn_norm.w"], w0.get(p + "ffn_norm.b"), cfg.norm_eps, cfg.layer_norm))
            parts = []
            if cfg.is_moe:
                idx, gate_w = ops.moe_router(hn, w0[p + "router"], cfg.top_k)
                for r in self.local_ranks:
                    w = self.w[r]
                    acc = torch.zeros_like(x)
                    f = cfg.ffn
                    for j, e in enumerate(plan.rank(r).experts):
                        sel = (idx == e)
                        rows = sel.any(-1).nonzero().squeeze(-1)
                        if rows.numel() == 0:
                            continue
                        w_gu = w[p + "w_gate_up"][j * 2 * f:(j + 1) * 2 * f]
                        w_dn = w[p + "w_down"][:, j * f:(j + 1) * f]
                        gw = (gate_w[rows] * sel[rows]).sum(-1, keepdim=True)
                        act = rq(ops.swiglu(ops.linear(hn[rows], w_gu)) * gw)
                        acc[rows] += ops.linear(act, w_dn)
                    parts.append(acc)
            else:
                for r in self.local_ranks:
                    w = self.w[r]
                    if cfg.gated:
                        act = ops.swiglu(ops.linear(hn, w[p + "w_gate_up"]))
                    else:
                        act = ops.gelu_tanh(ops.linear(hn, w[p + "w_up"], w[p + "b_up"]))
                    parts.append(ops.linear(rq(act), w[p + "w_down"]))
            x = x + self._all_reduce(parts)
            if cfg.bias:
                x = x + w0[p + "b_down"]
        self.pos += S
        last = x.reshape(B, S, -1)[:, -1]
        hn = rq(ops.norm(last, w0["final_norm.w"], w0.get("final_norm.b"), cfg.norm_eps, cfg.layer_norm))
        return self._all_gather([ops.linear(hn, self.w[r]["lm_head"]) for r in self.local_ranks])   # [B, V]

    # -- public API ---------------------------------------------------------------------------
    @torch.no_grad()
    def prefill(self, tokens: torch.Tensor) -> torch.Tensor:
        """tokens int32[B,S] -> fp32 logits[B,V] of the last position; (re)starts the KV cache."""
        self.reset()
        logits = self._forward(tokens)
        self.last_tokens = ops.argmax(logits)
        return logits

    @torch.no_grad()
    def decode(self, tokens: torch.Tensor) -> torch.Tensor:
        """tokens int32[B] (one new token per sequence) -> fp32 logits[B,V]."""
        logits = self._forward(tokens.reshape(-1, 1))
        self.last_tokens = ops.argmax(logits)
        return logits

    @torch.no_grad()
    def decode_greedy(self, n_steps: int) -> torch.Tensor:
        """Free-running greedy decode from the last logits: int32[B, n_steps]."""
        out = []
        for _ in range(n_steps):
            tok = self.last_tokens
            out.append(tok)
            self.decode(tok)
        return torch.stack(out, dim=1)
