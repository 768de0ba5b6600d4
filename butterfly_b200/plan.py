"""Partition plan: how one transformer block is split over `tp` GPUs of one box.

Stands in for the reference's intended "Model Partitioning" component, which is named
but not implemented (ref:CLAUDE.md:21 "Algorithms to intelligently divide transformer
layers/attention heads"; 0 lines of source).  The split is Megatron-style tensor
parallelism:

  * Q/K/V heads and FFN columns are column-parallel (each rank owns whole heads /
    a contiguous slice of FFN features), so the attention and the activation need no
    communication;
  * attention-out and FFN-down are row-parallel (each rank owns the K-slice matching
    its heads / FFN features) and produce partial sums that are all-reduced;
  * the KV cache is split by KV head; the LM head is split by vocabulary rows and
    all-gathered; the embedding table is replicated;
  * Mixtral experts are placed whole on ranks (E/tp experts per rank); activations are
    TP-replicated, so the expert combine rides on the same all-reduce.

Pure integers, no device, no torch.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

from .config import ModelConfig

DEFAULT_PAGE_SIZE = 32  # tokens per KV page


class PlanError(ValueError):
    pass


@dataclass(frozen=True)
class RankPlan:
    rank: int
    q_heads: Tuple[int, int]      # [begin, end) query heads owned
    kv_heads: Tuple[int, int]     # [begin, end) KV heads owned
    ffn_cols: Tuple[int, int]     # [begin, end) FFN features owned (dense) or full range (MoE)
    vocab_rows: Tuple[int, int]   # [begin, end) LM-head rows owned
    experts: Tuple[int, ...]      # global expert ids owned (MoE), else ()

    @property
    def n_q(self) -> int:
        return self.q_heads[1] - self.q_heads[0]

    @property
    def n_kv(self) -> int:
        return self.kv_heads[1] - self.kv_heads[0]

    @property
    def n_ffn(self) -> int:
        return self.ffn_cols[1] - self.ffn_cols[0]

    @property
    def n_vocab(self) -> int:
        return self.vocab_rows[1] - self.vocab_rows[0]


@dataclass(frozen=True)
class PartitionPlan:
    cfg: ModelConfig
    tp: int
    page_size: int
    vocab_shard: int              # rows per rank before clipping = ceil(V / tp)
    ranks: Tuple[RankPlan, ...]

    def rank(self, r: int) -> RankPlan:
        return self.ranks[r]

    # sizes every rank shares (the split is uniform)
    @property
    def q_local(self) -> int:
        return self.cfg.n_heads // self.tp

    @property
    def kv_local(self) -> int:
        return self.cfg.n_kv_heads // self.tp

    @property
    def qkv_rows(self) -> int:
        return (self.q_local + 2 * self.kv_local) * self.cfg.head_dim

    @property
    def ffn_local(self) -> int:
        return self.cfg.ffn if self.cfg.is_moe else self.cfg.ffn // self.tp

    @property
    def experts_local(self) -> int:
        return self.cfg.n_experts // self.tp if self.cfg.is_moe else 0

    def kv_pages_per_seq(self, max_ctx: int) -> int:
        return (max_ctx + self.page_size - 1) // self.page_size

    def weight_bytes_per_rank(self) -> int:
        """bf16 bytes of layer weights + LM-head shard one rank streams per decode step."""
        c = self.cfg
        d = c.head_dim
        attn = self.qkv_rows * c.hidden + c.hidden * self.q_local * d
        if c.is_moe:
            per_expert = (3 * c.hidden * c.ffn)
            ffn = self.experts_local * per_expert + c.n_experts * c.hidden
        else:
            ffn = (3 if c.gated else 2) * c.hidden * self.ffn_local
        head = self.vocab_shard * c.hidden
        return 2 * (c.n_layers * (attn + ffn) + head)

    def prefill_flops(self, batch: int, seq: int) -> int:
        """Algorithmic flops of one prefill over the whole TP group: 2 * tokens * the
        parameters of the linear layers a token passes through (top-k experts for MoE) + causal attention
        (2 * n_heads * head_dim * S^2 per layer and sequence: QK^T and PV over S^2/2 pairs) + the LM head on the
        last token of every sequence."""
        c = self.cfg
        d = c.head_dim
        attn_lin = (c.n_heads + 2 * c.n_kv_heads) * d * c.hidden + c.hidden * c.n_heads * d
        if c.is_moe:
            ffn = c.top_k * 3 * c.hidden * c.ffn + c.n_experts * c.hidden
        else:
            ffn = (3 if c.gated else 2) * c.hidden * c.ffn
        linear = 2 * batch * seq * c.n_layers * (attn_lin + ffn)
        attention = 2 * c.n_heads * d * seq * seq * c.n_layers * batch
        head = 2 * batch * c.vocab * c.hidden
        return linear + attention + head

    def kv_bytes_per_token_per_rank(self) -> int:
        """bf16 K+V bytes one rank holds per cached token over all layers."""
        return 2 * 2 * self.cfg.n_layers * self.kv_local * self.cfg.head_dim


def make_plan(cfg: ModelConfig, tp: int, page_size: int = DEFAULT_PAGE_SIZE) -> PartitionPlan:
    """Build the tensor-parallel plan for `cfg` on `tp` ranks (tp in {1,2,4,8} on one box)."""
    if tp < 1:
        raise PlanError(f"tp must be >= 1, got {tp}")
    if cfg.n_heads % tp:
        raise PlanError(f"{cfg.name}: n_heads={cfg.n_heads} not divisible by tp={tp}")
    if cfg.n_kv_heads % tp:
        raise PlanError(f"{cfg.name}: n_kv_heads={cfg.n_kv_heads} not divisible by tp={tp}")
    if cfg.n_heads % cfg.n_kv_heads:
        raise PlanError(f"{cfg.name}: n_heads must be a multiple of n_kv_heads")
    if cfg.is_moe:
        if cfg.n_experts % tp:
            raise PlanError(f"{cfg.name}: n_experts={cfg.n_experts} not divisible by tp={tp}")
    elif cfg.ffn % tp:
        raise PlanError(f"{cfg.name}: ffn={cfg.ffn} not divisible by tp={tp}")
    if page_size < 32 or page_size & (page_size - 1):
        raise PlanError("page_size must be a power of two >= 32 (attention streams 32-token tiles)")
    if (cfg.head_dim * 2) % 16 or cfg.hidden % 8:
        raise PlanError("head_dim and hidden must keep rows 16-byte aligned in bf16")

    ql, kl = cfg.n_heads // tp, cfg.n_kv_heads // tp
    vs = (cfg.vocab + tp - 1) // tp
    ranks = []
    for r in range(tp):
        if cfg.is_moe:
            el = cfg.n_experts // tp
            experts = tuple(range(r * el, (r + 1) * el))
            ffn_cols = (0, cfg.ffn)
        else:
            fl = cfg.ffn // tp
            experts = ()
            ffn_cols = (r * fl, (r + 1) * fl)
        ranks.append(RankPlan(
            rank=r,
            q_heads=(r * ql, (r + 1) * ql),
            kv_heads=(r * kl, (r + 1) * kl),
            ffn_cols=ffn_cols,
            vocab_rows=(min(cfg.vocab, r * vs), min(cfg.vocab, (r + 1) * vs)),
            experts=experts))
    return PartitionPlan(cfg=cfg, tp=tp, page_size=page_size, vocab_shard=vs, ranks=tuple(ranks))
