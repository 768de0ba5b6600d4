"""Model shapes served by the partitioned block forward.

The reference repository (TensorHusker/Butterfly) ships no source, so no shape is
specified there (ref:CLAUDE.md:13-23 only lists intended components).  The four
shapes below are the public architectures named by BASELINE.json `configs`
(shape table).  Everything is a frozen dataclass; there is
no YAML / env-flag layer on purpose.
"""
from __future__ import annotations

from dataclasses import dataclass, replace

FAMILY_GPT2 = "gpt2"
FAMILY_LLAMA = "llama"
FAMILY_MIXTRAL = "mixtral"


@dataclass(frozen=True)
class ModelConfig:
    name: str
    family: str            # gpt2 | llama | mixtral
    hidden: int            # h
    n_layers: int          # L
    n_heads: int           # nh  (query heads)
    n_kv_heads: int        # nkv (GQA groups; == nh for MHA)
    head_dim: int          # d
    ffn: int               # f
    vocab: int             # V
    n_experts: int = 0     # E  (0 = dense FFN)
    top_k: int = 0
    norm_eps: float = 1e-5
    rope_theta: float = 0.0   # 0 => learned absolute positions (GPT-2)
    max_pos: int = 8192
    tied_head: bool = False
    bias: bool = False     # GPT-2: biases on every linear + LayerNorm beta

    # ---- derived -----------------------------------------------------
    @property
    def q_dim(self) -> int:
        return self.n_heads * self.head_dim

    @property
    def kv_dim(self) -> int:
        return self.n_kv_heads * self.head_dim

    @property
    def is_moe(self) -> bool:
        return self.n_experts > 0

    @property
    def gated(self) -> bool:
        """SwiGLU (gate+up) vs a single up-projection with GELU-tanh."""
        return self.family != FAMILY_GPT2

    @property
    def layer_norm(self) -> bool:
        return self.family == FAMILY_GPT2

    def n_params(self) -> int:
        h, f = self.hidden, self.ffn
        attn = h * (self.q_dim + 2 * self.kv_dim) + self.q_dim * h
        ffn1 = (3 if self.gated else 2) * h * f
        ffn = ffn1 * max(1, self.n_experts) + (self.n_experts * h if self.is_moe else 0)
        norms = 2 * h * (2 if self.layer_norm else 1)
        per_layer = attn + ffn + norms
        if self.bias:
            per_layer += (self.q_dim + 2 * self.kv_dim) + h + f + h
        emb = self.vocab * h + (self.max_pos * h if self.rope_theta == 0.0 else 0)
        head = 0 if self.tied_head else self.vocab * h
        return self.n_layers * per_layer + emb + head + h * (2 if self.layer_norm else 1)

    def with_layers(self, n_layers: int) -> "ModelConfig":
        """Same width, reduced depth (parity runs whose fp32 oracle would not fit host RAM)."""
        return replace(self, n_layers=n_layers, name=f"{self.name}-L{n_layers}")


GPT2_SMALL = ModelConfig(
    name="gpt2-small", family=FAMILY_GPT2, hidden=768, n_layers=12, n_heads=12, n_kv_heads=12,
    head_dim=64, ffn=3072, vocab=50257, norm_eps=1e-5, rope_theta=0.0, max_pos=1024,
    tied_head=True, bias=True)

LLAMA3_8B = ModelConfig(
    name="llama3-8b", family=FAMILY_LLAMA, hidden=4096, n_layers=32, n_heads=32, n_kv_heads=8,
    head_dim=128, ffn=14336, vocab=128256, norm_eps=1e-5, rope_theta=500000.0, max_pos=8192)

LLAMA3_70B = ModelConfig(
    name="llama3-70b", family=FAMILY_LLAMA, hidden=8192, n_layers=80, n_heads=64, n_kv_heads=8,
    head_dim=128, ffn=28672, vocab=128256, norm_eps=1e-5, rope_theta=500000.0, max_pos=8192)

MIXTRAL_8X7B = ModelConfig(
    name="mixtral-8x7b", family=FAMILY_MIXTRAL, hidden=4096, n_layers=32, n_heads=32, n_kv_heads=8,
    head_dim=128, ffn=14336, vocab=32000, n_experts=8, top_k=2, norm_eps=1e-5,
    rope_theta=1000000.0, max_pos=32768)

PRESETS = {c.name: c for c in (GPT2_SMALL, LLAMA3_8B, LLAMA3_70B, MIXTRAL_8X7B)}


def tiny(family: str, *, hidden: int = 256, n_layers: int = 2, n_heads: int = 8, n_kv_heads: int | None = None,
         head_dim: int = 32, ffn: int = 512, vocab: int = 1000, n_experts: int = 8, max_pos: int = 512) -> ModelConfig:
    """Small same-architecture shapes for CPU tests and the GPU smoke run."""
    if family == FAMILY_GPT2:
        return ModelConfig(name="gpt2-tiny", family=family, hidden=hidden, n_layers=n_layers, n_heads=n_heads,
                           n_kv_heads=n_heads, head_dim=head_dim, ffn=ffn, vocab=vocab, rope_theta=0.0,
                           max_pos=max_pos, tied_head=True, bias=True)
    nkv = n_kv_heads if n_kv_heads is not None else max(1, n_heads // 2)
    if family == FAMILY_LLAMA:
        return ModelConfig(name="llama-tiny", family=family, hidden=hidden, n_layers=n_layers, n_heads=n_heads,
                           n_kv_heads=nkv, head_dim=head_dim, ffn=ffn, vocab=vocab, rope_theta=500000.0,
                           max_pos=max_pos)
    if family == FAMILY_MIXTRAL:
        return ModelConfig(name="mixtral-tiny", family=family, hidden=hidden, n_layers=n_layers, n_heads=n_heads,
                           n_kv_heads=nkv, head_dim=head_dim, ffn=ffn, vocab=vocab, n_experts=n_experts, top_k=2,
                           rope_theta=1000000.0, max_pos=max_pos)
    raise ValueError(f"unknown family {family!r}")
