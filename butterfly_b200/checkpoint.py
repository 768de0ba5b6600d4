"""Checkpoint hand-over: HuggingFace-layout tensors (safetensors files or a state dict) -> DictSource.

SURVEY.md section 8f item 1 ("real-checkpoint loader").  The reference has no weight format
(ref:CLAUDE.md:21 names a partitioner only), so the accepted layout is the de-facto public one of the
three model families in BASELINE.json; the result feeds `shard_weights` / `Model` unchanged:

    src = load_checkpoint(cfg, "/path/to/model_dir")          # *.safetensors inside
    model = Model(plan, rank, shard_weights(src, plan, rank))

Only the host side: tensors are converted to bf16 logical [out, in] matrices (GPT-2's Conv1D
[in, out] weights are transposed, its fused c_attn is split into q/k/v).
"""
from __future__ import annotations

import glob
import os
from typing import Dict, Tuple

import torch

from .config import FAMILY_GPT2, FAMILY_MIXTRAL, ModelConfig
from .weights import GLOBAL_LAYER, DictSource


def read_safetensors_dir(path: str) -> Dict[str, torch.Tensor]:
    """All tensors of every *.safetensors file under `path` (or of the single file `path`)."""
    from safetensors.torch import load_file
    files = [path] if os.path.isfile(path) else sorted(glob.glob(os.path.join(path, "*.safetensors")))
    if not files:
        raise FileNotFoundError(f"no .safetensors file under {path!r}")
    out: Dict[str, torch.Tensor] = {}
    for f in files:
        out.update(load_file(f))
    return out


def _pick(sd: Dict[str, torch.Tensor], *names: str) -> torch.Tensor:
    for n in names:
        if n in sd:
            return sd[n]
    raise KeyError(f"none of {names} in checkpoint")


def from_hf_state_dict(cfg: ModelConfig, sd: Dict[str, torch.Tensor]) -> DictSource:
    """Map a HuggingFace state dict (Llama / Mixtral / GPT-2 naming) onto the logical tensor set."""
    t: Dict[Tuple[int, str], torch.Tensor] = {}
    G = GLOBAL_LAYER
    if cfg.family == FAMILY_GPT2:
        pre = "transformer." if any(k.startswith("transformer.") for k in sd) else ""
        t[(G, "embed")] = sd[pre + "wte.weight"]
        t[(G, "pos_embed")] = sd[pre + "wpe.weight"]
        t[(G, "final_norm.w")], t[(G, "final_norm.b")] = sd[pre + "ln_f.weight"], sd[pre + "ln_f.bias"]
        qd, kd = cfg.q_dim, cfg.kv_dim
        for l in range(cfg.n_layers):
            p = f"{pre}h.{l}."
            t[(l, "attn_norm.w")], t[(l, "attn_norm.b")] = sd[p + "ln_1.weight"], sd[p + "ln_1.bias"]
            t[(l, "ffn_norm.w")], t[(l, "ffn_norm.b")] = sd[p + "ln_2.weight"], sd[p + "ln_2.bias"]
            w = sd[p + "attn.c_attn.weight"].t()                 # Conv1D stores [in, out]
            b = sd[p + "attn.c_attn.bias"]
            t[(l, "wq")], t[(l, "wk")], t[(l, "wv")] = w[:qd], w[qd:qd + kd], w[qd + kd:]
            t[(l, "bq")], t[(l, "bk")], t[(l, "bv")] = b[:qd], b[qd:qd + kd], b[qd + kd:]
            t[(l, "wo")], t[(l, "bo")] = sd[p + "attn.c_proj.weight"].t(), sd[p + "attn.c_proj.bias"]
            t[(l, "w_up")], t[(l, "b_up")] = sd[p + "mlp.c_fc.weight"].t(), sd[p + "mlp.c_fc.bias"]
            t[(l, "w_down")], t[(l, "b_down")] = sd[p + "mlp.c_proj.weight"].t(), sd[p + "mlp.c_proj.bias"]
        return DictSource(cfg, {k: v.contiguous() for k, v in t.items()})

    t[(G, "embed")] = sd["model.embed_tokens.weight"]
    t[(G, "final_norm.w")] = sd["model.norm.weight"]
    t[(G, "lm_head")] = sd.get("lm_head.weight", sd["model.embed_tokens.weight"])
    for l in range(cfg.n_layers):
        p = f"model.layers.{l}."
        t[(l, "attn_norm.w")] = sd[p + "input_layernorm.weight"]
        t[(l, "ffn_norm.w")] = sd[p + "post_attention_layernorm.weight"]
        for ours, theirs in (("wq", "q_proj"), ("wk", "k_proj"), ("wv", "v_proj"), ("wo", "o_proj")):
            t[(l, ours)] = sd[p + f"self_attn.{theirs}.weight"]
        if cfg.family == FAMILY_MIXTRAL:
            t[(l, "router")] = _pick(sd, p + "block_sparse_moe.gate.weight", p + "mlp.gate.weight")
            fused = [k for k in sd if k.startswith(p) and k.endswith("experts.gate_up_proj")]
            if fused:                                            # transformers >= 5: [E, 2f, h] and [E, h, f]
                gu = sd[fused[0]]
                dn = sd[fused[0].replace("gate_up_proj", "down_proj")]
                if gu.shape[1] != 2 * cfg.ffn:
                    gu, dn = gu.transpose(1, 2), dn.transpose(1, 2)
                for e in range(cfg.n_experts):
                    t[(l, f"e{e}.w_gate")], t[(l, f"e{e}.w_up")] = gu[e, :cfg.ffn], gu[e, cfg.ffn:]
                    t[(l, f"e{e}.w_down")] = dn[e]
            else:                                                # hub layout: experts.{e}.w1 (gate) / w3 (up) / w2 (down)
                for e in range(cfg.n_experts):
                    q = p + f"block_sparse_moe.experts.{e}."
                    t[(l, f"e{e}.w_gate")], t[(l, f"e{e}.w_up")] = sd[q + "w1.weight"], sd[q + "w3.weight"]
                    t[(l, f"e{e}.w_down")] = sd[q + "w2.weight"]
        else:
            t[(l, "w_gate")], t[(l, "w_up")] = sd[p + "mlp.gate_proj.weight"], sd[p + "mlp.up_proj.weight"]
            t[(l, "w_down")] = sd[p + "mlp.down_proj.weight"]
    return DictSource(cfg, {k: v.contiguous() for k, v in t.items()})


def load_checkpoint(cfg: ModelConfig, path: str) -> DictSource:
    return from_hf_state_dict(cfg, read_safetensors_dir(path))
