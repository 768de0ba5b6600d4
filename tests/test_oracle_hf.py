"""Pins the oracle against HuggingFace transformers reference models built offline from config
with identical weights (the reference itself has no implementation)."""
import pytest
import torch

from butterfly_b200 import SyntheticSource, make_plan, shard_weights, synthetic_tokens, tiny
from butterfly_b200.weights import GLOBAL_LAYER
from oracle import OracleModel

transformers = pytest.importorskip("transformers")


def _full(src, name, layer=GLOBAL_LAYER):
    t = src.full(name, layer).float()
    return t.reshape(-1) if t.shape[1] == 1 else t


def _hf_llama_like(cfg, src):
    if cfg.is_moe:
        from transformers import MixtralConfig, MixtralForCausalLM
        hc = MixtralConfig(vocab_size=cfg.vocab, hidden_size=cfg.hidden, intermediate_size=cfg.ffn,
                           num_hidden_layers=cfg.n_layers, num_attention_heads=cfg.n_heads,
                           num_key_value_heads=cfg.n_kv_heads, head_dim=cfg.head_dim, rms_norm_eps=cfg.norm_eps,
                           rope_theta=cfg.rope_theta, num_local_experts=cfg.n_experts,
                           num_experts_per_tok=cfg.top_k, max_position_embeddings=cfg.max_pos,
                           tie_word_embeddings=False, sliding_window=None, attention_dropout=0.0)
        m = MixtralForCausalLM(hc)
    else:
        from transformers import LlamaConfig, LlamaForCausalLM
        hc = LlamaConfig(vocab_size=cfg.vocab, hidden_size=cfg.hidden, intermediate_size=cfg.ffn,
                         num_hidden_layers=cfg.n_layers, num_attention_heads=cfg.n_heads,
                         num_key_value_heads=cfg.n_kv_heads, head_dim=cfg.head_dim, rms_norm_eps=cfg.norm_eps,
                         rope_theta=cfg.rope_theta, max_position_embeddings=cfg.max_pos,
                         tie_word_embeddings=False, attention_bias=False, mlp_bias=False)
        m = LlamaForCausalLM(hc)
    sd = m.state_dict()
    new = {"model.embed_tokens.weight": _full(src, "embed"), "model.norm.weight": _full(src, "final_norm.w"),
           "lm_head.weight": _full(src, "lm_head")}
    for l in range(cfg.n_layers):
        p = f"model.layers.{l}."
        new[p + "input_layernorm.weight"] = _full(src, "attn_norm.w", l)
        new[p + "post_attention_layernorm.weight"] = _full(src, "ffn_norm.w", l)
        for a, b in (("q_proj", "wq"), ("k_proj", "wk"), ("v_proj", "wv"), ("o_proj", "wo")):
            new[p + f"self_attn.{a}.weight"] = _full(src, b, l)
        if cfg.is_moe:
            gate_key = [k for k in sd if k.startswith(p) and "gate" in k and "experts" not in k]
            assert len(gate_key) == 1, gate_key
            new[gate_key[0]] = _full(src, "router", l)
            ek = [k for k in sd if k.startswith(p) and "experts" in k]
            if any("gate_up_proj" in k for k in ek):        # fused expert tensors (transformers >= 5)
                gu = torch.stack([torch.cat([_full(src, f"e{e}.w_gate", l), _full(src, f"e{e}.w_up", l)], 0)
                                  for e in range(cfg.n_experts)])
                dn = torch.stack([_full(src, f"e{e}.w_down", l) for e in range(cfg.n_experts)])
                kgu = [k for k in ek if "gate_up_proj" in k][0]
                kdn = [k for k in ek if "down_proj" in k][0]
                new[kgu] = gu if tuple(sd[kgu].shape) == tuple(gu.shape) else gu.transpose(1, 2).contiguous()
                new[kdn] = dn if tuple(sd[kdn].shape) == tuple(dn.shape) else dn.transpose(1, 2).contiguous()
            else:
                for e in range(cfg.n_experts):
                    new[p + f"block_sparse_moe.experts.{e}.w1.weight"] = _full(src, f"e{e}.w_gate", l)
                    new[p + f"block_sparse_moe.experts.{e}.w3.weight"] = _full(src, f"e{e}.w_up", l)
                    new[p + f"block_sparse_moe.experts.{e}.w2.weight"] = _full(src, f"e{e}.w_down", l)
        else:
            new[p + "mlp.gate_proj.weight"] = _full(src, "w_gate", l)
            new[p + "mlp.up_proj.weight"] = _full(src, "w_up", l)
            new[p + "mlp.down_proj.weight"] = _full(src, "w_down", l)
    missing = [k for k in sd if k not in new and "rotary" not in k]
    assert not missing, missing
    for k, v in new.items():
        assert tuple(sd[k].shape) == tuple(v.shape), (k, sd[k].shape, v.shape)
    m.load_state_dict(new, strict=False)
    return m.float().eval()


def _hf_gpt2(cfg, src):
    from transformers import GPT2Config, GPT2LMHeadModel
    hc = GPT2Config(vocab_size=cfg.vocab, n_positions=cfg.max_pos, n_embd=cfg.hidden, n_layer=cfg.n_layers,
                    n_head=cfg.n_heads, n_inner=cfg.ffn, activation_function="gelu_new",
                    layer_norm_epsilon=cfg.norm_eps, resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
    m = GPT2LMHeadModel(hc)
    new = {"transformer.wte.weight": _full(src, "embed"), "transformer.wpe.weight": _full(src, "pos_embed"),
           "transformer.ln_f.weight": _full(src, "final_norm.w"), "transformer.ln_f.bias": _full(src, "final_norm.b"),
           "lm_head.weight": _full(src, "embed")}
    for l in range(cfg.n_layers):
        p = f"transformer.h.{l}."
        new[p + "ln_1.weight"], new[p + "ln_1.bias"] = _full(src, "attn_norm.w", l), _full(src, "attn_norm.b", l)
        new[p + "ln_2.weight"], new[p + "ln_2.bias"] = _full(src, "ffn_norm.w", l), _full(src, "ffn_norm.b", l)
        # HF Conv1D stores [in, out]
        new[p + "attn.c_attn.weight"] = torch.cat([_full(src, n, l) for n in ("wq", "wk", "wv")], 0).t().contiguous()
        new[p + "attn.c_attn.bias"] = torch.cat([_full(src, n, l) for n in ("bq", "bk", "bv")], 0)
        new[p + "attn.c_proj.weight"] = _full(src, "wo", l).t().contiguous()
        new[p + "attn.c_proj.bias"] = _full(src, "bo", l)
        new[p + "mlp.c_fc.weight"] = _full(src, "w_up", l).t().contiguous()
        new[p + "mlp.c_fc.bias"] = _full(src, "b_up", l)
        new[p + "mlp.c_proj.weight"] = _full(src, "w_down", l).t().contiguous()
        new[p + "mlp.c_proj.bias"] = _full(src, "b_down", l)
    sd = m.state_dict()
    missing = [k for k in sd if k not in new and not k.endswith(".attn.bias") and not k.endswith("masked_bias")]
    assert not missing, missing
    m.load_state_dict(new, strict=False)
    return m.float().eval()


@pytest.mark.parametrize("family,tp", [("llama", 1), ("llama", 2), ("gpt2", 2), ("mixtral", 1), ("mixtral", 4)])
def test_oracle_matches_hf(family, tp):
    torch.manual_seed(0)
    cfg = tiny(family)
    src = SyntheticSource(cfg, 0)
    hf = _hf_gpt2(cfg, src) if family == "gpt2" else _hf_llama_like(cfg, src)
    plan = make_plan(cfg, tp)
    om = OracleModel(plan, [shard_weights(src, plan, r) for r in range(tp)])
    toks = synthetic_tokens(cfg, 2, 9)
    logits = om.prefill(toks)
    seq = toks.long()
    with torch.no_grad():
        ref = hf(input_ids=seq).logits[:, -1].float()
    assert (logits - ref).abs().max().item() < 2e-5
    # three decode steps against a full recompute by HF (pins KV-cache positions and RoPE offsets)
    for _ in range(3):
        nxt = om.last_tokens
        seq = torch.cat([seq, nxt.long()[:, None]], 1)
        logits = om.decode(nxt)
        with torch.no_grad():
            ref = hf(input_ids=seq).logits[:, -1].float()
        assert (logits - ref).abs().max().item() < 2e-5
        assert torch.equal(logits.argmax(-1), ref.argmax(-1))
