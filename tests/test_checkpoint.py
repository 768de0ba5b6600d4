"""Checkpoint hand-over: HF models saved as safetensors load through
butterfly_b200.checkpoint into the same logical tensors, bit for bit, and the oracle reproduces HF."""
import pytest
import torch

from butterfly_b200 import SyntheticSource, make_plan, shard_weights, synthetic_tokens, tiny
from butterfly_b200.checkpoint import load_checkpoint
from butterfly_b200.weights import logical_tensors, GLOBAL_LAYER
from oracle import OracleModel

pytest.importorskip("safetensors")
from test_oracle_hf import _hf_gpt2, _hf_llama_like  # noqa: E402


@pytest.mark.parametrize("family", ["llama", "gpt2", "mixtral"])
def test_safetensors_roundtrip(family, tmp_path):
    cfg = tiny(family)
    src = SyntheticSource(cfg, 0)
    hf = _hf_gpt2(cfg, src) if family == "gpt2" else _hf_llama_like(cfg, src)
    hf.save_pretrained(str(tmp_path), safe_serialization=True)
    ck = load_checkpoint(cfg, str(tmp_path))
    g_info, l_info = logical_tensors(cfg)
    for name in g_info:
        assert torch.equal(ck.full(name, GLOBAL_LAYER).view(torch.int16), src.full(name, GLOBAL_LAYER).view(torch.int16)), name
    for l in range(cfg.n_layers):
        for name in l_info:
            assert torch.equal(ck.full(name, l).view(torch.int16), src.full(name, l).view(torch.int16)), (l, name)
    plan = make_plan(cfg, 2)
    om = OracleModel(plan, [shard_weights(ck, plan, r) for r in range(2)])
    toks = synthetic_tokens(cfg, 2, 7)
    with torch.no_grad():
        ref = hf(input_ids=toks.long()).logits[:, -1].float()
    assert (om.prefill(toks) - ref).abs().max().item() < 2e-5
