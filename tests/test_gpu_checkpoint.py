"""SURVEY.md section 8f-1 on the GPU: an HF model saved as safetensors -> butterfly_b200.checkpoint.load_checkpoint
-> shard_weights -> Model on the B200, compared with the HuggingFace model's OWN fp32 forward (prefill + 3
decode steps).  This is a GPU-vs-independent-implementation check: the oracle is not involved at all."""
import pytest
import torch

from butterfly_b200 import SyntheticSource, make_plan, shard_weights, synthetic_tokens, tiny
from butterfly_b200.checkpoint import load_checkpoint
from parity_gates import Gate

pytest.importorskip("safetensors")
pytest.importorskip("transformers")
from test_oracle_hf import _hf_gpt2, _hf_llama_like  # noqa: E402

pytestmark = pytest.mark.gpu


def hf_reference_run(cfg, tmp_path, B=3, S=12, steps=3):
    """Builds the HF model offline from config (weights = the synthetic source, so sizes and scales are the usual
    ones), saves it as safetensors and returns (checkpoint source, tokens, [HF logits of prefill and each decode
    step], [token fed at each decode step])."""
    src = SyntheticSource(cfg, 0)
    hf = _hf_gpt2(cfg, src) if cfg.family == "gpt2" else _hf_llama_like(cfg, src)
    hf.save_pretrained(str(tmp_path), safe_serialization=True)
    ck = load_checkpoint(cfg, str(tmp_path))
    toks = synthetic_tokens(cfg, B, S)
    seq, want, fed = toks.long(), [], []
    with torch.no_grad():
        want.append(hf(input_ids=seq).logits[:, -1].float())
        for _ in range(steps):
            nxt = want[-1].argmax(-1)
            fed.append(nxt.to(torch.int32))
            seq = torch.cat([seq, nxt[:, None]], 1)
            want.append(hf(input_ids=seq).logits[:, -1].float())     # full recompute: no HF cache involved
    return ck, toks, want, fed


@pytest.mark.parametrize("family", ["llama", "gpt2", "mixtral"])
def test_gpu_through_loader_matches_hf(cuda, family, tmp_path):
    from butterfly_b200 import Model
    cfg = tiny(family)
    ck, toks, want, fed = hf_reference_run(cfg, tmp_path)
    plan = make_plan(cfg, 1)
    gm = Model(plan, 0, shard_weights(ck, plan, 0), max_batch=toks.shape[0], max_ctx=toks.shape[1] + 8)
    gate = Gate()
    gate.add(gm.prefill(toks), want[0])
    for i, t in enumerate(fed):
        gate.add(gm.decode(t), want[i + 1])
    gm.close()
    print(f"\n[{family}] GPU through load_checkpoint vs HuggingFace fp32: {gate.summary()}")
    gate.check(f"{family} via checkpoint loader vs HF")
