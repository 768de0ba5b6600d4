"""Host logic: partition plan and shard properties (hypothesis), synthetic weight determinism."""
import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

from butterfly_b200 import (GPT2_SMALL, LLAMA3_8B, LLAMA3_70B, MIXTRAL_8X7B, PlanError, SyntheticSource, make_plan,
                            shard_spec, shard_weights, tiny)
from butterfly_b200.weights import GLOBAL_LAYER, logical_tensors, synth_block, synth_scale, tensor_key


@pytest.mark.parametrize("cfg,tps", [(GPT2_SMALL, (1, 2, 4)), (LLAMA3_8B, (1, 2, 4, 8)), (LLAMA3_70B, (1, 2, 4, 8)),
                                     (MIXTRAL_8X7B, (1, 2, 4, 8))])
def test_plans_of_the_baseline_configs(cfg, tps):
    for tp in tps:
        plan = make_plan(cfg, tp)
        q = [h for r in plan.ranks for h in range(*r.q_heads)]
        kv = [h for r in plan.ranks for h in range(*r.kv_heads)]
        voc = [v for r in plan.ranks for v in (r.vocab_rows,)]
        assert q == list(range(cfg.n_heads)) and kv == list(range(cfg.n_kv_heads))
        assert voc[0][0] == 0 and voc[-1][1] == cfg.vocab and all(a[1] == b[0] for a, b in zip(voc, voc[1:]))
        if cfg.is_moe:
            assert sorted(e for r in plan.ranks for e in r.experts) == list(range(cfg.n_experts))
        else:
            cols = [r.ffn_cols for r in plan.ranks]
            assert cols[0][0] == 0 and cols[-1][1] == cfg.ffn and all(a[1] == b[0] for a, b in zip(cols, cols[1:]))
    assert abs(LLAMA3_8B.n_params() - 8.03e9) < 0.01e9 and abs(LLAMA3_70B.n_params() - 70.55e9) < 0.05e9
    assert abs(MIXTRAL_8X7B.n_params() - 46.70e9) < 0.05e9 and abs(GPT2_SMALL.n_params() - 124.4e6) < 0.5e6


def test_survey_byte_table():
    """8B decode B=32 ctx=2048 = 23.60 GB/step; per-GPU 2.950 GB at tp=8."""
    for tp, per_gpu in ((1, 23.604), (2, 11.802), (4, 5.901), (8, 2.950)):
        plan = make_plan(LLAMA3_8B, tp)
        b = plan.weight_bytes_per_rank() + plan.kv_bytes_per_token_per_rank() * 32 * 2048
        assert abs(b / 1e9 - per_gpu) < 0.02 * per_gpu


def test_invalid_plans_raise():
    with pytest.raises(PlanError):
        make_plan(GPT2_SMALL, 8)            # 12 heads do not divide by 8
    with pytest.raises(PlanError):
        make_plan(LLAMA3_8B, 16)            # 8 KV heads
    with pytest.raises(PlanError):
        make_plan(LLAMA3_8B, 0)
    with pytest.raises(PlanError):
        make_plan(LLAMA3_8B, 2, page_size=16)


@settings(max_examples=25, deadline=None)
@given(family=st.sampled_from(["llama", "gpt2", "mixtral"]), tp=st.sampled_from([1, 2, 4]),
       vocab=st.integers(min_value=37, max_value=700))
def test_shards_tile_every_logical_tensor_exactly_once(family, tp, vocab):
    cfg = tiny(family, hidden=64, n_layers=1, n_heads=4, n_kv_heads=4, head_dim=32, ffn=128, vocab=vocab, n_experts=4)
    plan = make_plan(cfg, tp)
    g_info, l_info = logical_tensors(cfg)
    cover = {}
    replicated = {"embed", "pos_embed", "final_norm.w", "final_norm.b", "attn_norm.w", "attn_norm.b", "ffn_norm.w",
                  "ffn_norm.b", "bo", "b_down", "router"}
    for r in range(tp):
        for stt in shard_spec(plan, r):
            filled = np.zeros((stt.rows, stt.cols), dtype=np.int32)
            for pc in stt.pieces:
                info = g_info[pc.name] if pc.layer == GLOBAL_LAYER else l_info[pc.name]
                assert 0 <= pc.r0 <= pc.r1 <= info.rows and 0 <= pc.c0 <= pc.c1 <= info.cols
                filled[pc.dst_row:pc.dst_row + pc.r1 - pc.r0, pc.dst_col:pc.dst_col + pc.c1 - pc.c0] += 1
                if pc.name in replicated or (pc.name == "embed" and stt.name == "embed"):
                    continue
                key = (pc.layer, pc.name if stt.name != "lm_head" else "lm_head:" + pc.name)
                c = cover.setdefault(key, np.zeros((info.rows, info.cols), dtype=np.int32))
                c[pc.r0:pc.r1, pc.c0:pc.c1] += 1
            assert (filled == 1).all(), stt.name          # every element of a rank tensor written once
    for key, c in cover.items():
        assert (c == 1).all(), key                         # every logical element owned by exactly one rank


def test_synthetic_weights_are_a_pure_function_of_position():
    cfg = tiny("llama")
    src = SyntheticSource(cfg, 0)
    full = src.full("wq", 1)
    assert torch.equal(src.block("wq", 1, 17, 40, 5, 99), full[17:40, 5:99])
    assert not torch.equal(full, SyntheticSource(cfg, 1).full("wq", 1))
    assert not torch.equal(full, src.full("wq", 0))
    assert abs(full.float().std().item() - 0.02) < 0.002 and abs(full.float().mean().item()) < 1e-3
    g = src.full("attn_norm.w", 0).float()
    assert abs(g.mean().item() - 1.0) < 0.01
    # the bits of the hash (splitmix64 / Irwin-Hall) are pinned by the golden file: test_oracle.test_golden_fixture
    assert tensor_key(0, 0, "wq") == tensor_key(0, 0, "wq") != tensor_key(0, 0, "wk")


def test_rank_weights_shapes_8b_tp8_without_materialising():
    plan = make_plan(LLAMA3_8B, 8)
    spec = {s.name: (s.rows, s.cols) for s in shard_spec(plan, 3)}
    assert spec["layers.0.wqkv"] == (768, 4096) and spec["layers.0.wo"] == (4096, 512)
    assert spec["layers.0.w_gate_up"] == (3584, 4096) and spec["layers.0.w_down"] == (4096, 1792)
    assert spec["lm_head"] == (16032, 4096)
    mx = {s.name: (s.rows, s.cols) for s in shard_spec(make_plan(MIXTRAL_8X7B, 8), 5)}
    assert mx["layers.0.w_gate_up"] == (2 * 14336, 4096) and mx["layers.0.w_down"] == (4096, 14336)
