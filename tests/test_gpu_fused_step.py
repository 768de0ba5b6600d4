"""The fused decode step (csrc/decode_step.cu: one persistent kernel per step, grid-wide phase counters,
stream-K tcgen05 GEMMs, in-kernel RoPE / paged attention / SwiGLU / greedy keys) against the CPU oracle, and
against the kernel-per-op path it replaces (BFLY_FUSED_STEP=0), on shapes that exercise its corner cases:
GQA groups 1..8, batches 1..64, vocabularies that are not a multiple of the 128-row weight tile, context
lengths around tile and page boundaries, split contexts, shuffled pages, ragged batches."""
import os

import pytest
import torch

from butterfly_b200 import LLAMA3_8B, SyntheticSource, make_plan, shard_weights, synthetic_tokens, tiny
from oracle import OracleModel, ops
from parity_gates import Gate

pytestmark = pytest.mark.gpu


def small(n_heads=4, n_kv_heads=2, hidden=512, ffn=1024, vocab=1000, n_layers=2):
    return tiny("llama", hidden=hidden, n_layers=n_layers, n_heads=n_heads, n_kv_heads=n_kv_heads, head_dim=128,
                ffn=ffn, vocab=vocab, max_pos=4096)


def make_model(cfg, fused, **kw):
    from butterfly_b200 import Model
    old = os.environ.get("BFLY_FUSED_STEP")
    os.environ["BFLY_FUSED_STEP"] = "1" if fused else "0"
    try:
        return Model(make_plan(cfg, 1), 0, SyntheticSource(cfg, 0), **kw)
    finally:
        if old is None:
            os.environ.pop("BFLY_FUSED_STEP", None)
        else:
            os.environ["BFLY_FUSED_STEP"] = old


def run_case(cfg, B, S, steps, max_ctx=None, page_table=None):
    plan = make_plan(cfg, 1)
    om = OracleModel(plan, [shard_weights(SyntheticSource(cfg, 0), plan, 0)])
    toks = synthetic_tokens(cfg, B, S)
    want = [om.prefill(toks)]
    teacher = []
    for _ in range(steps):
        t = ops.argmax(want[-1])
        teacher.append(t)
        want.append(om.decode(t))
    outs = {}
    for fused in (True, False):
        gm = make_model(cfg, fused, max_batch=B, max_ctx=max_ctx or (S + steps + 2))
        if page_table is not None:
            gm.set_page_table(page_table)
        got = [gm.prefill(toks)] + [gm.decode(t) for t in teacher]
        nodes = gm.stats()["graph_nodes_decode"]
        gm.prefill(toks)
        greedy = gm.decode_greedy(steps)
        gm.close()
        gate = Gate(cfg.n_layers)
        for a, b in zip(got, want):
            gate.add(a, b)
        print(f"\n[{cfg.name} nq={cfg.n_heads} nkv={cfg.n_kv_heads} V={cfg.vocab}] B={B} S={S} fused={fused} graph nodes={nodes}: {gate.summary()}")
        gate.check(f"fused={fused}")
        assert (nodes <= 4) == fused, f"fused={fused} but the captured decode step has {nodes} kernel nodes"
        outs[fused] = (got, greedy)
    # the two CUDA paths agree with each other as closely as each agrees with the oracle
    g2 = Gate(cfg.n_layers)
    for a, b in zip(outs[True][0], outs[False][0]):
        g2.add(a, b)
    g2.check("fused vs kernel-per-op")
    return outs


@pytest.mark.parametrize("nq,nkv", [(4, 4), (4, 2), (4, 1), (8, 1)])
def test_gqa_groups(cuda, nq, nkv):
    run_case(small(n_heads=nq, n_kv_heads=nkv, hidden=128 * nq), B=3, S=37, steps=4)


@pytest.mark.parametrize("B", [1, 7, 32, 33, 64])
def test_batch_sizes(cuda, B):
    run_case(small(), B=B, S=20, steps=3)


def test_vocab_not_a_multiple_of_the_weight_tile(cuda):
    run_case(small(vocab=1003), B=4, S=9, steps=3)
    run_case(small(vocab=130), B=2, S=5, steps=2)


@pytest.mark.parametrize("S", [1, 31, 32, 33, 63, 64, 65])
def test_context_around_tile_and_page_boundaries(cuda, S):
    run_case(small(), B=2, S=S, steps=3)


def test_split_context_and_shuffled_pages(cuda):
    cfg = small()
    B, max_ctx = 3, 2048                                            # 64 tiles of capacity: the step plan splits the context
    pps = max_ctx // 32
    perm = torch.randperm(B * pps, generator=torch.Generator().manual_seed(5)).to(torch.int32).reshape(B, pps)
    run_case(cfg, B=B, S=700, steps=3, max_ctx=max_ctx, page_table=perm)


def test_deeper_and_wider(cuda):
    run_case(small(n_heads=8, n_kv_heads=2, hidden=1024, ffn=2816, vocab=5000, n_layers=4), B=5, S=50, steps=4)


def test_ragged_batch_through_the_fused_step(cuda):
    """Sequences of different lengths decoded together (per-sequence RoPE position, KV slot and attention span)."""
    cfg = small()
    plan = make_plan(cfg, 1)
    rw = shard_weights(SyntheticSource(cfg, 0), plan, 0)
    groups = [(0, 2, 7), (2, 1, 45), (3, 2, 33)]
    gm = make_model(cfg, True, max_batch=5, max_ctx=64)
    oracles, lo_parts = [], []
    for seq0, n, S in groups:
        toks = synthetic_tokens(cfg, n, S, seed=10 + S)
        om = OracleModel(plan, [rw])
        lo_parts.append(om.prefill(toks))
        gm.prefill_at(seq0, toks)
        oracles.append(om)
    gate = Gate(cfg.n_layers)
    for _ in range(5):
        fed = [ops.argmax(lo) for lo in lo_parts]
        lo_parts = [om.decode(t) for om, t in zip(oracles, fed)]
        gate.add(gm.decode(torch.cat(fed)), torch.cat(lo_parts))
    assert gm.stats()["graph_nodes_decode"] <= 4
    gm.close()
    print(f"\nragged fused step: {gate.summary()}")
    gate.check("ragged")


def test_free_running_greedy_equals_step_by_step(cuda):
    cfg = small()
    toks = synthetic_tokens(cfg, 4, 9)
    gm = make_model(cfg, True, max_batch=4, max_ctx=64)
    gm.prefill(toks)
    fed = gm.decode_greedy(12)
    gm.prefill(toks)
    ref = []
    for _ in range(12):
        ref.append(gm.last_tokens.clone())
        gm.decode(None)
    gm.close()
    assert torch.equal(fed, torch.stack(ref, dim=1))


def test_llama3_8b_width_both_paths(cuda):
    run_case(LLAMA3_8B.with_layers(2), B=32, S=40, steps=3)
