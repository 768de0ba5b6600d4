"""Oracle self-consistency on CPU: TP invariance, KV cache vs recompute, bf16-emulation sanity, golden
fixtures, and the distributed (gloo, world_size 2) path against the in-process simulation."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from butterfly_b200 import SyntheticSource, make_plan, shard_weights, synthetic_tokens, tiny
from oracle import OracleModel, ops

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "oracle_tiny.npz")


def _model(cfg, tp, **kw):
    src = SyntheticSource(cfg, 0)
    plan = make_plan(cfg, tp)
    return OracleModel(plan, [shard_weights(src, plan, r) for r in range(tp)], **kw)


@pytest.mark.parametrize("family", ["llama", "gpt2", "mixtral"])
def test_tp_invariance_and_cache(family):
    cfg = tiny(family)
    toks = synthetic_tokens(cfg, 3, 10)
    ref = _model(cfg, 1)
    l1 = ref.prefill(toks)
    g1 = ref.decode_greedy(4)
    for tp in (2, 4):
        m = _model(cfg, tp)
        assert (m.prefill(toks) - l1).abs().max().item() < 1e-5
        assert torch.equal(m.decode_greedy(4), g1)
    # decode with cache == prefill over the extended prompt
    m = _model(cfg, 2)
    m.prefill(toks)
    for i in range(3):
        last = m.decode(g1[:, i])
    full = _model(cfg, 2).prefill(torch.cat([toks, g1[:, :3]], dim=1))
    assert (last - full).abs().max().item() < 1e-5


def test_bf16_emulation_is_close_but_not_identical():
    cfg = tiny("llama")
    toks = synthetic_tokens(cfg, 2, 12)
    a, b = _model(cfg, 1).prefill(toks), _model(cfg, 1, emulate_bf16=True).prefill(toks)
    err = (a - b).abs().max().item()
    assert 0 < err < 1e-2


def test_argmax_tie_break_lowest_index():
    x = torch.tensor([[0.0, 2.0, 2.0, 1.0], [5.0, 5.0, 5.0, 5.0]])
    assert ops.argmax(x).tolist() == [1, 0]


def test_golden_fixture():
    """Frozen outputs of the oracle (tests/golden/make_golden.py): any later change to the oracle, the
    init recipe or the hash shows up here."""
    z = np.load(GOLDEN)
    for family in ("llama", "gpt2", "mixtral"):
        cfg = tiny(family)
        toks = synthetic_tokens(cfg, 2, 8)
        assert np.array_equal(toks.numpy(), z[f"{family}_tokens"])
        m = _model(cfg, 2)
        logits = m.prefill(toks)
        assert np.abs(logits.numpy() - z[f"{family}_logits"]).max() < 1e-5
        assert np.array_equal(m.decode_greedy(6).numpy(), z[f"{family}_greedy"])
    src = SyntheticSource(tiny("llama"), 0)
    assert np.array_equal(src.block("wq", 0, 0, 2, 0, 8).view(torch.int16).numpy(), z["kat_wq_bits"])


def _gloo_worker(rank, world, port, family, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    cfg = tiny(family)
    plan = make_plan(cfg, world)
    m = OracleModel(plan, [shard_weights(SyntheticSource(cfg, 0), plan, rank)], dist_group=True)
    toks = synthetic_tokens(cfg, 2, 9)
    logits = m.prefill(toks)
    greedy = m.decode_greedy(3)
    if rank == 0:
        torch.save({"logits": logits, "greedy": greedy}, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("family", ["llama", "mixtral"])
def test_distributed_ranks_over_gloo_match_in_process_tp(family, tmp_path):
    out = str(tmp_path / "r0.pt")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_gloo_worker, args=(2, port, family, out), nprocs=2, join=True)
    got = torch.load(out)
    cfg = tiny(family)
    ref = _model(cfg, 2)
    want = ref.prefill(synthetic_tokens(cfg, 2, 9))
    assert (got["logits"] - want).abs().max().item() < 1e-5
    assert torch.equal(got["greedy"], ref.decode_greedy(3))
