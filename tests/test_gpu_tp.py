"""Tensor-parallel parity on >= 2 B200s of one box: one process per GPU, NCCL all-reduce after each
row-parallel GEMM, key all-gather for the greedy token, logits all-gather on request -- against the
oracle's in-process TP simulation.  Skipped (explicitly) on a single-GPU box; the CPU suite covers the
multi-rank host logic over gloo (tests/test_oracle.py)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from butterfly_b200 import GPT2_SMALL, LLAMA3_8B, LLAMA3_70B, MIXTRAL_8X7B, SyntheticSource, make_plan, shard_weights, synthetic_tokens, tiny
from oracle import OracleModel, ops

pytestmark = pytest.mark.gpu
LOGIT_TOL = 1e-2


def _worker(rank, world, port, cfg, B, S, steps, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    from butterfly_b200 import Model
    plan = make_plan(cfg, world)
    gm = Model(plan, rank, SyntheticSource(cfg, 0), device=rank, max_batch=B, max_ctx=S + steps + 8)
    gm.connect()
    toks = synthetic_tokens(cfg, B, S)
    res = {"prefill": gm.prefill(toks), "tok0": gm.last_tokens.clone()}
    teacher = torch.load(out + ".teacher")
    res["decode"] = [gm.decode(teacher[i]) for i in range(steps)]
    gm.prefill(toks)
    res["greedy"] = gm.decode_greedy(steps)
    torch.save(res, f"{out}.{rank}")
    dist.barrier()
    gm.close()
    dist.destroy_process_group()


def _run(cfg, world, B, S, steps, tmp_path):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, this box has {torch.cuda.device_count()}")
    plan = make_plan(cfg, world)
    src = SyntheticSource(cfg, 0)
    om = OracleModel(plan, [shard_weights(src, plan, r) for r in range(world)])
    toks = synthetic_tokens(cfg, B, S)
    want_prefill = om.prefill(toks)
    teacher, want_decode = [], []
    for _ in range(steps):
        t = om.last_tokens.clone()
        teacher.append(t)
        want_decode.append(om.decode(t))
    out = str(tmp_path / "tp")
    torch.save(teacher, out + ".teacher")
    mp.spawn(_worker, args=(world, 29600 + os.getpid() % 1000, cfg, B, S, steps, out), nprocs=world, join=True)
    results = [torch.load(f"{out}.{r}") for r in range(world)]
    for r, res in enumerate(results):
        worst = (res["prefill"] - want_prefill).abs().max().item()
        for got, want in zip(res["decode"], want_decode):
            worst = max(worst, (got - want).abs().max().item())
        assert worst <= LOGIT_TOL, f"rank {r}: logits differ from the oracle by {worst}"
        assert torch.equal(res["greedy"], results[0]["greedy"]), "ranks disagree on the greedy tokens"
        assert torch.equal(res["tok0"], ops.argmax(res["prefill"]))
    print(f"\n[{cfg.name}] tp={world}: max|dlogit|={worst:.3e}")


@pytest.mark.parametrize("family", ["llama", "gpt2", "mixtral"])
def test_tp2_tiny(cuda, family, tmp_path):
    _run(tiny(family), 2, B=3, S=10, steps=4, tmp_path=tmp_path)


def test_tp2_mixtral_grouped_prefill(cuda, tmp_path):
    """A 300-row prefill takes the grouped expert path (4 local experts per rank), its output all-reduced by NCCL."""
    _run(tiny("mixtral"), 2, B=2, S=150, steps=2, tmp_path=tmp_path)


def test_tp2_llama3_8b_width(cuda, tmp_path):
    _run(LLAMA3_8B.with_layers(2), 2, B=4, S=40, steps=3, tmp_path=tmp_path)


def test_tp2_gpt2_small_full(cuda, tmp_path):
    """BASELINE.json configs[0]: GPT-2 small (124M) TP=2 greedy decode vs the CPU fp32 oracle."""
    _run(GPT2_SMALL, 2, B=1, S=64, steps=32, tmp_path=tmp_path)


def test_tp4_tiny(cuda, tmp_path):
    _run(tiny("llama"), 4, B=2, S=9, steps=3, tmp_path=tmp_path)


def test_tp8_mixtral_tiny_one_expert_per_rank(cuda, tmp_path):
    _run(tiny("mixtral", n_heads=8, n_kv_heads=8), 8, B=2, S=9, steps=3, tmp_path=tmp_path)


def test_tp8_llama3_70b_width(cuda, tmp_path):
    """BASELINE.json configs[3] shape at TP=8 (8 q heads + 1 kv head per rank), full width, depth 1: the fp32
    oracle of all 80 layers would need 282 GB of host RAM."""
    _run(LLAMA3_70B.with_layers(1), 8, B=4, S=70, steps=3, tmp_path=tmp_path)


def test_tp8_mixtral_8x7b_width(cuda, tmp_path):
    """BASELINE.json configs[4] shape at TP=8: one whole expert per rank, full width, depth 1."""
    _run(MIXTRAL_8X7B.with_layers(1), 8, B=6, S=40, steps=3, tmp_path=tmp_path)
