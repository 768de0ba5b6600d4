"""Tensor-parallel parity on >= 2 B200s of one box: one process per GPU, the fused peer-memory all-reduce
(one-shot AND two-shot) after each row-parallel GEMM at decode, NCCL all-reduce at prefill, key all-gather
for the greedy token, logits all-gather on request -- against the oracle's in-process TP simulation.
Skipped (explicitly) when the box has fewer GPUs than the case needs; the CPU suite covers the multi-rank
host logic over gloo (tests/test_oracle.py), and `bench.py --gpus N` runs its own parity leg on the N GPUs
it is given, so the driver's scaling run carries parity evidence even when this file is skipped."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from butterfly_b200 import GPT2_SMALL, LLAMA3_8B, LLAMA3_70B, MIXTRAL_8X7B, SyntheticSource, make_plan, shard_weights, synthetic_tokens, tiny
from oracle import OracleModel, ops
from parity_gates import Gate

pytestmark = pytest.mark.gpu

AR_AUTO, AR_ONE_SHOT, AR_TWO_SHOT = -1, 0, 1


def _worker(rank, world, port, cfg, B, S, steps, out, ar_modes, alternate):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    from butterfly_b200 import Model
    plan = make_plan(cfg, world)
    gm = Model(plan, rank, SyntheticSource(cfg, 0), device=rank, max_batch=B, max_ctx=S + steps + 8)
    gm.connect()
    toks = synthetic_tokens(cfg, B, S)
    teacher = torch.load(out + ".teacher")
    res = {}
    for mode in ar_modes:
        gm.set_ar_mode(mode)
        r = {"prefill": gm.prefill(toks), "tok0": gm.last_tokens.clone(), "decode": []}
        for i in range(steps):
            if alternate:                      # the two protocols interleaved call by call on the same buffers
                dist.barrier()
                gm.set_ar_mode((mode + i) % 2)
            r["decode"].append(gm.decode(teacher[i]))
        gm.set_ar_mode(mode)
        gm.prefill(toks)
        r["greedy"] = gm.decode_greedy(steps)
        res[mode] = r
    torch.save(res, f"{out}.{rank}")
    dist.barrier()
    gm.close()
    dist.destroy_process_group()


def _run(cfg, world, B, S, steps, tmp_path, ar_modes=(AR_AUTO,), alternate=False):
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
    mp.spawn(_worker, args=(world, 29600 + os.getpid() % 1000, cfg, B, S, steps, out, tuple(ar_modes), alternate),
             nprocs=world, join=True)
    results = [torch.load(f"{out}.{r}") for r in range(world)]
    for mode in ar_modes:
        gate = Gate(cfg.n_layers)
        for r, per_mode in enumerate(results):
            res = per_mode[mode]
            gate.add(res["prefill"], want_prefill)
            for got, want in zip(res["decode"], want_decode):
                gate.add(got, want)
            assert torch.equal(res["greedy"], results[0][mode]["greedy"]), "ranks disagree on the greedy tokens"
            assert torch.equal(res["tok0"], ops.argmax(res["prefill"]))
            for a, b in zip(res["decode"], results[0][mode]["decode"]):
                assert torch.equal(a, b), "ranks disagree on the gathered logits"
        print(f"\n[{cfg.name}] tp={world} B={B} ar_mode={mode}{' alternating' if alternate else ''}: {gate.summary()}")
        gate.check(f"{cfg.name} tp={world} ar_mode={mode}")
    # one-shot and two-shot add the tp partial rows in the same (rank) order: their logits are the same bits
    first = results[0][ar_modes[0]]
    for mode in ar_modes[1:]:
        other = results[0][mode]
        for a, b in zip(first["decode"], other["decode"]):
            assert torch.equal(a, b), f"all-reduce modes {ar_modes[0]} and {mode} differ bitwise"
        assert torch.equal(first["greedy"], other["greedy"])


@pytest.mark.parametrize("family", ["llama", "gpt2", "mixtral"])
def test_tp2_tiny(cuda, family, tmp_path):
    _run(tiny(family), 2, B=3, S=10, steps=4, tmp_path=tmp_path, ar_modes=(AR_ONE_SHOT, AR_TWO_SHOT))


def test_tp2_mixtral_grouped_prefill(cuda, tmp_path):
    """A 300-row prefill takes the grouped expert path (4 local experts per rank), its output all-reduced by NCCL."""
    _run(tiny("mixtral"), 2, B=2, S=150, steps=2, tmp_path=tmp_path)


def test_tp2_llama3_8b_width(cuda, tmp_path):
    _run(LLAMA3_8B.with_layers(2), 2, B=4, S=40, steps=3, tmp_path=tmp_path, ar_modes=(AR_ONE_SHOT, AR_TWO_SHOT))


def test_tp2_llama3_8b_width_long_prefill(cuda, tmp_path):
    """320 prefill rows: the bandwidth-regime all-reduce (bf16 payload over NCCL) instead of the fp32 one."""
    _run(LLAMA3_8B.with_layers(2), 2, B=4, S=80, steps=2, tmp_path=tmp_path)


def test_tp2_gpt2_small_full(cuda, tmp_path):
    """BASELINE.json configs[0]: GPT-2 small (124M) TP=2 greedy decode vs the CPU fp32 oracle."""
    _run(GPT2_SMALL, 2, B=1, S=64, steps=32, tmp_path=tmp_path)


def test_tp4_tiny(cuda, tmp_path):
    _run(tiny("llama"), 4, B=2, S=9, steps=3, tmp_path=tmp_path)


@pytest.mark.parametrize("B", [3, 6])
def test_tp4_two_shot_ragged_rows(cuda, B, tmp_path):
    """Two-shot with fewer rows than ranks (B=3) and a row count that is not a multiple of tp (B=6); the modes
    also alternate call by call on the same buffers and flags."""
    _run(tiny("llama"), 4, B=B, S=9, steps=4, tmp_path=tmp_path, ar_modes=(AR_ONE_SHOT, AR_TWO_SHOT))
    _run(tiny("llama"), 4, B=B, S=9, steps=4, tmp_path=tmp_path, ar_modes=(AR_ONE_SHOT,), alternate=True)


def test_tp4_llama3_8b_width(cuda, tmp_path):
    _run(LLAMA3_8B.with_layers(2), 4, B=4, S=48, steps=4, tmp_path=tmp_path, ar_modes=(AR_ONE_SHOT, AR_TWO_SHOT))


def test_tp8_llama3_8b_width(cuda, tmp_path):
    """The north-star shape's width at TP=8 (4 q heads + 1 kv head per rank), depth 2, both all-reduce modes."""
    _run(LLAMA3_8B.with_layers(2), 8, B=4, S=48, steps=4, tmp_path=tmp_path, ar_modes=(AR_ONE_SHOT, AR_TWO_SHOT))


def test_tp8_mixtral_tiny_one_expert_per_rank(cuda, tmp_path):
    _run(tiny("mixtral", n_heads=8, n_kv_heads=8), 8, B=2, S=9, steps=3, tmp_path=tmp_path)


def test_tp8_llama3_70b_width(cuda, tmp_path):
    """BASELINE.json configs[3] shape at TP=8 (8 q heads + 1 kv head per rank), full width, depth 1: the fp32
    oracle of all 80 layers would need 282 GB of host RAM."""
    _run(LLAMA3_70B.with_layers(1), 8, B=4, S=70, steps=3, tmp_path=tmp_path)


def test_tp8_two_shot_70b_width_b64(cuda, tmp_path):
    """The 64 x 8192 decode rows of the 70B configuration (2 MiB per call): the size rule picks the two-shot
    all-reduce by itself (AR_AUTO); the same case is run with one-shot forced and must give the same bits."""
    _run(LLAMA3_70B.with_layers(1), 8, B=64, S=8, steps=3, tmp_path=tmp_path, ar_modes=(AR_AUTO, AR_ONE_SHOT, AR_TWO_SHOT))


def test_tp8_mixtral_8x7b_width(cuda, tmp_path):
    """BASELINE.json configs[4] shape at TP=8: one whole expert per rank, full width, depth 1."""
    _run(MIXTRAL_8X7B.with_layers(1), 8, B=6, S=40, steps=3, tmp_path=tmp_path)


def _ckpt_worker(rank, world, port, cfg, ckpt_dir, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    from butterfly_b200 import Model
    from butterfly_b200.checkpoint import load_checkpoint
    plan = make_plan(cfg, world)
    toks, fed = torch.load(out + ".inputs")
    gm = Model(plan, rank, shard_weights(load_checkpoint(cfg, ckpt_dir), plan, rank), device=rank,
               max_batch=toks.shape[0], max_ctx=toks.shape[1] + 8)
    gm.connect()
    got = [gm.prefill(toks)] + [gm.decode(t) for t in fed]
    torch.save(got, f"{out}.{rank}")
    dist.barrier()
    gm.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("family", ["llama", "mixtral"])
def test_tp2_through_loader_matches_hf(cuda, family, tmp_path):
    """SURVEY.md 8f-1 at tp=2: safetensors -> load_checkpoint -> shard_weights per rank -> two GPUs, against the
    HuggingFace model's own fp32 forward (no oracle involved)."""
    if torch.cuda.device_count() < 2:
        pytest.skip(f"needs 2 GPUs, this box has {torch.cuda.device_count()}")
    pytest.importorskip("safetensors")
    pytest.importorskip("transformers")
    from test_gpu_checkpoint import hf_reference_run
    cfg = tiny(family)
    ckpt = tmp_path / "ckpt"
    ckpt.mkdir()
    _, toks, want, fed = hf_reference_run(cfg, ckpt)
    out = str(tmp_path / "tp")
    torch.save((toks, fed), out + ".inputs")
    mp.spawn(_ckpt_worker, args=(2, 29600 + os.getpid() % 1000, cfg, str(ckpt), out), nprocs=2, join=True)
    for r in range(2):
        gate = Gate()
        for got, w in zip(torch.load(f"{out}.{r}"), want):
            gate.add(got, w)
        print(f"\n[{family}] tp=2 rank {r} through load_checkpoint vs HuggingFace fp32: {gate.summary()}")
        gate.check(f"{family} tp=2 via checkpoint loader vs HF")
