#!/usr/bin/env python
"""Headline benchmark: Llama-3-8B-shaped bf16 greedy decode, batch 32, context 2048, TP = --gpus.

    python bench.py --gpus 1 --steps K --warmup W                 (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference ...                          (the CPU oracle on the host cores)

A "step" is one decode step of the partitioned block forward over the whole batch (32 new tokens).
`value` = tokens/s with tokens, weights and KV cache resident in HBM (K CUDA-graph replays timed with
CUDA events on the engine's stream, max over ranks).  `e2e` = the same metric through the public
Model.decode() call with HOST token buffers: every step copies the input tokens host->device and the
greedy tokens device->host inside the timed region.  `roofline` is measured live for the dominant
kernel (the gate/up projection GEMM).  `cpu_baseline` times the oracle on a bounded sample.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "decode_tokens_per_s"
UNIT = "tokens/s"
BATCH = 32
CTX = 2048
WORKLOAD = "llama3-8b bf16 greedy decode, batch 32, ctx 2048 (BASELINE.json configs[1]/[2])"
# --model selects another BASELINE.json shape (not the headline line): name -> (batch, ctx)
OTHER_WORKLOADS = {"llama3-70b": (64, 4096), "mixtral-8x7b": (32, 2048), "gpt2-small": (1, 64)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------------
# CPU arm: the oracle on the host cores, bounded sample
# --------------------------------------------------------------------------------------------------
def cpu_decode_sample(n_layers_sample: int = 4, steps: int = 3, ctx: int = CTX, batch: int = BATCH):
    """Times the oracle's decode step at full Llama-3-8B width on `n_layers_sample` of the 32 layers
    plus the LM head, with a KV cache of `ctx` tokens per sequence already in place, and scales the
    layer part to 32 layers.  Returns (tokens/s extrapolated, description, cores)."""
    from butterfly_b200 import LLAMA3_8B, make_plan
    from butterfly_b200.weights import RankWeights, shard_spec
    from oracle import OracleModel

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    full = LLAMA3_8B
    cfg = full.with_layers(n_layers_sample)
    plan = make_plan(cfg, 1)
    g = torch.Generator().manual_seed(0)
    tensors = {}
    base = (torch.randn(1 << 22, generator=g) * 0.02).to(torch.bfloat16)
    for st in shard_spec(plan, 0):            # same shapes as the GPU arm; values are irrelevant to timing
        n = st.rows * st.cols
        t = base.repeat((n + base.numel() - 1) // base.numel())[:n].reshape(st.rows, st.cols)
        tensors[st.name] = t.reshape(-1) if st.cols == 1 else t
    om = OracleModel(plan, [RankWeights(plan, 0, tensors)])
    for l in range(cfg.n_layers):              # a cache of `ctx` tokens per sequence, with room to append
        om.k_cache[l][0] = torch.randn((batch, ctx + 16, cfg.n_kv_heads, cfg.head_dim), generator=g)
        om.v_cache[l][0] = torch.randn((batch, ctx + 16, cfg.n_kv_heads, cfg.head_dim), generator=g)
    om.pos = ctx
    tok = torch.randint(0, cfg.vocab, (batch,), generator=g, dtype=torch.int32)
    om.decode(tok)                             # warm-up
    om.pos = ctx
    t0 = time.perf_counter()
    for _ in range(steps):
        om.pos = ctx                           # every timed step attends over exactly `ctx` cached tokens
        om.decode(tok)
    t_sample = (time.perf_counter() - t0) / steps
    # LM head + embedding share, timed separately so only the layer part is scaled
    hn = torch.randn((batch, cfg.hidden))
    w = om.w[0]["lm_head"]
    t1 = time.perf_counter()
    for _ in range(steps):
        _ = hn @ w.t()
    t_head = (time.perf_counter() - t1) / steps
    t_layers = max(t_sample - t_head, 1e-9) / n_layers_sample
    t_full = t_layers * full.n_layers + t_head
    sample = (f"oracle decode step, full width, {n_layers_sample}/32 layers + LM head timed over {steps} steps at "
              f"B={batch} ctx={ctx}; layer time scaled x{full.n_layers // n_layers_sample} (extrapolated)")
    return batch / t_full, sample, cores, t_full


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 3))
    value, sample, cores, t_full = cpu_decode_sample(steps=steps)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t_full * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": BATCH, "ctx": CTX, "parallelism": "cpu"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the reference repository has no implementation; this is the CPU fp32 oracle that stands in for it",
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
def measure_dominant_gemm(lib, M, N, K, iters=20):
    """Live roofline of the dominant kernel: the gate/up projection GEMM of one layer at decode."""
    dev = torch.device("cuda", torch.cuda.current_device())
    n_buf = 3                                            # rotate weights so no launch re-reads L2-resident data
    Ws = [torch.randn((M, K), device=dev, dtype=torch.float32).mul_(0.02).to(torch.bfloat16) for _ in range(n_buf)]
    X = torch.randn((N, K), device=dev, dtype=torch.float32).to(torch.bfloat16)
    splits, bn, stages = C.c_int32(0), C.c_int32(0), C.c_int32(0)
    lib.bfly_k_gemm_plan(M, N, K, C.byref(splits), C.byref(bn), C.byref(stages))
    out = torch.empty((splits.value, N, M), device=dev, dtype=torch.float32)
    used = C.c_int32(0)
    st = torch.cuda.current_stream()

    def launch(i):
        rc = lib.bfly_k_gemm(Ws[i % n_buf].data_ptr(), X.data_ptr(), M, N, K, K, 0, 0, 0, out.data_ptr(), N * M, None,
                             M, C.byref(used), st.cuda_stream)
        assert rc == 0, lib.bfly_last_error(None)

    for i in range(3):
        launch(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(iters):
        launch(i)
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    algo_bytes = 2 * M * K + 2 * N * K + 4 * N * M * used.value
    return {"kernel": "gemm_tn_kernel (gate/up projection, tcgen05)", "shape": [M, N, K], "splits": used.value,
            "ms": ms, "algo_bytes": algo_bytes, "gbs": algo_bytes / ms / 1e6}


def measure_decode_attention(lib, batch, ctx, nq, nk, d, max_ctx, iters=12):
    """Live roofline of the second-largest kernel: paged decode attention of one layer (K and V of every
    cached token read once).  Same split rule as the engine (csrc/attn_decode.cu: attn_decode_splits)."""
    dev = torch.device("cuda", torch.cuda.current_device())
    ps = 32
    pps = (ctx + ps - 1) // ps
    n_pages = batch * pps
    n_buf = 3                                            # rotate caches so no launch re-reads L2-resident data
    pages = [torch.randn((n_pages, 2, nk, ps * d), device=dev, dtype=torch.float32).to(torch.bfloat16)
             for _ in range(n_buf)]
    table = torch.arange(n_pages, dtype=torch.int32, device=dev).reshape(batch, pps)
    q = torch.randn((batch, nq * d), device=dev, dtype=torch.float32).to(torch.bfloat16)
    lens = torch.full((batch,), ctx - 1, dtype=torch.int32, device=dev)
    out = torch.empty((batch, nq * d), device=dev, dtype=torch.bfloat16)
    n_splits, tiles = 1, max(1, (max_ctx + 31) // 32)
    while batch * nk * n_splits < 96 and n_splits < 32 and tiles // (n_splits + 1) >= 8:
        n_splits += 1
    part_o = torch.empty((batch * nq * n_splits * d,), device=dev, dtype=torch.float32)
    part_ml = torch.empty((batch * nq * n_splits * 2,), device=dev, dtype=torch.float32)
    st = torch.cuda.current_stream()

    def launch(i):
        rc = lib.bfly_k_attn_decode(q.data_ptr(), pages[i % n_buf].data_ptr(), table.data_ptr(), pps, ps, nk, d,
                                    lens.data_ptr(), 1, batch, nq, part_o.data_ptr(), part_ml.data_ptr(), n_splits,
                                    out.data_ptr(), st.cuda_stream)
        assert rc == 0, lib.bfly_last_error(None)

    for i in range(3):
        launch(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(iters):
        launch(i)
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    algo_bytes = batch * ctx * nk * d * 2 * 2
    return {"kernel": "attn_decode_kernel (+ attn_combine_kernel when split)", "n_splits": n_splits, "kernel_ms": ms,
            "algo_bytes_per_launch": algo_bytes, "achieved": algo_bytes / ms / 1e6, "unit": "GB/s", "bound": "hbm"}


def run_gpu(args):
    from butterfly_b200 import LLAMA3_8B, Model, SyntheticSource, _lib, make_plan, synthetic_tokens
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.shard_of > 1:            # profiling aid: one rank's shard on one GPU, collectives skipped
        os.environ["BFLY_SKIP_COLLECTIVES"] = "1"
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch N>1 with torchrun)"
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    global BATCH, CTX, WORKLOAD
    base_cfg = LLAMA3_8B
    if args.model != "llama3-8b":
        from butterfly_b200 import PRESETS
        base_cfg = PRESETS[args.model]
        BATCH, CTX = OTHER_WORKLOADS[args.model]
        WORKLOAD = f"{args.model} bf16 greedy decode, batch {BATCH}, ctx {CTX} (BASELINE.json configs; not the headline line)"
    cfg = base_cfg if args.layers <= 0 else base_cfg.with_layers(args.layers)
    tp = world if args.shard_of <= 1 else args.shard_of
    plan = make_plan(cfg, tp)
    W, K = args.warmup, args.steps
    max_ctx = CTX + 2 * (max(W, 3) + K) + 16
    t_setup = time.perf_counter()
    model = Model(plan, rank, SyntheticSource(cfg, 0), device=local, max_batch=BATCH, max_ctx=max_ctx)
    if args.shard_of <= 1:
        model.connect()
    toks = synthetic_tokens(cfg, BATCH, CTX)
    lib, stream = model.lib, model.stream

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- prefill of the 2048-token prompts (also reported) ----
    model.prefill(toks[:1, :128].contiguous(), want_logits=False)      # warm-up: module load, NCCL channels
    barrier()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record(stream)
    model.prefill(toks, want_logits=False)
    p1.record(stream)
    barrier()
    prefill_ms = p0.elapsed_time(p1)
    setup_s = time.perf_counter() - t_setup

    # ---- device-resident decode: W warm-up + K timed CUDA-graph replays ----
    model.decode_enqueue(BATCH, max(W, 3))
    barrier()
    launches0 = model.stats()["kernel_launches"]
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    model.decode_enqueue(BATCH, K)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    launches = model.stats()["kernel_launches"] - launches0

    # ---- end to end through the public API with host buffers ----
    pinned_in = torch.zeros((BATCH,), dtype=torch.int32).pin_memory()
    pinned_in.copy_(model.decode_greedy(1)[:, 0])
    for _ in range(2):
        model.decode(pinned_in, want_logits=False)
        pinned_in.copy_(model.last_tokens)
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        model.decode(pinned_in, want_logits=False)       # H2D tokens, graph replay, D2H greedy tokens, sync
        pinned_in.copy_(model.last_tokens)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0

    t = torch.tensor([ms, e2e_s * 1e3, prefill_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms, prefill_ms = t.tolist()

    if rank == 0:
        hbm_peak, peak_src = peaks()
        ms_step = ms / K
        value = BATCH * K / (ms / 1e3)
        ctx_mid = CTX + max(W, 3) + K / 2.0
        step_bytes = plan.weight_bytes_per_rank() + plan.kv_bytes_per_token_per_rank() * BATCH * (ctx_mid + 1)
        gemm = measure_dominant_gemm(lib, plan.ffn_local * (2 if cfg.gated else 1) * max(1, plan.experts_local), BATCH,
                                     cfg.hidden)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(W, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "model": cfg.name, "global_batch": BATCH, "ctx": CTX,
                       "parallelism": f"tp{tp}", "l2": "inputs_exceed_l2 (weights+KV per step >> 126 MB)",
                       "layers": cfg.n_layers},
            "e2e": {"value": BATCH * K / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": BATCH * 4,
                    "d2h_bytes_per_step": BATCH * 4},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": gemm["gbs"], "peak": hbm_peak, "unit": "GB/s",
                         "frac": gemm["gbs"] / hbm_peak,
                         # dram__bytes_read.sum + dram__bytes_write.sum of this launch from the committed
                         # `ncu --set full` capture (profiles/r01_ncu_full_gemm_attn.txt): 235.17 + 3.96 MB
                         "traffic": 239129344 if (tp == 1 and cfg.n_layers == 32) else None,
                         "traffic_source": "profiles/r01_ncu_full_gemm_attn.txt", "peak_source": peak_src,
                         "kernel": gemm["kernel"], "shape_MNK": gemm["shape"], "splits": gemm["splits"],
                         "kernel_ms": gemm["ms"], "algo_bytes_per_launch": gemm["algo_bytes"]},
            "step_roofline": {"algo_bytes_per_step_per_gpu": step_bytes, "achieved_gbs_per_gpu": step_bytes / ms_step / 1e6,
                              "frac_of_hbm_peak": step_bytes / ms_step / 1e6 / hbm_peak},
            "prefill": {"tokens_per_s": BATCH * CTX / (prefill_ms / 1e3), "ms": prefill_ms, "tokens": BATCH * CTX},
            "setup_s": setup_s,
        }
        try:                                             # second-largest kernel; never allowed to cost the line
            attn = measure_decode_attention(lib, BATCH, CTX, plan.q_local, plan.kv_local, cfg.head_dim, model.max_ctx)
            attn["peak"], attn["frac"] = hbm_peak, attn["achieved"] / hbm_peak
            line["roofline_attention"] = attn
        except Exception as exc:                         # noqa: BLE001
            line["roofline_attention"] = {"error": str(exc)[:200]}
        if args.shard_of > 1:
            line["INVALID"] = (f"profiling run: rank 0's shard of a tp={tp} plan on one GPU with the collectives "
                               "skipped; per-rank compute time only, results are not a model output")
        if not args.no_cpu and world == 1 and args.shard_of <= 1 and args.model == "llama3-8b":
            v, sample, cores, _ = cpu_decode_sample()
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    model.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--layers", type=int, default=0, help="debug: reduce depth (the result is then NOT the headline config)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--model", default="llama3-8b", choices=["llama3-8b", "llama3-70b", "mixtral-8x7b", "gpt2-small"],
                    help="shape to run; anything but llama3-8b is an extra (non-headline) line")
    ap.add_argument("--shard-of", type=int, default=0, help="profiling aid: run rank 0's shard of a tp=N plan on one "
                    "GPU with the collectives skipped (the JSON line is marked INVALID)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
