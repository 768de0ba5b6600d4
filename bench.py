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


def tensor_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        if "bf16_tflops_sustained" in d:
            return float(d["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
    return 2250.0, "fallback (nominal dense bf16 2.25 PFLOP/s)"


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
def _mem_available_gb() -> float:
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


class CpuDecode:
    """The oracle's decode step of the FULL Llama-3-8B shape (32 layers + LM head, B=32, a KV cache of `ctx`
    tokens per sequence already in place) on the host cores.  Weight VALUES do not matter for timing, so they are
    a repeated random block; when the host cannot hold the 32 GB of fp32 weights (+17 GB of fp32 KV), the 32
    layer applications cycle through `distinct` sets of weights and caches (4 sets = 3.5 GB of weights, far
    beyond any CPU cache), which leaves flops and bytes per step unchanged."""

    def __init__(self, ctx: int = CTX, batch: int = BATCH, full=None, distinct=None):
        """`full` (default: the Llama-3-8B shape) and `distinct` (default: by available host memory) exist so the CPU
        suite can run this class on a tiny shape."""
        from butterfly_b200 import LLAMA3_8B, make_plan
        from butterfly_b200.weights import RankWeights, shard_spec
        from oracle import OracleModel

        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        full = full or LLAMA3_8B
        self.avail_gb = _mem_available_gb()
        self.distinct = distinct or (full.n_layers if self.avail_gb >= 80.0 else min(4, full.n_layers))
        cfg = full.with_layers(self.distinct)
        plan = make_plan(cfg, 1)
        g = torch.Generator().manual_seed(0)
        tensors = {}
        base = (torch.randn(1 << 22, generator=g) * 0.02).to(torch.bfloat16)
        for st in shard_spec(plan, 0):            # same shapes as the GPU arm
            n = st.rows * st.cols
            t = base.repeat((n + base.numel() - 1) // base.numel())[:n].reshape(st.rows, st.cols)
            tensors[st.name] = t.reshape(-1) if st.cols == 1 else t
        om = OracleModel(plan, [RankWeights(plan, 0, tensors)])
        del tensors
        for l in range(cfg.n_layers):              # a cache of `ctx` tokens per sequence, with room to append
            om.k_cache[l][0] = torch.randn((batch, ctx + 16, cfg.n_kv_heads, cfg.head_dim), generator=g)
            om.v_cache[l][0] = torch.randn((batch, ctx + 16, cfg.n_kv_heads, cfg.head_dim), generator=g)
        if self.distinct < full.n_layers:          # layer l runs on the weights and cache of set l % distinct
            w = om.w[0]
            for l in range(self.distinct, full.n_layers):
                for name in [k for k in w if k.startswith("layers.0.")]:
                    w[f"layers.{l}." + name[len("layers.0."):]] = w[f"layers.{l % self.distinct}." + name[len("layers.0."):]]
                om.k_cache.append(om.k_cache[l % self.distinct])
                om.v_cache.append(om.v_cache[l % self.distinct])
            om.cfg = full
            om.plan = make_plan(full, 1)
        self.om, self.ctx, self.batch, self.layers = om, ctx, batch, full.n_layers
        self.tok = torch.randint(0, full.vocab, (batch,), generator=g, dtype=torch.int32)

    def step(self) -> float:
        self.om.pos = self.ctx                     # every step attends over exactly `ctx` cached tokens
        t0 = time.perf_counter()
        self.om.decode(self.tok)
        return time.perf_counter() - t0

    def describe(self, steps: int) -> str:
        how = ("32 distinct layers" if self.distinct == self.layers else
               f"32 layer applications cycling through {self.distinct} distinct weight/KV sets "
               f"(host has {self.avail_gb:.0f} GB available; same flops and bytes per step)")
        return (f"oracle (CPU fp32 PyTorch) decode step of the full llama3-8b shape, {how}, + LM head, B={self.batch} "
                f"ctx={self.ctx}, random weights, {steps} timed steps, {self.cores} threads")


def run_reference(args):
    """The stand-in reference (the reference repository has no code): every one of the --steps steps is a real,
    full-cost decode step of the oracle on the host cores; nothing is extrapolated."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cpu = CpuDecode()
    for _ in range(args.warmup):
        cpu.step()
    times = [cpu.step() for _ in range(args.steps)]
    total = sum(times)
    value = BATCH * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": gpu_config("llama3-8b", 32, max(1, args.gpus)),
        "extrapolated": False, "layers_timed": cpu.layers, "distinct_weight_sets": cpu.distinct,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu.cores, "kind": "port", "sample": cpu.describe(args.steps)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the reference repository has no implementation; this is the CPU fp32 oracle that stands in for it "
                "(config.parallelism names the GPU arm's layout; this arm runs on the host cores only)",
    }
    print(json.dumps(line), flush=True)


def gpu_config(model: str, layers: int, tp: int) -> dict:
    return {"workload": WORKLOAD, "model": model, "global_batch": BATCH, "ctx": CTX, "parallelism": f"tp{tp}",
            "l2": "inputs_exceed_l2 (weights+KV per step >> 126 MB)", "layers": layers}


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


def source_sha16(*names) -> str:
    """Identifies the kernel source a committed ncu traffic figure belongs to (git is not on the GPU box)."""
    import hashlib
    h = hashlib.sha256()
    for n in names:
        with open(os.path.join(ROOT, "butterfly_b200", "csrc", n), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def committed_traffic(kernel_key: str, shape, sources):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
    (profiles/r02_traffic.json: one entry per (kernel, shape), with the sha of the kernel source it was taken
    from).  Returns (bytes or None, source file or None, stale flag)."""
    path = os.path.join(ROOT, "profiles", "r02_traffic.json")
    try:
        with open(path) as f:
            table = json.load(f)
    except (OSError, ValueError):
        return None, None, None
    for e in table.get("entries", []):
        if e.get("kernel") == kernel_key and list(e.get("shape", [])) == list(shape):
            return e.get("dram_bytes"), e.get("from"), e.get("source_sha16") != source_sha16(*sources)
    return None, None, None


def parity_leg(dist, world, rank, local):
    """Before anything is timed: the same engine, communicator layout and all-reduce modes as the timed run,
    against the CPU fp32 oracle (rank 0 holds all `world` rank shards in-process and broadcasts the teacher
    tokens).  Uses the oracle strictly as the checker.  Returns the `parity` object; any failure aborts."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from parity_gates import LOGIT_TOL, REL_MAX, REL_RMS, Gate
    from butterfly_b200 import LLAMA3_8B, LLAMA3_70B, Model, SyntheticSource, make_plan, shard_weights, synthetic_tokens
    import hashlib
    cases = [("llama3-8b width, 2 layers, B=4, prefill 48 + 4 teacher-forced steps", LLAMA3_8B.with_layers(2), 4, 48, 4,
              (0, 1) if world > 1 else (-1,))]
    if world >= 4:     # 64 x 8192 fp32 decode rows = 2 MiB per call: the size rule (-1) picks two-shot by itself
        cases.append(("llama3-70b width, 1 layer, B=64 (2 MiB per all-reduce), prefill 8 + 3 steps",
                      LLAMA3_70B.with_layers(1), 64, 8, 3, (-1, 0, 1)))
    out, worst = [], Gate(2)
    mode_name = {-1: "by_size", 0: "one_shot", 1: "two_shot"}
    for name, cfg, B, S, steps, modes in cases:
        plan = make_plan(cfg, world)
        src = SyntheticSource(cfg, 0)
        toks = synthetic_tokens(cfg, B, S)
        gm = Model(plan, rank, src, device=local, max_batch=B, max_ctx=S + steps + 8)
        gm.connect()
        want = None
        if rank == 0:
            from oracle import OracleModel
            om = OracleModel(plan, [shard_weights(src, plan, r) for r in range(world)])
            want, teacher = [om.prefill(toks)], []
            for _ in range(steps):
                t = om.last_tokens.clone()
                teacher.append(t)
                want.append(om.decode(t))
            del om
        else:
            teacher = None
        if dist is not None:
            box = [teacher]
            dist.broadcast_object_list(box, src=0)
            teacher = box[0]
        got = {}
        for m in modes:
            gm.set_ar_mode(m)
            got[m] = [gm.prefill(toks)] + [gm.decode(t) for t in teacher]
        gm.close()
        digest = {m: hashlib.sha256(b"".join(x.numpy().tobytes() for x in got[m])).hexdigest() for m in modes}
        all_digests = [digest]
        if dist is not None:
            all_digests = [None] * world
            dist.all_gather_object(all_digests, digest)
        if rank == 0:
            entry = {"case": name, "tp": world, "ar_modes": [mode_name[m] for m in modes] if world > 1 else ["none (tp=1)"]}
            for m in modes:
                g = Gate(cfg.n_layers)
                for a, b in zip(got[m], want):
                    g.add(a, b)
                    worst.add(a, b)
                entry[f"max_abs_dlogit_{mode_name[m]}"] = g.max_abs
                entry[f"max_rel_sigma_{mode_name[m]}"] = g.max_rel
                entry[f"rms_rel_sigma_{mode_name[m]}"] = g.rms_rel
                g.check(f"bench parity leg: {name}, all-reduce {mode_name[m]}")
            ranks_equal = all(dg == all_digests[0] for dg in all_digests)
            modes_equal = len({digest[m] for m in modes if m != modes[0]} | {digest[modes[0]]}) == 1
            entry["ranks_bit_identical"], entry["ar_modes_bit_identical"] = ranks_equal, modes_equal
            assert ranks_equal, f"bench parity leg: ranks returned different logits ({name})"
            assert modes_equal, f"bench parity leg: all-reduce modes differ bitwise ({name})"
            out.append(entry)
    if rank != 0:
        return None
    modes_seen = sorted({m for e in out for m in e["ar_modes"]})
    return {"vs": "cpu oracle (fp32, all rank shards in-process on rank 0)", "cases": out,
            "max_abs_dlogit": worst.max_abs, "max_rel_sigma": worst.max_rel, "rms_rel_sigma": worst.rms_rel,
            "gates": {"max_abs": LOGIT_TOL, "max_rel_sigma": REL_MAX, "rms_rel_sigma": REL_RMS},
            "ar_modes": modes_seen}


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
    parity = None
    if not args.no_parity and args.shard_of <= 1:
        try:
            parity = parity_leg(dist, world, rank, local)
        except AssertionError as exc:               # a fast wrong answer is not a result
            print(json.dumps({"parity_failure": str(exc)[:400], "n_gpus": world}), flush=True)
            os._exit(3)
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
    # the greedy tokens of the K timed steps must be the same bits on every rank
    import hashlib
    hist = hashlib.sha256(model.read_history(BATCH, K).numpy().tobytes()).hexdigest()
    hists = [hist]
    if dist is not None:
        hists = [None] * world
        dist.all_gather_object(hists, hist)
    if any(h != hists[0] for h in hists):
        print(json.dumps({"parity_failure": "ranks disagree on the greedy tokens of the timed steps", "n_gpus": world}), flush=True)
        os._exit(3)

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
        gemm_shape = (plan.ffn_local * (2 if cfg.gated else 1) * max(1, plan.experts_local), BATCH, cfg.hidden)
        gemm = measure_dominant_gemm(lib, *gemm_shape)
        attn = measure_decode_attention(lib, BATCH, CTX, plan.q_local, plan.kv_local, cfg.head_dim, model.max_ctx)
        # the roofline object describes the kernel with the largest share of the step
        share_gemm = gemm["ms"] * cfg.n_layers / ms_step
        share_attn = attn["kernel_ms"] * cfg.n_layers / ms_step
        if share_attn >= share_gemm:
            kname, kms, kbytes, kshape = attn["kernel"], attn["kernel_ms"], attn["algo_bytes_per_launch"], \
                [BATCH, CTX, plan.q_local, plan.kv_local, cfg.head_dim]
            traffic, tsrc, stale = committed_traffic("attn_decode", kshape, ["attn_decode.cu", "attn_common.cuh"])
            kshare, extra = share_attn, {"n_splits": attn["n_splits"]}
        else:
            kname, kms, kbytes, kshape = gemm["kernel"], gemm["ms"], gemm["algo_bytes"], list(gemm["shape"])
            traffic, tsrc, stale = committed_traffic("gemm_gate_up", kshape, ["gemm_tcgen05.cu", "tcgen05_common.cuh"])
            kshare, extra = share_gemm, {"splits": gemm["splits"]}
        achieved = kbytes / kms / 1e6
        flops = plan.prefill_flops(BATCH, CTX) if hasattr(plan, "prefill_flops") else None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(W, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": gpu_config(cfg.name, cfg.n_layers, tp),
            "e2e": {"value": BATCH * K / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": BATCH * 4,
                    "d2h_bytes_per_step": BATCH * 4},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved / hbm_peak, "traffic": traffic, "traffic_source": tsrc,
                         "traffic_stale": stale, "peak_source": peak_src, "kernel": kname, "shape": kshape,
                         "kernel_ms": kms, "algo_bytes_per_launch": kbytes, "share_of_step": kshare,
                         "launches_per_step": cfg.n_layers, **extra},
            "roofline_gemm": {"kernel": gemm["kernel"], "shape_MNK": gemm["shape"], "splits": gemm["splits"],
                              "kernel_ms": gemm["ms"], "algo_bytes_per_launch": gemm["algo_bytes"],
                              "achieved": gemm["gbs"], "frac": gemm["gbs"] / hbm_peak, "share_of_step": share_gemm},
            "roofline_attention": {**attn, "peak": hbm_peak, "frac": attn["achieved"] / hbm_peak,
                                   "share_of_step": share_attn},
            "step_roofline": {"algo_bytes_per_step_per_gpu": step_bytes, "achieved_gbs_per_gpu": step_bytes / ms_step / 1e6,
                              "frac_of_hbm_peak": step_bytes / ms_step / 1e6 / hbm_peak},
            "prefill": {"tokens_per_s": BATCH * CTX / (prefill_ms / 1e3), "ms": prefill_ms, "tokens": BATCH * CTX},
            "parity": parity,
            "ranks_agree_on_timed_tokens": True,
            "setup_s": setup_s,
        }
        if flops is not None:
            tf_peak = tensor_peak()
            tf = flops / world / (prefill_ms / 1e3) / 1e12
            line["roofline_prefill"] = {"bound": "tensor", "algo_flops": flops, "achieved": tf, "peak": tf_peak[0],
                                        "unit": "TFLOP/s per GPU", "frac": tf / tf_peak[0], "peak_source": tf_peak[1]}
        if args.shard_of > 1:
            line["INVALID"] = (f"profiling run: rank 0's shard of a tp={tp} plan on one GPU with the collectives "
                               "skipped; per-rank compute time only, results are not a model output")
        if not args.no_cpu and world == 1 and args.shard_of <= 1 and args.model == "llama3-8b":
            cpu = CpuDecode()                            # bounded: 1 warm-up + 2 timed full-cost steps (~20 s)
            cpu.step()
            t_cpu = [cpu.step() for _ in range(2)]
            line["cpu_baseline"] = {"value": BATCH * len(t_cpu) / sum(t_cpu), "unit": UNIT, "cores": cpu.cores,
                                    "kind": "port", "sample": cpu.describe(len(t_cpu))}
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
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle parity leg (profiling runs only)")
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
