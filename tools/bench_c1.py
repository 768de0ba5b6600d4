#!/usr/bin/env python
"""Isolated latency of C1, the decode all-reduce fused with residual add + norm.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_c1.py

For the row shapes of the BASELINE.json configs (32 x 4096 = Llama-3-8B / Mixtral decode, 64 x 8192 = Llama-3-70B)
prints the average time of ar_add_norm_kernel (one-shot and two-shot) next to add_norm_kernel on the same rows;
the difference is what the exchange over NVLink costs.  Rank 0 prints; max over ranks."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from butterfly_b200 import LLAMA3_8B, LLAMA3_70B, Model, SyntheticSource, make_plan  # noqa: E402


def main():
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    variants = [("epoch flags", {"BFLY_AR_PROTO": "flags", "BFLY_AR_CLUSTER": "", "BFLY_AR_BF16": ""}),
                ("flag-free (self-validating words)", {"BFLY_AR_PROTO": "", "BFLY_AR_CLUSTER": "", "BFLY_AR_BF16": ""})]
    if os.environ.get("C1_ALL"):
        variants += [("epoch flags, rows split over a cluster", {"BFLY_AR_PROTO": "flags", "BFLY_AR_CLUSTER": "a", "BFLY_AR_BF16": ""}),
                     ("epoch flags, cluster, bf16 payload", {"BFLY_AR_PROTO": "flags", "BFLY_AR_CLUSTER": "a", "BFLY_AR_BF16": "1"})]
    for cfg, T in ((LLAMA3_8B.with_layers(1), 32), (LLAMA3_70B.with_layers(1), 64)):
        for vname, env in variants:
            for k, v in env.items():                 # the engine reads these knobs when the context is created
                if v:
                    os.environ[k] = v
                else:
                    os.environ.pop(k, None)
            m = Model(make_plan(cfg, world), rank, SyntheticSource(cfg, 0), device=local, max_batch=T, max_ctx=64)
            m.connect()
            for mode, name in ((0, "one-shot"), (1, "two-shot")):
                m.set_ar_mode(mode)
                dist.barrier()
                a, b = C.c_float(0), C.c_float(0)
                m._check(m.lib.bfly_bench_allreduce(m.ctx, T, int(os.environ.get('C1_ITERS', '2000')), C.byref(a), C.byref(b)))
                t = torch.tensor([a.value, b.value], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                if rank == 0:
                    print(f"tp={world} rows {T} x {cfg.hidden} ({T * cfg.hidden * 4 // 1024} KiB fp32/rank) {vname}, {name}: "
                          f"ar_add_norm {t[0].item():.2f} us, add_norm {t[1].item():.2f} us, exchange {t[0].item() - t[1].item():.2f} us",
                          flush=True)
            m.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
