"""Host-side pieces of bench.py that need no GPU: the CPU reference arm's step (the oracle with layer
applications cycling through fewer distinct weight sets when host memory is short), the committed DRAM-traffic
table, and the shared config object that makes the two arms comparable."""
import importlib.util
import json
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)


def test_cpu_arm_cycles_weight_sets_without_changing_the_work():
    from butterfly_b200 import tiny
    cfg = tiny("llama", n_layers=4)
    full = bench.CpuDecode(ctx=40, batch=3, full=cfg, distinct=4)
    cyc = bench.CpuDecode(ctx=40, batch=3, full=cfg, distinct=2)
    assert full.layers == cyc.layers == 4 and cyc.distinct == 2
    assert len(cyc.om.k_cache) == 4 and cyc.om.k_cache[2] is cyc.om.k_cache[0]            # aliased, not copied
    assert cyc.om.w[0]["layers.3.wqkv"] is cyc.om.w[0]["layers.1.wqkv"]
    for arm in (full, cyc):
        assert arm.step() > 0.0
        logits = arm.om.decode(arm.tok)
        assert logits.shape == (3, cfg.vocab) and torch.isfinite(logits).all()
    assert "distinct" in cyc.describe(2)


def test_both_arms_share_one_config_object():
    a, b = bench.gpu_config("llama3-8b", 32, 8), bench.gpu_config("llama3-8b", 32, 8)
    assert a == b and a["parallelism"] == "tp8" and a["layers"] == 32 and "workload" in a


def test_traffic_table_is_well_formed_and_keyed_by_the_kernel_source():
    with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
        table = json.load(f)
    for e in table["entries"]:
        assert e["dram_bytes"] >= e["algo_bytes"] > 0 and e["dram_bytes"] < 1.1 * e["algo_bytes"]   # no wasted re-reads
        assert len(e["source_sha16"]) == 16
    got, src, stale = bench.committed_traffic("attn_decode", [32, 2048, 32, 8, 128], ["attn_decode.cu", "attn_common.cuh"])
    assert got == 276874240 and src and stale in (True, False)
    assert bench.committed_traffic("attn_decode", [1, 2, 3, 4, 5], ["attn_decode.cu"]) == (None, None, None)
