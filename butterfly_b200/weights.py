"""Weight sources and per-rank sharding (`shard_weights`).

The reference names a partitioner but has no weight format or loader (ref:CLAUDE.md:21,
0 lines of source), so the logical tensor set below is this repo's definition
(SURVEY.md section 8a rows a2/a5/a9/a11/a12/a13/a14).

Logical (unsharded) tensors, all [out_features, in_features] row-major, bf16:
  global : embed[V,h]  pos_embed[P,h] (GPT-2)  final_norm.w[h] (.b GPT-2)  lm_head[V,h] (untied)
  layer  : attn_norm.w  wq[nh*d,h] wk[nkv*d,h] wv[nkv*d,h]  wo[h,nh*d]  ffn_norm.w
           dense : w_gate[f,h] (gated only)  w_up[f,h]  w_down[h,f]
           MoE   : router[E,h]  e{i}.w_gate  e{i}.w_up  e{i}.w_down   (a rank's experts are stored
                   concatenated: w_gate_up[E_l*2f,h] = [gate_e|up_e] blocks, w_down[h,E_l*f])
           GPT-2 : attn_norm.b bq bk bv bo ffn_norm.b b_up b_down

`shard_spec` says how each rank tensor is assembled from rectangular blocks of the logical
tensors; `shard_weights` executes that spec on the host, and the B200 engine executes the same
spec on the device (butterfly_b200/model.py), so both sides hold identical bits.

Synthetic init (frozen recipe, seeds: weights 0, tokens 1 -- BASELINE.md "Parity protocol"):
every element is a pure function of (seed, layer, tensor name, row, col): a splitmix64 hash of the
element index, the four 16-bit fields of which are summed (Irwin-Hall, n=4: bell-shaped, bounded
at +-3.46 sigma), centred, converted to fp32, multiplied by one fp32 scale, optionally added to
1.0 (norm gains) and rounded to bf16.  Integer arithmetic plus one IEEE multiply/add means the
CUDA fill kernel (csrc/synth.cu) reproduces the host values bit for bit, so a 70B-shaped rank can
create its shard directly in HBM without any host copy of the full model.
"""
from __future__ import annotations

import math
import os
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from .config import ModelConfig
from .plan import PartitionPlan

_IH4_STD = math.sqrt((2.0 ** 32 - 1.0) / 3.0)   # std of the sum of four uniform 16-bit integers
_GOLDEN = 0x9E3779B97F4A7C15
_M1 = 0xBF58476D1CE4E5B9
_M2 = 0x94D049BB133111EB
_MASK64 = (1 << 64) - 1


def tensor_key(seed: int, layer: int, name: str) -> int:
    """Stable 64-bit key for one logical tensor (FNV-1a over 'seed/layer/name')."""
    h = 0xCBF29CE484222325
    for b in f"{seed}/{layer}/{name}".encode():
        h = ((h ^ b) * 0x100000001B3) & _MASK64
    return h


def synth_scale(std: float) -> np.float32:
    return np.float32(std / _IH4_STD)


def synth_block(key: int, ncols_full: int, r0: int, r1: int, c0: int, c1: int,
                scale: np.float32, offset: float = 0.0) -> torch.Tensor:
    """bf16 values of rows [r0,r1) x cols [c0,c1) of the logical tensor identified by `key`."""
    out = torch.empty((r1 - r0, c1 - c0), dtype=torch.bfloat16)
    if out.numel() == 0:
        return out
    cols = np.arange(c0, c1, dtype=np.uint64)[None, :]
    step = max(1, (1 << 21) // max(1, c1 - c0))

    def fill(a: int):
        b = min(r1, a + step)
        with np.errstate(over="ignore"):
            rows = np.arange(a, b, dtype=np.uint64)[:, None]
            z = np.uint64(key) + (rows * np.uint64(ncols_full) + cols) * np.uint64(_GOLDEN)
            z = (z ^ (z >> np.uint64(30))) * np.uint64(_M1)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(_M2)
            z = z ^ (z >> np.uint64(31))
            m = np.uint64(0xFFFF)
            s = (z & m) + ((z >> np.uint64(16)) & m) + ((z >> np.uint64(32)) & m) + (z >> np.uint64(48))
            v = (s.astype(np.int64) - 131070).astype(np.float32) * scale
            if offset != 0.0:
                v = v + np.float32(offset)
        out[a - r0:b - r0] = torch.from_numpy(v).to(torch.bfloat16)

    starts = list(range(r0, r1, step))
    if len(starts) > 4:                        # numpy releases the GIL inside its ufuncs
        with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as ex:
            list(ex.map(fill, starts))
    else:
        for a in starts:
            fill(a)
    return out


# ----------------------------------------------------------------------------------------------
# logical tensor catalogue
# ----------------------------------------------------------------------------------------------

@dataclass(frozen=True)
class TensorInfo:
    rows: int
    cols: int          # 1-D tensors are [rows, 1]
    std: float
    offset: float = 0.0


def logical_tensors(cfg: ModelConfig) -> Tuple[Dict[str, TensorInfo], Dict[str, TensorInfo]]:
    """(global tensors, per-layer tensors) with their init statistics."""
    h, f, d = cfg.hidden, cfg.ffn, cfg.head_dim
    out_std = 0.02 / math.sqrt(2.0 * cfg.n_layers)
    # Oracle logit std ~0.1.  Calibrated on CPU with OracleModel(emulate_bf16=True) against the fp32
    # oracle: the bf16 numerics contract costs ~1-1.4 % of the logit std (rms) at Llama-3-8B width,
    # i.e. a max over B x 128256 logits of ~7 % of the std, so 0.1 keeps the absolute north-star bar
    # (max |dlogit| <= 1e-2) meaningful at full depth (BASELINE.md "Parity protocol").
    head_std = 0.1 / math.sqrt(h)
    g: Dict[str, TensorInfo] = {
        "embed": TensorInfo(cfg.vocab, h, head_std if cfg.tied_head else 0.02),
        "final_norm.w": TensorInfo(h, 1, 0.02, 1.0),
    }
    if cfg.rope_theta == 0.0:
        g["pos_embed"] = TensorInfo(cfg.max_pos, h, 0.01)
    if cfg.layer_norm:
        g["final_norm.b"] = TensorInfo(h, 1, 0.02)
    if not cfg.tied_head:
        g["lm_head"] = TensorInfo(cfg.vocab, h, head_std)
    l: Dict[str, TensorInfo] = {
        "attn_norm.w": TensorInfo(h, 1, 0.02, 1.0),
        "ffn_norm.w": TensorInfo(h, 1, 0.02, 1.0),
        "wq": TensorInfo(cfg.q_dim, h, 0.02),
        "wk": TensorInfo(cfg.kv_dim, h, 0.02),
        "wv": TensorInfo(cfg.kv_dim, h, 0.02),
        "wo": TensorInfo(h, cfg.q_dim, out_std),
    }
    if cfg.layer_norm:
        l["attn_norm.b"] = TensorInfo(h, 1, 0.02)
        l["ffn_norm.b"] = TensorInfo(h, 1, 0.02)
    if cfg.bias:
        l.update({"bq": TensorInfo(cfg.q_dim, 1, 0.02), "bk": TensorInfo(cfg.kv_dim, 1, 0.02),
                  "bv": TensorInfo(cfg.kv_dim, 1, 0.02), "bo": TensorInfo(h, 1, 0.02),
                  "b_up": TensorInfo(f, 1, 0.02), "b_down": TensorInfo(h, 1, 0.02)})
    if cfg.is_moe:
        l["router"] = TensorInfo(cfg.n_experts, h, 0.02)
        for e in range(cfg.n_experts):
            l[f"e{e}.w_gate"] = TensorInfo(f, h, 0.02)
            l[f"e{e}.w_up"] = TensorInfo(f, h, 0.02)
            l[f"e{e}.w_down"] = TensorInfo(h, f, out_std)
    else:
        if cfg.gated:
            l["w_gate"] = TensorInfo(f, h, 0.02)
        l["w_up"] = TensorInfo(f, h, 0.02)
        l["w_down"] = TensorInfo(h, f, out_std)
    return g, l


GLOBAL_LAYER = -1   # `layer` value used for global tensors


class SyntheticSource:
    """Seeded synthetic weights; any block of any logical tensor on demand."""

    def __init__(self, cfg: ModelConfig, seed: int = 0):
        self.cfg, self.seed = cfg, seed
        self._g, self._l = logical_tensors(cfg)

    def info(self, name: str, layer: int) -> TensorInfo:
        return self._g[name] if layer == GLOBAL_LAYER else self._l[name]

    def key(self, name: str, layer: int) -> int:
        return tensor_key(self.seed, layer, name)

    def block(self, name: str, layer: int, r0: int, r1: int, c0: int, c1: int) -> torch.Tensor:
        t = self.info(name, layer)
        return synth_block(self.key(name, layer), t.cols, r0, r1, c0, c1, synth_scale(t.std), t.offset)

    def full(self, name: str, layer: int = GLOBAL_LAYER) -> torch.Tensor:
        t = self.info(name, layer)
        return self.block(name, layer, 0, t.rows, 0, t.cols)


class DictSource:
    """Weights given as full logical tensors: {(layer, name): tensor[rows, cols]} (any float dtype).

    This is the entry for externally supplied weights (HF cross-check in tests; a checkpoint
    loader would produce the same mapping)."""

    def __init__(self, cfg: ModelConfig, tensors: Dict[Tuple[int, str], torch.Tensor]):
        self.cfg = cfg
        self._t = {k: (v if v.dim() == 2 else v.reshape(-1, 1)).to(torch.bfloat16) for k, v in tensors.items()}
        self._g, self._l = logical_tensors(cfg)
        for (layer, name), v in self._t.items():
            t = self.info(name, layer)
            if tuple(v.shape) != (t.rows, t.cols):
                raise ValueError(f"{name}@{layer}: shape {tuple(v.shape)} != {(t.rows, t.cols)}")

    def info(self, name: str, layer: int) -> TensorInfo:
        return self._g[name] if layer == GLOBAL_LAYER else self._l[name]

    def block(self, name: str, layer: int, r0: int, r1: int, c0: int, c1: int) -> torch.Tensor:
        return self._t[(layer, name)][r0:r1, c0:c1].contiguous()

    def full(self, name: str, layer: int = GLOBAL_LAYER) -> torch.Tensor:
        return self._t[(layer, name)]


# ----------------------------------------------------------------------------------------------
# shard spec
# ----------------------------------------------------------------------------------------------

@dataclass(frozen=True)
class Piece:
    """dst rows [dst_row, dst_row + r1-r0) of a rank tensor = block [r0:r1, c0:c1] of a logical tensor."""
    name: str
    layer: int
    r0: int
    r1: int
    c0: int
    c1: int
    dst_row: int
    dst_col: int = 0


@dataclass(frozen=True)
class ShardTensor:
    name: str                 # rank-tensor name, e.g. "layers.3.wqkv"
    rows: int
    cols: int
    pieces: Tuple[Piece, ...]


def shard_spec(plan: PartitionPlan, rank: int) -> List[ShardTensor]:
    cfg, rp = plan.cfg, plan.rank(rank)
    h, d, f = cfg.hidden, cfg.head_dim, cfg.ffn
    g_info, l_info = logical_tensors(cfg)
    out: List[ShardTensor] = []

    def whole(dst: str, name: str, layer: int):
        t = g_info[name] if layer == GLOBAL_LAYER else l_info[name]
        out.append(ShardTensor(dst, t.rows, t.cols, (Piece(name, layer, 0, t.rows, 0, t.cols, 0),)))

    whole("embed", "embed", GLOBAL_LAYER)
    if "pos_embed" in g_info:
        whole("pos_embed", "pos_embed", GLOBAL_LAYER)
    whole("final_norm.w", "final_norm.w", GLOBAL_LAYER)
    if cfg.layer_norm:
        whole("final_norm.b", "final_norm.b", GLOBAL_LAYER)
    v0, v1 = rp.vocab_rows
    out.append(ShardTensor("lm_head", v1 - v0, h,
                           (Piece("embed" if cfg.tied_head else "lm_head", GLOBAL_LAYER, v0, v1, 0, h, 0),)))

    q0, q1 = rp.q_heads[0] * d, rp.q_heads[1] * d
    k0, k1 = rp.kv_heads[0] * d, rp.kv_heads[1] * d
    nq, nk = q1 - q0, k1 - k0
    for l in range(cfg.n_layers):
        p = f"layers.{l}."
        whole(p + "attn_norm.w", "attn_norm.w", l)
        whole(p + "ffn_norm.w", "ffn_norm.w", l)
        if cfg.layer_norm:
            whole(p + "attn_norm.b", "attn_norm.b", l)
            whole(p + "ffn_norm.b", "ffn_norm.b", l)
        out.append(ShardTensor(p + "wqkv", nq + 2 * nk, h, (
            Piece("wq", l, q0, q1, 0, h, 0), Piece("wk", l, k0, k1, 0, h, nq),
            Piece("wv", l, k0, k1, 0, h, nq + nk))))
        out.append(ShardTensor(p + "wo", h, nq, (Piece("wo", l, 0, h, q0, q1, 0),)))
        if cfg.bias:
            out.append(ShardTensor(p + "bqkv", nq + 2 * nk, 1, (
                Piece("bq", l, q0, q1, 0, 1, 0), Piece("bk", l, k0, k1, 0, 1, nq),
                Piece("bv", l, k0, k1, 0, 1, nq + nk))))
            whole(p + "bo", "bo", l)
            whole(p + "b_down", "b_down", l)
        if cfg.is_moe:
            whole(p + "router", "router", l)
            # local experts are concatenated: w_gate_up rows = [gate_e | up_e] per expert block, w_down
            # columns = expert blocks, so the expert FFNs run as ONE column- and ONE row-parallel GEMM
            ne = len(rp.experts)
            gu, dn = [], []
            for j, e in enumerate(rp.experts):
                gu += [Piece(f"e{e}.w_gate", l, 0, f, 0, h, j * 2 * f), Piece(f"e{e}.w_up", l, 0, f, 0, h, j * 2 * f + f)]
                dn.append(Piece(f"e{e}.w_down", l, 0, h, 0, f, 0, j * f))
            out.append(ShardTensor(p + "w_gate_up", ne * 2 * f, h, tuple(gu)))
            out.append(ShardTensor(p + "w_down", h, ne * f, tuple(dn)))
        else:
            f0, f1 = rp.ffn_cols
            nf = f1 - f0
            if cfg.gated:
                out.append(ShardTensor(p + "w_gate_up", 2 * nf, h, (
                    Piece("w_gate", l, f0, f1, 0, h, 0), Piece("w_up", l, f0, f1, 0, h, nf))))
            else:
                out.append(ShardTensor(p + "w_up", nf, h, (Piece("w_up", l, f0, f1, 0, h, 0),)))
                out.append(ShardTensor(p + "b_up", nf, 1, (Piece("b_up", l, f0, f1, 0, 1, 0),)))
            out.append(ShardTensor(p + "w_down", h, nf, (Piece("w_down", l, 0, h, f0, f1, 0),)))
    return out


class RankWeights:
    """One rank's bf16 shards on the host, keyed by rank-tensor name."""

    def __init__(self, plan: PartitionPlan, rank: int, tensors: Dict[str, torch.Tensor]):
        self.plan, self.rank, self.tensors = plan, rank, tensors

    def __getitem__(self, name: str) -> torch.Tensor:
        return self.tensors[name]

    def get(self, name: str) -> Optional[torch.Tensor]:
        return self.tensors.get(name)

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in self.tensors.values())


def shard_weights(source, plan: PartitionPlan, rank: int) -> RankWeights:
    """Materialise rank `rank`'s shards of `source` (SyntheticSource | DictSource) on the host."""
    tensors: Dict[str, torch.Tensor] = {}
    for st in shard_spec(plan, rank):
        t = torch.empty((st.rows, st.cols), dtype=torch.bfloat16)
        for pc in st.pieces:
            t[pc.dst_row:pc.dst_row + (pc.r1 - pc.r0), pc.dst_col:pc.dst_col + (pc.c1 - pc.c0)] = \
                source.block(pc.name, pc.layer, pc.r0, pc.r1, pc.c0, pc.c1)
        tensors[st.name] = t.reshape(-1) if st.cols == 1 else t
    return RankWeights(plan, rank, tensors)


def synthetic_tokens(cfg: ModelConfig, batch: int, seq: int, seed: int = 1) -> torch.Tensor:
    """Seeded prompt tokens int32[batch, seq] (tokens seed 1 is the frozen protocol value)."""
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, cfg.vocab, (batch, seq), generator=g, dtype=torch.int64).to(torch.int32)
