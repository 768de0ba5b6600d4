"""Host-side Model: the thin Python layer above the C ABI.

    plan  = make_plan(cfg, tp)
    model = Model(plan, rank, shard_weights(src, plan, rank) | SyntheticSource(cfg), ...)
    logits = model.prefill(tokens[B,S]);  logits = model.decode(tokens[B]);  toks = model.decode_greedy(n)

Same call shapes as oracle.OracleModel so the parity tests read the same on both sides.  All
compute is in libbutterfly_b200.so; this file only marshals pointers.  One process per GPU rank.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Union

import numpy as np
import torch

from . import _lib
from .config import FAMILY_GPT2, FAMILY_LLAMA, FAMILY_MIXTRAL
from .plan import PartitionPlan
from .weights import RankWeights, SyntheticSource, shard_spec, synth_scale

_FAMILY = {FAMILY_GPT2: 0, FAMILY_LLAMA: 1, FAMILY_MIXTRAL: 2}


class Model:
    def __init__(self, plan: PartitionPlan, rank: int, weights: Union[RankWeights, SyntheticSource], *,
                 device: int = 0, max_batch: int = 32, max_ctx: int = 2048, max_tokens: Optional[int] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("butterfly_b200.Model needs a CUDA device (B200); there is no CPU fallback")
        self.lib = _lib.load()
        self.plan, self.cfg, self.rank, self.device = plan, plan.cfg, rank, device
        self.max_batch, self.max_ctx = max_batch, max_ctx
        cfg = plan.cfg
        if max_tokens is None:
            max_tokens = max(max_batch, min(8192, max_batch * max_ctx))
        self.max_tokens = max_tokens
        desc = _lib.ModelDesc(
            family=_FAMILY[cfg.family], hidden=cfg.hidden, n_layers=cfg.n_layers, n_heads=cfg.n_heads,
            n_kv_heads=cfg.n_kv_heads, head_dim=cfg.head_dim, ffn=cfg.ffn, vocab=cfg.vocab,
            n_experts=cfg.n_experts, top_k=cfg.top_k, norm_eps=cfg.norm_eps, rope_theta=cfg.rope_theta,
            max_pos=cfg.max_pos, tied_head=int(cfg.tied_head), bias=int(cfg.bias), tp=plan.tp, rank=rank,
            page_size=plan.page_size, max_batch=max_batch, max_ctx=max_ctx, max_tokens=max_tokens)
        ctx = C.c_void_p()
        _lib.check(self.lib.bfly_create(C.byref(desc), device, C.byref(ctx)))
        self.ctx = ctx
        self.last_tokens: Optional[torch.Tensor] = None
        try:
            if isinstance(weights, RankWeights):
                self._upload(weights)
            else:
                self._synthesize(weights)
            self._check(self.lib.bfly_weights_ready(self.ctx))
        except Exception:
            self.close()
            raise

    # ---- lifecycle ------------------------------------------------------------------------------
    def _check(self, rc: int):
        _lib.check(rc, self.ctx)

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.bfly_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- weights --------------------------------------------------------------------------------
    def _upload(self, rw: RankWeights):
        assert rw.rank == self.rank and rw.plan.tp == self.plan.tp
        for name, t in rw.tensors.items():
            t2 = (t if t.dim() == 2 else t.reshape(-1, 1)).contiguous()
            assert t2.dtype == torch.bfloat16
            if t2.numel() == 0:
                continue
            self._check(self.lib.bfly_upload_weight(self.ctx, name.encode(), t2.data_ptr(), t2.shape[0], t2.shape[1]))

    def _synthesize(self, src: SyntheticSource):
        """Create this rank's shards directly in HBM from the synthetic source (no host copy)."""
        for st in shard_spec(self.plan, self.rank):
            if st.rows * st.cols == 0:
                continue
            self._check(self.lib.bfly_alloc_weight(self.ctx, st.name.encode(), st.rows, st.cols))
            for pc in st.pieces:
                info = src.info(pc.name, pc.layer)
                self._check(self.lib.bfly_fill_weight_block(
                    self.ctx, st.name.encode(), pc.dst_row, pc.dst_col, src.key(pc.name, pc.layer), info.cols,
                    pc.r0, pc.c0, pc.r1 - pc.r0, pc.c1 - pc.c0, float(synth_scale(info.std)), float(info.offset)))

    def read_weight(self, name: str, rows: int, cols: int) -> torch.Tensor:
        out = torch.empty((rows, cols), dtype=torch.bfloat16)
        self._check(self.lib.bfly_read_weight(self.ctx, name.encode(), out.data_ptr(), out.numel() * 2))
        return out

    # ---- communication --------------------------------------------------------------------------
    def connect(self, group=None):
        """Create the NCCL communicator of the TP group; the unique id travels over torch.distributed."""
        if self.plan.tp == 1:
            return
        import torch.distributed as dist
        assert dist.is_initialized() and dist.get_world_size(group) == self.plan.tp
        ident = np.zeros(128, dtype=np.uint8)
        if dist.get_rank(group) == 0:
            _lib.check(self.lib.bfly_comm_unique_id(ident.ctypes.data))
        obj = [ident.tobytes()]
        dist.broadcast_object_list(obj, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        buf = (C.c_char * 128).from_buffer_copy(obj[0])
        self._check(self.lib.bfly_comm_init(self.ctx, buf, self.plan.tp, self.rank))

    def set_ar_mode(self, mode: int):
        """Decode all-reduce over peer memory: -1 by size, 0 one-shot, 1 two-shot; the same on every rank."""
        self._check(self.lib.bfly_set_ar_mode(self.ctx, int(mode)))

    # ---- forward --------------------------------------------------------------------------------
    def set_page_table(self, table: torch.Tensor):
        t = table.to(torch.int32).contiguous()
        self._check(self.lib.bfly_set_page_table(self.ctx, t.data_ptr(), t.shape[0], t.shape[1]))

    def prefill(self, tokens: torch.Tensor, want_logits: bool = True) -> Optional[torch.Tensor]:
        """tokens int32[B,S] -> fp32 logits[B,V] of the last position (host tensor)."""
        tok = tokens.to(torch.int32).contiguous()
        B, S = tok.shape
        logits = torch.empty((B, self.cfg.vocab), dtype=torch.float32) if want_logits else None
        nxt = torch.empty((B,), dtype=torch.int32)
        self._check(self.lib.bfly_prefill(self.ctx, tok.data_ptr(), B, S,
                                          logits.data_ptr() if want_logits else None, nxt.data_ptr()))
        self.last_tokens = nxt
        return logits

    def prefill_at(self, seq0: int, tokens: torch.Tensor, want_logits: bool = True) -> Optional[torch.Tensor]:
        """Ragged batches: restart sequences seq0 .. seq0+n-1 with tokens int32[n,S]; the others keep their cache."""
        tok = tokens.to(torch.int32).contiguous()
        n, S = tok.shape
        logits = torch.empty((n, self.cfg.vocab), dtype=torch.float32) if want_logits else None
        nxt = torch.empty((n,), dtype=torch.int32)
        self._check(self.lib.bfly_prefill_at(self.ctx, seq0, tok.data_ptr(), n, S,
                                             logits.data_ptr() if want_logits else None, nxt.data_ptr()))
        if self.last_tokens is None or self.last_tokens.numel() < seq0 + n:
            grown = torch.zeros((seq0 + n,), dtype=torch.int32)
            if self.last_tokens is not None:
                grown[:self.last_tokens.numel()] = self.last_tokens
            self.last_tokens = grown
        self.last_tokens[seq0:seq0 + n] = nxt
        return logits

    def decode(self, tokens: Optional[torch.Tensor], want_logits: bool = True) -> Optional[torch.Tensor]:
        """tokens int32[B] -> fp32 logits[B,V]; tokens=None continues from the engine's greedy tokens."""
        B = tokens.numel() if tokens is not None else self.last_tokens.numel()
        tok = tokens.to(torch.int32).contiguous() if tokens is not None else None
        logits = torch.empty((B, self.cfg.vocab), dtype=torch.float32) if want_logits else None
        nxt = torch.empty((B,), dtype=torch.int32)
        self._check(self.lib.bfly_decode(self.ctx, tok.data_ptr() if tok is not None else None, B,
                                         logits.data_ptr() if want_logits else None, nxt.data_ptr()))
        self.last_tokens = nxt
        return logits

    def decode_greedy(self, n_steps: int, batch: Optional[int] = None) -> torch.Tensor:
        """n free-running greedy steps on the device: int32[B, n_steps] = the token fed at each step."""
        B = batch if batch is not None else self.last_tokens.numel()
        out = torch.empty((B, n_steps), dtype=torch.int32)
        self._check(self.lib.bfly_decode_greedy(self.ctx, B, n_steps, out.data_ptr()))
        return out

    def decode_enqueue(self, batch: int, n_steps: int):
        self._check(self.lib.bfly_decode_enqueue(self.ctx, batch, n_steps))

    def read_history(self, batch: int, n_steps: int) -> torch.Tensor:
        """Tokens fed at each step of the last decode_enqueue / decode_greedy: int32[B, n_steps]."""
        out = torch.empty((batch, n_steps), dtype=torch.int32)
        self._check(self.lib.bfly_read_history(self.ctx, batch, n_steps, out.data_ptr()))
        return out

    def sync(self):
        self._check(self.lib.bfly_sync(self.ctx))

    @property
    def stream(self) -> torch.cuda.ExternalStream:
        return torch.cuda.ExternalStream(self.lib.bfly_stream(self.ctx), device=self.device)

    def stats(self) -> dict:
        st = _lib.Stats()
        self._check(self.lib.bfly_get_stats(self.ctx, C.byref(st)))
        return {k: getattr(st, k) for k, _ in st._fields_}
