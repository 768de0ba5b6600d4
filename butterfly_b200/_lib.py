"""ctypes binding of libbutterfly_b200.so (include/butterfly_b200.h).  No fallback: if the library is
absent it is built with nvcc, and if that fails the import of the compute path raises."""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

i32, i64, u64, f32 = C.c_int32, C.c_int64, C.c_uint64, C.c_float
vp, cp = C.c_void_p, C.c_char_p


class BflyError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"butterfly_b200 error {code}: {msg}")
        self.code = code


class ModelDesc(C.Structure):
    _fields_ = [("family", i32), ("hidden", i32), ("n_layers", i32), ("n_heads", i32), ("n_kv_heads", i32),
                ("head_dim", i32), ("ffn", i32), ("vocab", i32), ("n_experts", i32), ("top_k", i32),
                ("norm_eps", f32), ("rope_theta", f32), ("max_pos", i32), ("tied_head", i32), ("bias", i32),
                ("tp", i32), ("rank", i32), ("page_size", i32), ("max_batch", i32), ("max_ctx", i32),
                ("max_tokens", i32)]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", u64), ("weight_bytes", u64), ("kv_bytes", u64), ("workspace_bytes", u64),
                ("sm_count", i32), ("graph_nodes_decode", i32)]


# name -> (restype, argtypes); must list every symbol declared in include/butterfly_b200.h
SIGNATURES = {
    "bfly_version": (i32, []),
    "bfly_last_error": (cp, [vp]),
    "bfly_create": (i32, [C.POINTER(ModelDesc), i32, C.POINTER(vp)]),
    "bfly_destroy": (i32, [vp]),
    "bfly_get_stats": (i32, [vp, C.POINTER(Stats)]),
    "bfly_upload_weight": (i32, [vp, cp, vp, i64, i64]),
    "bfly_alloc_weight": (i32, [vp, cp, i64, i64]),
    "bfly_fill_weight_block": (i32, [vp, cp, i64, i64, u64, i64, i64, i64, i64, i64, f32, f32]),
    "bfly_read_weight": (i32, [vp, cp, vp, i64]),
    "bfly_weights_ready": (i32, [vp]),
    "bfly_comm_unique_id": (i32, [vp]),
    "bfly_comm_init": (i32, [vp, vp, i32, i32]),
    "bfly_set_page_table": (i32, [vp, vp, i32, i32]),
    "bfly_set_ar_mode": (i32, [vp, i32]),
    "bfly_bench_allreduce": (i32, [vp, i32, i32, C.POINTER(f32), C.POINTER(f32)]),
    "bfly_prefill": (i32, [vp, vp, i32, i32, vp, vp]),
    "bfly_prefill_at": (i32, [vp, i32, vp, i32, i32, vp, vp]),
    "bfly_decode": (i32, [vp, vp, i32, vp, vp]),
    "bfly_decode_greedy": (i32, [vp, i32, i32, vp]),
    "bfly_decode_enqueue": (i32, [vp, i32, i32]),
    "bfly_read_history": (i32, [vp, i32, i32, vp]),
    "bfly_read_step_trace": (i32, [vp, vp, i32, C.POINTER(i32), C.POINTER(i32)]),
    "bfly_sync": (i32, [vp]),
    "bfly_stream": (vp, [vp]),
    "bfly_k_fill_synth": (i32, [vp, i64, u64, i64, i64, i64, i64, i64, f32, f32, vp]),
    "bfly_k_embed": (i32, [vp, vp, vp, vp, i32, vp, i32, i32, i32, vp]),
    "bfly_k_add_norm": (i32, [vp, vp, i32, i64, vp, vp, vp, f32, i32, vp, i32, i32, vp]),
    "bfly_k_rope_kv": (i32, [vp, i32, i64, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, i32, i32, vp]),
    "bfly_k_attn_decode": (i32, [vp, vp, vp, i32, i32, i32, i32, vp, i32, i32, i32, vp, vp, i32, vp, vp]),
    "bfly_k_attn_decode_rope": (i32, [vp, i32, i64, vp, vp, vp, vp, i32, i32, i32, i32, vp, i32, i32, vp, vp, i32, vp, vp]),
    "bfly_k_step_gemm_table": (i32, [i32, i32, i32, vp, vp, C.POINTER(i32)]),
    "bfly_k_attn_decode_splits": (i32, [i32, i32, i32]),
    "bfly_k_attn_prefill": (i32, [vp, vp, vp, i32, i32, i32, i32, vp, i32, i32, i32, vp, vp]),
    "bfly_k_attn_decode_static": (i32, [vp, vp, vp, i32, i32, i32, i32, vp, i32, i32, i32, i32, vp, vp]),
    "bfly_k_act": (i32, [vp, i32, i64, vp, i32, i32, i32, vp, vp, i32, vp]),
    "bfly_k_lm_finalize": (i32, [vp, i32, i64, i32, i32, i32, vp, i64, vp, vp, vp]),
    "bfly_k_moe_router": (i32, [vp, vp, i32, i32, i32, i32, i32, i32, vp, vp]),
    "bfly_k_gemm": (i32, [vp, vp, i32, i32, i32, i64, i32, i32, i32, vp, i64, vp, i64, C.POINTER(i32), vp]),
    "bfly_k_gemm_swiglu": (i32, [vp, vp, i32, i32, i32, i32, vp, i32, vp, vp]),
    "bfly_k_gemm_plan": (i32, [i32, i32, i32, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
}

_lock = threading.Lock()
_lib = None


def lib_path() -> str:
    return _build.LIB


def load():
    """Load (building first if needed) and type every entry point.  Raises if anything is missing."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = _build.LIB
        override = os.environ.get("BFLY_LIB")     # A/B profiling of another build of the same ABI
        if override:
            path = override
        elif not os.path.exists(path):
            path = _build.build_library()
        lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
        for name, (res, args) in SIGNATURES.items():
            if override and not hasattr(lib, name):
                continue
            fn = getattr(lib, name)          # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        _lib = lib
        return lib


def check(rc: int, ctx=None):
    if rc != 0:
        msg = load().bfly_last_error(ctx)
        raise BflyError(rc, msg.decode() if msg else "unknown error")
