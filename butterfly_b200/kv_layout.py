"""Host-side mirror of the paged KV cache layout (csrc/kernels.cuh: kv_k_offset / kv_v_offset).

One rank's cache of one layer is [n_pages][2 (K,V)][kv_local][page_size*head_dim] bf16.  Inside a
(page, K|V, head) block:
  K : row-major [slot][d]; for d >= 64 the 16-byte chunks of odd slots are XOR-swizzled (chunk ^ 4)
  V : transposed per 32-token sub-block: [slot // 32][d][slot % 32]
The engine writes this layout itself (rope_kv kernel); these helpers exist so tests and tools can build
or inspect pages without going through it.
"""
from __future__ import annotations

import torch


def _odd_row_index(d: int) -> torch.Tensor:
    dd = torch.arange(d)
    return (((dd >> 3) ^ 4) << 3) | (dd & 7)


def pack_pages(k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """k, v: [n_pages, kv_heads, page_size, d] (logical) -> pages [n_pages, 2, kv_heads, page_size*d]."""
    n_pages, kvh, ps, d = k.shape
    assert ps % 32 == 0, "page_size must be a multiple of 32"
    kk = k.clone()
    if d >= 64:
        kk[:, :, 1::2, :] = k[:, :, 1::2, :][..., _odd_row_index(d)]
    vv = v.reshape(n_pages, kvh, ps // 32, 32, d).permute(0, 1, 2, 4, 3).contiguous()
    return torch.stack([kk.reshape(n_pages, kvh, ps * d), vv.reshape(n_pages, kvh, ps * d)], dim=1).contiguous()


def unpack_pages(pages: torch.Tensor, page_size: int, d: int):
    """Inverse of pack_pages: returns (k, v) as [n_pages, kv_heads, page_size, d]."""
    n_pages, _, kvh, _ = pages.shape
    kk = pages[:, 0].reshape(n_pages, kvh, page_size, d).clone()
    if d >= 64:
        kk[:, :, 1::2, :] = pages[:, 0].reshape(n_pages, kvh, page_size, d)[:, :, 1::2, :][..., _odd_row_index(d)]
    vv = pages[:, 1].reshape(n_pages, kvh, page_size // 32, d, 32).permute(0, 1, 2, 4, 3).reshape(n_pages, kvh, page_size, d)
    return kk, vv.contiguous()
