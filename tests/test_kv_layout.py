"""Host mirror of the paged KV layout (butterfly_b200/kv_layout.py) against the offset formulas the kernels use
(csrc/kernels.cuh: kv_k_offset / kv_v_offset, restated here in plain Python) -- CPU only."""
import pytest
import torch

from butterfly_b200.kv_layout import pack_pages, unpack_pages


def k_offset(slot, dd, d):
    chunk = dd >> 3
    if d >= 64:
        chunk ^= (slot & 1) << 2
    return slot * d + (chunk << 3) + (dd & 7)


def v_offset(slot, dd, d):
    return (slot >> 5) * d * 32 + dd * 32 + (slot & 31)


@pytest.mark.parametrize("d", [32, 64, 128])
@pytest.mark.parametrize("ps", [32, 64])
def test_pack_matches_the_device_offsets_and_round_trips(d, ps):
    g = torch.Generator().manual_seed(d * 1000 + ps)
    n_pages, kvh = 3, 2
    k = torch.randn((n_pages, kvh, ps, d), generator=g)
    v = torch.randn((n_pages, kvh, ps, d), generator=g)
    pages = pack_pages(k, v)
    assert pages.shape == (n_pages, 2, kvh, ps * d)
    for page in range(n_pages):
        for head in range(kvh):
            for slot in range(ps):
                for dd in range(0, d, 5):
                    assert pages[page, 0, head, k_offset(slot, dd, d)] == k[page, head, slot, dd]
                    assert pages[page, 1, head, v_offset(slot, dd, d)] == v[page, head, slot, dd]
    k2, v2 = unpack_pages(pages, ps, d)
    assert torch.equal(k2, k) and torch.equal(v2, v)


@pytest.mark.parametrize("d", [32, 64, 128])
def test_offsets_are_a_bijection_inside_a_block(d):
    ps = 64
    ks = {k_offset(s, x, d) for s in range(ps) for x in range(d)}
    vs = {v_offset(s, x, d) for s in range(ps) for x in range(d)}
    assert ks == set(range(ps * d)) and vs == set(range(ps * d))


def test_attention_scratch_sizing_covers_every_row_count():
    """The decode context-split rule (bfly_k_attn_decode_splits) is not monotone in the row count T, so a
    scratch sized from samples of T can be too small for the values in between (the engine now takes the
    maximum over every T; this sweep is the property it relies on, and shows that sampling was unsafe)."""
    from butterfly_b200 import _lib
    lib = _lib.load()
    nq, d = 8, 128
    unsafe = 0
    for kv_local, max_ctx, max_t in ((1, 4096, 192), (1, 640, 128), (2, 2048, 96), (8, 2048, 64)):
        need = [t * nq * lib.bfly_k_attn_decode_splits(t, kv_local, max_ctx) * d for t in range(1, max_t + 1)]
        assert min(need) > 0
        cap_all = max(need)
        sampled = [t for t in range(1, max_t + 1) if t < 64 or t % 64 == 0 or t == max_t]      # round-1 rule
        cap_sampled = max(need[t - 1] for t in sampled)
        assert all(n <= cap_all for n in need)
        unsafe += int(any(n > cap_sampled for n in need))
    assert unsafe >= 1, "expected at least one geometry where sampled sizing under-allocates"
    assert lib.bfly_k_attn_decode_splits(0, 1, 128) == 0
