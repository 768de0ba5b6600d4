"""Host logic of the fused decode step kernel, no GPU: the stream-K partition of a GEMM over the grid
(csrc/decode_step.h: sk_lo / sk_owner, exported as bfly_k_step_gemm_table) and the depth-aware parity gates."""
import ctypes as C

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from butterfly_b200 import _lib
from parity_gates import LOGIT_TOL, REL_MAX, REL_RMS, gates_for


def table(M, K, grid):
    lib = _lib.load()
    m_tiles = (M + 127) // 128
    first = np.zeros(m_tiles, dtype=np.uint16)
    nsl = np.zeros(m_tiles, dtype=np.uint8)
    mx = C.c_int32(0)
    _lib.check(lib.bfly_k_step_gemm_table(M, K, grid, first.ctypes.data, nsl.ctypes.data, C.byref(mx)))
    return first.astype(int), nsl.astype(int), mx.value


def ranges(M, K, grid):
    """The unit range of every CTA, restated independently: [c*U/g, (c+1)*U/g)."""
    m_tiles, nkb = (M + 127) // 128, (K + 63) // 64
    units = m_tiles * nkb
    g = min(grid, units)
    return [(c * units // g, (c + 1) * units // g) for c in range(g)], m_tiles, nkb, units


@settings(max_examples=150, deadline=None)
@given(M=st.integers(1, 40000), K=st.sampled_from([64, 200, 512, 768, 1792, 4096, 8192, 14336]), grid=st.sampled_from([1, 7, 132, 148]))
def test_stream_k_partition_properties(M, K, grid):
    first, nsl, mx = table(M, K, grid)
    rng, m_tiles, nkb, units = ranges(M, K, grid)
    # the ranges tile [0, units) without gaps or overlaps, none is empty, sizes differ by at most one unit
    assert rng[0][0] == 0 and rng[-1][1] == units
    assert all(a[1] == b[0] for a, b in zip(rng, rng[1:]))
    sizes = [b - a for a, b in rng]
    assert min(sizes) >= 1 and max(sizes) - min(sizes) <= 1
    # per tile: the CTAs that own some of its K blocks are exactly first[t] .. first[t] + nsl[t] - 1
    for t in range(m_tiles):
        owners = [c for c, (a, b) in enumerate(rng) if a < (t + 1) * nkb and b > t * nkb]
        assert owners == list(range(first[t], first[t] + nsl[t])), (t, owners, first[t], nsl[t])
    assert mx == max(nsl)


@pytest.mark.parametrize("M,K,tp", [(6144, 4096, 1), (4096, 4096, 1), (28672, 4096, 1), (4096, 14336, 1), (128256, 4096, 1),
                                    (768, 4096, 8), (4096, 512, 8), (3584, 4096, 8), (4096, 1792, 8), (16032, 4096, 8)])
def test_llama3_8b_shapes_keep_the_fold_short(M, K, tp):
    """Slices a consumer phase folds per tile on 148 SMs: a tile of nkb K blocks dealt out in runs of >= per_cta units
    is spread over at most ceil(nkb / per_cta) + 1 CTAs -- 2-6 at tp=1, <= 27 at tp=8 (QKV: 2-3 blocks per CTA)."""
    _, nsl, mx = table(M, K, 148)
    m_tiles, nkb = (M + 127) // 128, (K + 63) // 64
    per_cta = max(1, m_tiles * nkb // min(148, m_tiles * nkb))
    assert mx <= -(-nkb // per_cta) + 1, (M, K, mx)
    assert mx <= (6 if tp == 1 else 27), (M, K, mx)
    assert mx == max(nsl) and min(nsl) >= 1


def test_parity_gates_grow_with_the_square_root_of_depth():
    assert gates_for(1) == (LOGIT_TOL, REL_MAX, REL_RMS) and gates_for(2) == (LOGIT_TOL, REL_MAX, REL_RMS)
    a8, m8, r8 = gates_for(8)
    a32, m32, r32 = gates_for(32)
    assert a8 is None and a32 is None                       # the absolute 1e-2 bar only holds up to ~6 layers
    assert gates_for(6)[0] == LOGIT_TOL
    assert abs(r32 / r8 - 2.0) < 1e-9 and abs(m32 / m8 - 2.0) < 1e-9
    assert 0.05 < r32 < 0.07 and 0.3 < m32 < 0.4            # calibrated: emulation 4.2 % / 21 %, B200 4.6 % / 28 %
