"""Host-side dispatch of the conv1d kernels (``cs_conv1d_plan``: pure arithmetic, no GPU needed): resource and coverage
invariants of the chosen tile configuration over random shapes."""
import ctypes

import pytest
from hypothesis import given, settings, strategies as st

from crossscale.ops import _lib

KB = 1024


def plan(B, L, Cin, Cout, K, pad, direction):
    out = (ctypes.c_int * 16)()
    _lib.check(_lib.cuda_lib().cs_conv1d_plan(B, L, Cin, Cout, K, pad, direction, out), "cs_conv1d_plan")
    return list(out)


shapes = st.tuples(st.integers(1, 5000), st.integers(1, 2000), st.sampled_from([16, 32, 48, 64, 96, 128, 192, 256, 512, 768, 1024]),
                   st.sampled_from([16, 32, 64, 96, 128, 256, 384, 512, 768]), st.integers(1, 33))


@settings(max_examples=400, deadline=None)
@given(shapes, st.booleans())
def test_forward_and_dgrad_plans_fit_the_sm(shape, dgrad):
    B, L, Cin, Cout, K = shape
    pad = K // 2
    Lout = L + 2 * pad - K + 1
    if Lout < 1:
        return
    v = plan(B, L, Cin, Cout, K, pad, 1 if dgrad else 0)
    variant, NT, MT, AS, WS, slot, tma_store, smem, gx, gy, gz, cols, tiles, units = v[:14]
    red, outc, lout = (Cout, Cin, L) if dgrad else (Cin, Cout, Lout)            # the data gradient is a forward conv of dy
    if red % 64 != 0 or outc % 32 != 0 or lout + K - 1 <= 16 or (dgrad and outc % 64 != 0):
        assert variant == 0                                                   # narrow layers stay on the plane kernels
        return
    assert variant in (1, 2)
    assert smem <= 227 * KB and NT * MT <= cols <= 512 and cols & (cols - 1) == 0
    if cols <= 256 and smem <= 113 * KB:
        assert 2 * (smem + KB) <= 228 * KB                                    # two CTAs (and their TMEM) fit one SM
    assert AS in (1, 2) and 2 <= WS <= 8 and slot in (0, 32, 64)
    assert (slot > 0) == (lout + K - 1 <= 64)
    assert tiles == (B * ((lout + 127) // 128) if slot == 0 else -(-B // (128 // slot)))
    assert units * MT >= tiles and outc % NT == 0
    if tma_store:
        tile_stride = ((128 + K - 1) * 128 + 1023) // 1024 * 1024
        assert NT % 64 == 0 and AS * MT * tile_stride >= 32 * KB          # room for the four warps' store slabs
    if variant == 2:
        assert outc % 256 == 0 and NT == 256 and units % 2 == 0 and (gx, gy, gz) == (units, outc // NT, 1)
    else:
        assert gx == outc // NT and gy * gz >= units and gy <= 32768 and gz <= 65535


@settings(max_examples=400, deadline=None)
@given(shapes)
def test_weight_gradient_plans_cover_every_tap_and_tile(shape):
    B, L, Cin, Cout, K = shape
    pad = K // 2
    Lout = L + 2 * pad - K + 1
    if Lout < 1:
        return
    v = plan(B, L, Cin, Cout, K, pad, 2)
    variant, NC, TP, groups, splits, slot, _, smem, gx, gy, gz, cols, tiles = v[:13]
    if Cin % 64 != 0 or Cout % 64 != 0 or (Cout > 64 and Cout % 128 != 0) or Lout + K - 1 <= 16:
        assert variant == 0
        return
    assert variant == 1 and NC in (64, 128) and NC * TP == 512 and cols == 512
    assert groups * TP >= K > (groups - 1) * TP
    assert 1 <= splits <= tiles and gz == groups * splits and gz <= 65535
    assert gx * 128 >= Cout and gy * NC == Cin and smem <= 227 * KB
    assert gx * gy * gz <= max(2 * 148, gx * gy * groups)                      # at most two waves of one-CTA-per-SM kernels


def test_known_shapes_pick_the_measured_configurations():
    # the ResNet stages and the sweep's large case (see profiles/README.md)
    assert plan(4096, 500, 256, 256, 7, 3, 0)[:5] == [2, 256, 1, 2, 4]         # CTA pairs, N = 256
    assert plan(1024, 500, 128, 128, 7, 3, 0)[:3] == [1, 128, 2]               # two tiles per CTA, two CTAs per SM
    assert plan(64, 360, 128, 128, 7, 3, 0)[:3] == [1, 128, 1]                 # small problem: one tile per CTA
    assert plan(64, 45, 768, 768, 7, 3, 0)[5] == 64                            # short windows packed two per tile
    assert plan(256, 500, 16, 16, 5, 2, 0)[0] == 0                             # narrow layer: plane kernel
    assert plan(4096, 500, 256, 256, 3, 1, 2)[1:4] == [128, 4, 1]              # wgrad: 128 input channels, one tap group
    assert plan(256, 500, 256, 256, 31, 15, 2)[1:4] == [64, 8, 4]              # wide filter: 8 taps per CTA, 4 groups
