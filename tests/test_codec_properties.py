"""Property tests of the CPU golden model (csrc/oracle, csrc/common/bfp_codec.h) with hypothesis: the
invariants the device kernels and the fused all-reduce rely on.  No GPU.  The reference has no unit
tests for its codec cores; these pin the behaviour that hw/bf16_to_bfp_core.sv /
hw/bfp_to_bf16_core.sv and the MX extension are restated with."""
import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from fpga_ai_nic_b200.ops import oracle as O

finite = st.floats(min_value=-(2.0 ** 40), max_value=2.0 ** 40, allow_nan=False, allow_infinity=False, width=32)
blocks = st.lists(finite, min_size=16, max_size=16)
scales = st.integers(min_value=-60, max_value=60)


def _t(vals, k=0):
    return torch.tensor(vals, dtype=torch.float32) * (2.0 ** k)


@settings(derandomize=True, max_examples=200, deadline=None)
@given(blocks, scales)
def test_bfp_clean_error_is_half_a_step_and_requantisation_is_exact(vals, k):
    x = _t(vals, k)
    y = O.bfp_roundtrip(x, O.CLEAN)
    amax = x.abs().max().item()
    if amax == 0 or not np.isfinite(amax) or amax < 2.0 ** -100:
        return
    E = int(np.floor(np.log2(amax)))           # unbiased exponent of the block maximum
    step = 2.0 ** (E - 6)                       # int8 grid: 2^(E-133+127)
    err = (x.double() - y.double()).abs()
    # RNE onto the grid, except where +-127 clamps values in the top half-step of the binade
    assert bool((err <= step * 0.5 + 1e-45).logical_or(x.abs() >= 127.0 * step).all())
    assert bool((y.abs() <= 127.0 * step).all())
    # what every rank decodes is a fixed point: re-encoding it (the published shard of the two-shot
    # all-reduce, a checkpoint round trip) changes nothing
    assert torch.equal(O.bfp_roundtrip(y, O.CLEAN), y)


@settings(derandomize=True, max_examples=200, deadline=None)
@given(blocks, scales)
def test_bfp_hw_exact_truncates_toward_minus_infinity(vals, k):
    x = _t(vals, k)
    y = O.bfp_roundtrip(x, O.HW_EXACT)
    amax = x.abs().max().item()
    if amax < 2.0 ** -100:
        return
    E = int(np.floor(np.log2(amax)))
    step = 2.0 ** (E - 6)
    d = (x.double() - y.double())
    # floor rounding (hw/bf16_to_bfp_core.sv:121-126): the decoded value never exceeds the input by more
    # than the zero quirk (q == 0 decodes to +2^(E-150)) and is at most one step below it
    assert bool((d > -step * 2.0 ** -16).all()) and bool((d < step * (1 + 2.0 ** -16)).all())


@settings(derandomize=True, max_examples=100, deadline=None)
@given(st.lists(finite, min_size=32, max_size=32), scales)
def test_mx_roundtrip_bound_and_fixed_point(vals, k):
    x = _t(vals, k)
    buf = O.mx_compress(torch.cat([x, torch.zeros(512 - 32)]))
    y = O.mx_decompress(buf, 512)[:32]
    amax = x.abs().max().item()
    if amax < 2.0 ** -100 or amax > 2.0 ** 100:
        return
    # e4m3 has 3 mantissa bits: relative error <= 2^-4 for values in the normal range of the block scale;
    # the scale 2^(E-8) puts the block maximum in the top binade, where everything above 448 saturates
    E = int(np.floor(np.log2(amax)))
    scale = 2.0 ** (E - 8)
    sat = x.abs() > 448.0 * scale
    normal = (x.abs() >= scale * 2.0 ** -6) & ~sat
    rel = ((x - y).abs() / x.abs().clamp_min(1e-38))[normal]
    assert bool((rel <= 2.0 ** -4 + 1e-6).all())
    assert bool((y.abs()[sat] == 448.0 * scale).all())
    buf2 = O.mx_compress(torch.cat([y, torch.zeros(512 - 32)]))
    assert torch.equal(O.mx_decompress(buf2, 512)[:32], y)


@settings(derandomize=True, max_examples=30, deadline=None)
@given(st.integers(min_value=1, max_value=8), st.integers(min_value=1, max_value=3000),
       st.sampled_from(["oneshot", "twoshot"]), st.sampled_from([O.CODEC_BFP, O.CODEC_MXFP8]), st.integers(0, 2 ** 31 - 1))
def test_fused_allreduce_model_replicas_identical_and_close_to_exact(nranks, count, algo, codec, seed):
    g = torch.randn(nranks, count, generator=torch.Generator().manual_seed(seed)) * 1e-2
    out = O.allreduce(g, algo, codec, O.CLEAN)
    exact = O.allreduce(g, "exact")
    for r in range(1, nranks):
        assert torch.equal(out[r], out[0])                     # weight replicas cannot drift (SURVEY D11 fixed)
    tol = (0.02 if codec == O.CODEC_BFP else 0.09) * (2 if algo == "twoshot" else 1)
    # The quantisation noise scales with the magnitude of the CONTRIBUTIONS, not of their sum (a handful of
    # values that nearly cancel would make any bound relative to the exact sum meaningless): measure it
    # against the norm of all contributions.
    denom = g.norm().clamp_min(1e-20)
    assert ((out[0] - exact[0]).norm() / denom).item() <= tol * 1.5 + 1e-6


@settings(derandomize=True, max_examples=50, deadline=None)
@given(st.integers(min_value=1, max_value=5000), st.floats(min_value=0.0, max_value=1.0, width=32), st.integers(0, 2 ** 31 - 1))
def test_sgd_is_a_single_fma(count, lr, seed):
    gen = torch.Generator().manual_seed(seed)
    w, g = torch.randn(count, generator=gen), torch.randn(count, generator=gen)
    want = (w.double() - float(np.float32(lr)) * g.double()).float()    # one rounding, like the hard-FP ffma
    assert torch.equal(O.sgd(w, g, lr), want)
