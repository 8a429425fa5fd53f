"""Property tests: random shapes, ratios, ties, signed zeros, denormals and
all-zero rows through select -> quantise/pack -> dequantise, always against the oracle."""
import numpy as np
import pytest
import torch
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

import splitq
import splitq_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@st.composite
def case(draw):
    S = draw(st.integers(1, 300))
    D = draw(st.sampled_from([1, 3, 4, 20, 64, 130, 512, 896]))
    seed = draw(st.integers(0, 2**31 - 1))
    mode = draw(st.sampled_from([O.SCALE_GROUP, O.SCALE_TOKEN]))
    widths = draw(st.permutations([2, 4, 8]))
    n_rungs = draw(st.integers(0, 3))
    cuts = sorted(draw(st.lists(st.floats(0, 1), min_size=n_rungs, max_size=n_rungs)))
    ks, prev = [], 0.0
    for c in cuts:
        ks.append(int((c - prev) * S))
        prev = c
    ladder = list(zip(ks, widths[:n_rungs]))
    flavour = draw(st.sampled_from(["normal", "ties", "denormal", "zero_rows", "huge", "signed_zero"]))
    return S, D, seed, mode, ladder, flavour


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(case())
def test_boundary_roundtrip_matches_oracle(c):
    S, D, seed, mode, ladder, flavour = c
    gen = torch.Generator().manual_seed(seed)
    h = torch.randn(S, D, generator=gen)
    imp = torch.rand(S, generator=gen)
    if flavour == "ties":
        imp = torch.randint(0, 3, (S,), generator=gen).float()
    elif flavour == "denormal":
        h = h * 1e-41
        imp = imp * 1e-42
    elif flavour == "zero_rows":
        h[::2] = 0.0
    elif flavour == "huge":
        h = h * 1e30
    elif flavour == "signed_zero":
        imp[::2] = 0.0
        imp[1::4] = -0.0
        h[:, ::2] = -0.0
    nbytes = splitq.payload_bytes(S, D, ladder, mode)
    assert nbytes == O.payload_bytes(S, D, ladder, mode)
    bits, off = splitq.select_bits(imp.unsqueeze(0).to(DEV), ladder, D)
    want_bits, want_payload, want_deq = O.boundary_roundtrip(h, imp, ladder, mode)
    assert np.array_equal(bits[0].cpu().numpy(), want_bits)
    payload = splitq.quant_pack(h.unsqueeze(0).to(DEV), bits, off, mode, nbytes)
    assert np.array_equal(payload[0].cpu().numpy(), want_payload)
    deq = splitq.dequant(payload, S, D)[0].cpu()
    assert torch.equal(torch.isnan(deq), torch.isnan(want_deq))
    assert np.array_equal(torch.nan_to_num(deq).numpy().view(np.uint32), torch.nan_to_num(want_deq).numpy().view(np.uint32))


@settings(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(st.integers(1, 400), st.integers(1, 9), st.integers(0, 2**31 - 1), st.booleans())
def test_materialised_importance_property(S, H, seed, rectangular):
    gen = torch.Generator().manual_seed(seed)
    Sq = S + 3 if rectangular else S
    A = torch.softmax(torch.randn(1, H, Sq, S, generator=gen) * 3, dim=-1)
    col, last = splitq.importance_from_attention(A.to(DEV), want_last=True)
    want_c, want_l = O.importance_regular(A), O.importance_last_row(A)
    assert ((col[0].cpu() - want_c).abs() / want_c.abs().clamp_min(1e-30)).max().item() <= 1e-5
    assert ((last[0].cpu() - want_l).abs() / want_l.abs().clamp_min(1e-30)).max().item() <= 1e-5
    assert abs(col.sum().item() - 1.0) < 1e-4  # rows of A sum to one
